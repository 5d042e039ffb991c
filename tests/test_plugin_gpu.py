"""The drop-in boundary end to end: the UNMODIFIED reference host (oracle/_ref, built from
/root/reference) dlopens grok_b200/plugin/_build/libgrokj2k_plugin.so through its own
grk_plugin_load / grk_plugin_init, and
  * grk_plugin_compress: our plugin does DC shift, MCT, wavelet, quantisation and T1 on the GPU,
    the host does rate allocation, T2 and markers -> the codestream must be byte-identical to the
    one the stock CPU encoder writes for the same image and parameters;
  * grk_plugin_decompress: the host parses markers and packets (T2), our plugin does T1,
    dequantisation, inverse wavelet and inverse MCT on the GPU -> the samples must be identical
    to the stock CPU decoder's (bit-exact, 9/7 included)."""
import os

import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import codestream as CS
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def host(gpu):
    if not O.have_plugin():
        pytest.skip("reference host or plugin not built (needs /root/reference at build time)")
    O.ref_plugin_open(device=0)
    return O


CASES = [
    # w, h, ncomp, prec, numres, irreversible, ht
    (320, 200, 3, 12, 6, 0, 1),   # HTJ2K lossless RGB (BASELINE configs[2] in small)
    (257, 131, 1, 8, 4, 0, 1),    # HTJ2K lossless grey, odd sizes
    (320, 200, 3, 8, 6, 0, 0),    # J2K lossless RGB (configs[0] in small)
    (200, 333, 1, 16, 5, 0, 0),   # J2K lossless grey 16 bit
    (320, 200, 3, 12, 6, 1, 0),   # 9/7 + ICT, all passes (configs[1] in small)
]


# Through the unmodified host a plugin-driven encode only carries data when the layer is rate
# controlled: without a rate target pcrdBisect* calls makeSingleLosslessLayer() BEFORE
# compress_synch_with_plugin (tile/TileProcessor.cpp:1062 vs :1086), i.e. before the plugin's
# passes are known, and writes empty packets.  A ratio of 1 keeps every pass (the budget is the
# raw image size) and goes through rate control, so it is the lossless case of this boundary.
@pytest.mark.parametrize("ratio", [1.0, 6.0, 20.0])
@pytest.mark.parametrize("w,h,ncomp,prec,nr,irr,ht", CASES)
def test_plugin_encode_byte_identical(host, tmp_path, w, h, ncomp, prec, nr, irr, ht, ratio):
    if ht and ratio != 1.0:
        pytest.skip("HT blocks have one pass: nothing to truncate")
    img = H.synth_image(w, h, ncomp, prec, seed=11 + w)
    pnm = str(tmp_path / "in.pnm")
    O.write_pnm(pnm, img, prec)
    try:
        O.ref_set_rate(0)
        lossless = O.ref_codec_encode(img, prec, irr, ht=bool(ht), numres=nr)
        O.ref_set_rate(ratio)
        stock = O.ref_codec_encode(img, prec, irr, ht=bool(ht), numres=nr)
        ours = O.ref_plugin_encode(pnm, w, h, ncomp, irr, ht=bool(ht), numres=nr)
    finally:
        O.ref_set_rate(0)
    assert len(ours) == len(stock) and np.array_equal(ours, stock)
    if ratio == 1.0 and ht:
        # every block kept: the very codestream the stock encoder writes without a rate target
        assert len(ours) == len(lossless) and np.array_equal(ours, lossless)
    elif ratio == 1.0:
        # all passes kept (the stock encoder's own rate-1 stream differs from its rate-0 stream
        # only in how the last passes are terminated in the packet headers): still lossless
        got = O.ref_codec_decode(ours, ncomp, w, h)
        if not irr:
            assert all(np.array_equal(got[c], img[c]) for c in range(ncomp))
    else:
        assert len(ours) < len(lossless)


@pytest.mark.parametrize("w,h,ncomp,prec,nr,irr,ht", CASES)
def test_plugin_decode_identical_samples(host, tmp_path, w, h, ncomp, prec, nr, irr, ht):
    img = H.synth_image(w, h, ncomp, prec, seed=23 + h)
    cs = O.ref_codec_encode(img, prec, irr, ht=bool(ht), numres=nr)
    path = str(tmp_path / "in.j2k")
    cs.tofile(path)
    want = O.ref_codec_decode(cs, ncomp, w, h)
    got = O.ref_plugin_decode(path, ncomp, w, h)
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c])
        if not irr:
            assert np.array_equal(got[c], img[c])


@pytest.mark.parametrize("w,h,ncomp,prec,nr,irr,ht,ratio", [
    (320, 200, 3, 12, 6, 0, 1, 1.0),    # HTJ2K lossless
    (320, 200, 3, 12, 6, 1, 0, 8.0),    # 9/7 + ICT, rate controlled (BASELINE configs[5] in small)
])
def test_plugin_batch_encode_byte_identical(host, tmp_path, w, h, ncomp, prec, nr, irr, ht, ratio):
    """grk_plugin_batch_compress: a directory of frames through the plugin's worker pool (several
    device encoders in flight, host callbacks on several threads) -> every output file equals the
    stock CPU encoder's codestream for that frame."""
    indir, outdir = tmp_path / "in", tmp_path / "out"
    indir.mkdir()
    outdir.mkdir()
    frames = []
    for f in range(9):
        img = H.synth_image(w, h, ncomp, prec, seed=100 + f)
        frames.append(img)
        O.write_pnm(str(indir / ("f%03d.ppm" % f)), img, prec)
    try:
        O.ref_set_rate(ratio)
        n = O.ref_plugin_batch_encode(str(indir), str(outdir), irr, ht=bool(ht), numres=nr)
        stock = [O.ref_codec_encode(img, prec, irr, ht=bool(ht), numres=nr) for img in frames]
    finally:
        O.ref_set_rate(0)
    assert n == len(frames)
    for f in range(len(frames)):
        got = np.fromfile(str(outdir / ("f%03d.j2k" % f)), dtype=np.uint8)
        assert len(got) == len(stock[f]) and np.array_equal(got, stock[f]), "frame %d" % f


@pytest.mark.parametrize("w,h,ncomp,prec,nr,irr,ht", [
    (320, 200, 3, 12, 6, 0, 1),    # HTJ2K lossless
    (320, 200, 3, 12, 6, 1, 0),    # 9/7 + ICT, MQ (BASELINE configs[1] in small)
])
def test_plugin_batch_decode_identical_samples(host, tmp_path, w, h, ncomp, prec, nr, irr, ht):
    """grk_plugin_batch_decompress: a directory of codestreams through the plugin's worker pool
    (host T2 and output callbacks on several threads, one device decoder per worker) -> every
    frame's samples equal the stock CPU decoder's."""
    indir, outdir = tmp_path / "in", tmp_path / "out"
    indir.mkdir()
    outdir.mkdir()
    want = []
    for f in range(9):
        img = H.synth_image(w, h, ncomp, prec, seed=200 + f)
        cs = O.ref_codec_encode(img, prec, irr, ht=bool(ht), numres=nr)
        cs.tofile(str(indir / ("f%03d.j2k" % f)))
        want.append(O.ref_codec_decode(cs, ncomp, w, h))
    assert O.ref_plugin_batch_decode(str(indir), str(outdir)) == 9
    for f in range(9):
        got = O.read_plane_dump(str(outdir / ("f%03d.ppm" % f)))
        assert len(got) == ncomp
        for c in range(ncomp):
            assert np.array_equal(got[c], want[f][c]), "frame %d component %d" % (f, c)


# ---------------------------------------------------------------------------- own T2 (row N1)

def _plugin_decode(tmp_path, cs, ncomp, reduce=0, layers=0):
    path = str(tmp_path / "in.j2k")
    cs.tofile(path)
    c = CS.Codestream(cs)
    w, h = c.image_size(reduce)
    c.close()
    try:
        O.ref_set_decode_options(reduce, layers)
        want = O.ref_codec_decode(cs, ncomp, w, h)
        got = O.ref_plugin_decode(path, ncomp, w, h)
    finally:
        O.ref_set_decode_options(0, 0)
    return want, got


def _encode(img, prec, irr, opts="", rate=0.0, **kw):
    try:
        O.ref_set_rate(rate)
        O.ref_set_encode_options(opts)
        return O.ref_codec_encode(img, prec, irr, threads=os.cpu_count(), **kw)
    finally:
        O.ref_set_rate(0)
        O.ref_set_encode_options("")


@pytest.mark.parametrize("w,h,ncomp,prec,irr,ht,nr,cblk,tile,opts,rate", [
    (250, 190, 3, 8, 0, 0, 5, (32, 32), (96, 80), "off=13:7;torg=5:3;prog=PCRL;prc=64x64", 0.0),   # multi-tile
    (200, 150, 3, 8, 0, 0, 5, (32, 32), (0, 0), "prog=RPCL;prc=64x64:32x32;rates=40:15:1", 0.0),   # layers
    (130, 100, 1, 10, 0, 0, 5, (64, 64), (0, 0), "sty=5;rates=30:8:1", 0.0),                       # several segments
    (300, 260, 3, 12, 0, 1, 4, (64, 64), (128, 128), "prog=RPCL;prc=128x128", 0.0),                # HT, tiled
    (512, 270, 3, 12, 1, 0, 6, (32, 32), (0, 0), "cinema=4k24", 0.0),                              # CPRL + POC
    (160, 120, 3, 8, 0, 0, 5, (32, 32), (0, 0), "roi=0:5", 0.0),                                   # RGN
])
def test_plugin_decode_beyond_the_bridge(host, tmp_path, w, h, ncomp, prec, irr, ht, nr, cblk, tile, opts, rate):
    """grk_plugin_decompress with T2 on the plugin's side: streams a grk_plugin_tile cannot carry
    (several tiles, several segments per block, ROI) decode to the stock decoder's samples"""
    img = H.synth_image(w, h, ncomp, prec, seed=w + h)
    cs = _encode(img, prec, irr, opts, rate, ht=bool(ht), numres=nr, cblk=cblk, tile=tile)
    want, got = _plugin_decode(tmp_path, cs, ncomp)
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c])


@pytest.mark.parametrize("reduce,layers", [(1, 0), (3, 0), (0, 2)])
def test_plugin_decode_reduce_and_layers(host, tmp_path, reduce, layers):
    img = H.synth_image(250, 190, 3, 8, seed=12)
    cs = _encode(img, 8, 0, "rates=30:10:1;off=3:5", tile=(128, 96), numres=5, cblk=(32, 32))
    want, got = _plugin_decode(tmp_path, cs, 3, reduce, layers)
    for c in range(3):
        assert np.array_equal(got[c], want[c])


def test_plugin_decode_bridge_path_still_works(host, tmp_path):
    """GRKB_PLUGIN_HOST_T2=1: the host's T2 fills the grk_plugin_tile (rows P1, P2) -- and what that
    tree cannot express is refused, not decoded wrongly"""
    os.environ["GRKB_PLUGIN_HOST_T2"] = "1"
    try:
        img = H.synth_image(320, 200, 3, 12, seed=3)
        cs = _encode(img, 12, 0, ht=True, numres=6, cblk=(64, 64))
        want, got = _plugin_decode(tmp_path, cs, 3)
        assert all(np.array_equal(a, b) for a, b in zip(want, got))
        cs = _encode(img, 12, 0, "roi=0:4", numres=5, cblk=(64, 64))
        with pytest.raises(RuntimeError):
            _plugin_decode(tmp_path, cs, 3)
        cs = _encode(img, 12, 0, numres=5, cblk=(64, 64), tile=(128, 128))
        with pytest.raises(RuntimeError):
            _plugin_decode(tmp_path, cs, 3)
    finally:
        del os.environ["GRKB_PLUGIN_HOST_T2"]


# ---------------------------------------------------------------------------- BASELINE shapes

@pytest.fixture(scope="module")
def frame16():
    return H.synth_image(4096, 2160, 3, 16, seed=4096)


def test_configs1_frame_through_the_plugin_door(host, tmp_path, frame16):
    """BASELINE configs[1] (the frame bench.py times) through grk_plugin_decompress.  Its LL blocks
    carry 67 coding passes against the bridge's bound of 3 * (16 + 3) = 57
    (plugin/plugin_bridge.cpp:57-72): through the host's T2 the frame cannot be handed to any plugin
    (checked below: the bridge path fails it); with T2 on our side it decodes, bit-exact."""
    cs = _encode(frame16, 16, 1, numres=6, cblk=(64, 64))
    c = CS.Codestream(cs)
    t = c.parse_tile(0)
    c.close()
    assert int(t["segs"]["numpasses"].max()) > 3 * (16 + 3)
    want, got = _plugin_decode(tmp_path, cs, 3)
    for k in range(3):
        assert np.array_equal(got[k], want[k])
    os.environ["GRKB_PLUGIN_HOST_T2"] = "1"
    try:
        with pytest.raises(RuntimeError):
            _plugin_decode(tmp_path, cs, 3)
    finally:
        del os.environ["GRKB_PLUGIN_HOST_T2"]


def test_configs2_ht_4k_through_the_plugin_doors(host, tmp_path, frame16):
    """BASELINE configs[2]: 4096x2160 12-bit HTJ2K lossless.  Encode through grk_plugin_compress is
    byte-identical to the stock encoder's codestream; decode through grk_plugin_decompress gives
    the stock decoder's samples (= the image)."""
    img = [p >> 4 for p in frame16]
    pnm = str(tmp_path / "in.ppm")
    O.write_pnm(pnm, img, 12)
    lossless = _encode(img, 12, 0, ht=True, numres=6, cblk=(64, 64))
    try:
        O.ref_set_rate(1.0)
        ours = O.ref_plugin_encode(pnm, 4096, 2160, 3, 0, ht=True, numres=6)
    finally:
        O.ref_set_rate(0)
    assert len(ours) == len(lossless) and np.array_equal(ours, lossless)
    want, got = _plugin_decode(tmp_path, ours, 3)
    for k in range(3):
        assert np.array_equal(got[k], want[k]) and np.array_equal(got[k], img[k])


def test_configs3_8k_tiled_through_the_plugin_door(host, tmp_path, frame8k):
    """BASELINE configs[3]: 8192x8192 16-bit HTJ2K in 1024x1024 tiles -- 64 tiles through one
    grk_plugin_decompress call (impossible across grk_plugin_tile, which holds one tile)"""
    img = frame8k
    cs = _encode(img, 16, 0, ht=True, numres=6, cblk=(64, 64), tile=(1024, 1024))
    path = str(tmp_path / "in.j2k")
    cs.tofile(path)
    got = O.ref_plugin_decode(path, 3, 8192, 8192)
    for k in range(3):
        assert np.array_equal(got[k], img[k])


def test_plugin_encode_cinema_profile_small(host, tmp_path):
    """-x 24 / -w 24: the plugin partitions the tile as the host's grk_compress_init is going to
    (32x32 blocks, 256x256 precincts, 9/7, CPRL + POC on the host side) -> byte-identical"""
    for prof, (w, h) in (("cinema=4k24", (640, 360)), ("cinema=2k24", (512, 270))):
        img = H.synth_image(w, h, 3, 12, seed=w)
        pnm = str(tmp_path / "in.ppm")
        O.write_pnm(pnm, img, 12)
        try:
            O.ref_set_encode_options(prof)
            stock = O.ref_codec_encode(img, 12, 0, numres=6)
            ours = O.ref_plugin_encode(pnm, w, h, 3, 0, numres=6)
        finally:
            O.ref_set_encode_options("")
        assert len(ours) == len(stock) and np.array_equal(ours, stock), prof


def test_configs4_dci4k_frames_through_the_batch_door(host, tmp_path, frame16):
    """BASELINE configs[4]: DCI-4K 12-bit frames through grk_plugin_batch_compress with the
    cinema-4K profile (9/7, 32x32 blocks, 256x256 precincts, CPRL + POC, 1 302 083 bytes per frame):
    every file byte-identical to the stock encoder's"""
    indir, outdir = tmp_path / "in", tmp_path / "out"
    indir.mkdir()
    outdir.mkdir()
    base = [p >> 4 for p in frame16]
    frames = [base, [np.roll(p, 37, axis=1) for p in base]]
    for f, img in enumerate(frames):
        O.write_pnm(str(indir / ("f%03d.ppm" % f)), img, 12)
    try:
        O.ref_set_encode_options("cinema=4k24")
        n = O.ref_plugin_batch_encode(str(indir), str(outdir), 0, numres=6)
        stock = [O.ref_codec_encode(img, 12, 0, numres=6, threads=os.cpu_count()) for img in frames]
    finally:
        O.ref_set_encode_options("")
    assert n == len(frames)
    for f in range(len(frames)):
        got = np.fromfile(str(outdir / ("f%03d.j2k" % f)), dtype=np.uint8)
        assert len(stock[f]) <= 1302083
        assert len(got) == len(stock[f]) and np.array_equal(got, stock[f]), "frame %d" % f
