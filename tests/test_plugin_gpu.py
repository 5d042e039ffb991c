"""The drop-in boundary end to end: the UNMODIFIED reference host (oracle/_ref, built from
/root/reference) dlopens grok_b200/plugin/_build/libgrokj2k_plugin.so through its own
grk_plugin_load / grk_plugin_init, and
  * grk_plugin_compress: our plugin does DC shift, MCT, wavelet, quantisation and T1 on the GPU,
    the host does rate allocation, T2 and markers -> the codestream must be byte-identical to the
    one the stock CPU encoder writes for the same image and parameters;
  * grk_plugin_decompress: the host parses markers and packets (T2), our plugin does T1,
    dequantisation, inverse wavelet and inverse MCT on the GPU -> the samples must be identical
    to the stock CPU decoder's (bit-exact, 9/7 included)."""
import numpy as np
import pytest

import codec_helpers as H
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
