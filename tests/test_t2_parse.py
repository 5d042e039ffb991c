"""T2 on our side (grok_b200/csrc/codestream.cpp, t2_parse.cpp: host C++, no GPU):
codestreams written by the STOCK reference encoder with every packet-level feature it offers are
parsed by the product's own marker + packet parser, the resulting descriptor tables are decoded by
the oracle's block decoders / wavelet / MCT, and the samples must equal what the STOCK reference
decoder makes of the same codestream.  A wrong byte range, pass count, segment split, bit-plane
count, step size or block position anywhere shows up as different samples (lossless cases also
give the original image back)."""
import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import codestream as CS
from oracle import oracle as O

pytestmark = pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")


def _roundtrip(w, h, ncomp, prec, irr, opts, ht=False, numres=5, cblk=(32, 32), tile=(0, 0), rate=0.0,
               reduce=0, layers=0, seed=5, want_lossless=None):
    img = H.synth_image(w, h, ncomp, prec, seed)
    try:
        O.ref_set_rate(rate)
        O.ref_set_encode_options(opts)
        cs = O.ref_codec_encode(img, prec, irr, ht=ht, numres=numres, cblk=cblk, tile=tile)
        O.ref_set_decode_options(reduce, layers)
        rw, rh = (w + (1 << reduce) - 1) >> reduce, (h + (1 << reduce) - 1) >> reduce
        c = CS.Codestream(cs)
        rw, rh = c.image_size(reduce)
        c.close()
        want = O.ref_codec_decode(cs, ncomp, rw, rh)
    finally:
        O.ref_set_rate(0)
        O.ref_set_encode_options("")
        O.ref_set_decode_options(0, 0)
    got, tabs = H.decode_codestream_oracle(cs, reduce, layers)
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c]), "component %d differs (%s)" % (c, opts)
    if want_lossless is None:
        want_lossless = not irr and rate == 0 and reduce == 0 and layers == 0 and "rates" not in opts \
            and "roi" not in opts
    if want_lossless:
        for c in range(ncomp):
            assert np.array_equal(got[c], img[c])
    return cs, tabs


@pytest.mark.parametrize("prog", ["LRCP", "RLCP", "RPCL", "PCRL", "CPRL"])
@pytest.mark.parametrize("layers", ["", "rates=40:15:1"])
def test_progression_orders_precincts_layers(orc, prog, layers):
    opts = "prog=%s;prc=64x64:32x32" % prog + (";" + layers if layers else "")
    cs, tabs = _roundtrip(200, 150, 3, 8, 0, opts)
    assert tabs[0]["packets"] > 50
    if layers:
        assert not tabs[0]["zero_copy"]  # bytes of a block spread over the layers' packets: gathered


def test_sop_counters_after_a_tile_part_boundary(orc):
    """The reference ENCODER numbers SOP markers wrongly after the first tile part of a tile (the
    counter jumps by the packets already written) and the reference DECODER then stops at the first
    such packet and delivers the lowest resolution only.  Our parser does not depend on Nsop: the
    stream decodes completely -- to the original image, its last layer being lossless."""
    img = H.synth_image(180, 130, 3, 8, 5)
    try:
        O.ref_set_encode_options("prog=RPCL;prc=64x64;rates=50:20:1;sop;eph;tp=R")
        cs = O.ref_codec_encode(img, 8, 0, numres=5, cblk=(32, 32))
    finally:
        O.ref_set_encode_options("")
    got, tabs = H.decode_codestream_oracle(cs)
    assert not tabs[0]["truncated"] and tabs[0]["packets"] == 405
    assert all(np.array_equal(a, b) for a, b in zip(got, img))


def test_precinct_with_more_than_1024_blocks(orc):
    """64 x 34 = 2 176 code blocks in one precinct.  Our parser + the oracle give the image back; the
    stock decoder does not (reference defect, cache/SparseCache.h:40-54: after tryGet() misses a chunk
    the cached chunk pointer is stale and block 1024 + i lands in block i) -- so this case is pinned to
    the image, not to the stock decoder."""
    img = H.synth_image(2048, 1080, 1, 8, 3)
    cs = O.ref_codec_encode(img, 8, 0, numres=6, cblk=(16, 16), threads=4)
    got, tabs = H.decode_codestream_oracle(cs)
    assert len(tabs[0]["blocks"]) > 8000
    assert np.array_equal(got[0], img[0])
    want = O.ref_codec_decode(cs, 1, 2048, 1080, threads=4)
    if np.array_equal(want[0], img[0]):
        pytest.fail("the stock decoder handles > 1024 blocks per precinct now: pin this case to it")


def test_single_layer_is_zero_copy(orc):
    cs, tabs = _roundtrip(200, 150, 3, 8, 0, "")
    assert tabs[0]["zero_copy"] and tabs[0]["packets"] == 15


@pytest.mark.parametrize("opts", [
    "sop;eph", "sop", "eph;prog=RPCL;prc=128x128", "tp=R", "tp=C;prog=CPRL", "tp=L;rates=30:10:1", "plt;tlm",
    "prog=PCRL;prc=64x64;rates=50:20:5:1;sop;eph",     "prog=PCRL;prc=64x64;rates=50:20:5;eph;tp=C",
])
def test_markers_and_tile_parts(orc, opts):
    _roundtrip(180, 130, 3, 8, 0, opts)


@pytest.mark.parametrize("opts,tile", [
    ("", (96, 80)), ("off=13:7;torg=5:3", (96, 80)), ("prog=RPCL;prc=64x64;off=1:1", (100, 70)),
    ("rates=20:1;tp=R", (128, 128)), ("off=64:32;torg=64:32;sop", (64, 64)),
])
def test_multi_tile_offsets(orc, opts, tile):
    cs, tabs = _roundtrip(250, 190, 3, 8, 0, opts, tile=tile)
    assert len(tabs) >= 4


@pytest.mark.parametrize("sty", [1, 2, 4, 8, 16, 32, 5, 63])
@pytest.mark.parametrize("layers", ["", "rates=30:8:1"])
def test_code_block_styles(orc, sty, layers):
    """LAZY / TERMALL give several codeword segments per block: the segment split of
    T2Decompress::initSegment (t2/T2Decompress.cpp:29-56), across layers too"""
    opts = "sty=%d" % sty + (";" + layers if layers else "")
    cs, tabs = _roundtrip(130, 100, 1, 10, 0, opts, cblk=(64, 64))
    if sty & 5:
        assert tabs[0]["blocks"]["numsegs"].max() > 1


def test_poc(orc):
    _roundtrip(200, 150, 3, 8, 0, "poc=0,0,1,3,3,CPRL/3,0,1,5,3,RPCL;prc=64x64")
    _roundtrip(200, 150, 3, 8, 0, "rates=20:5:1;poc=0,0,2,5,3,LRCP/0,0,3,5,3,RLCP")


@pytest.mark.parametrize("w,h,ncomp,prec", [(320, 200, 3, 12), (257, 131, 1, 8), (200, 130, 3, 16)])
def test_ht(orc, w, h, ncomp, prec):
    _roundtrip(w, h, ncomp, prec, 0, "", ht=True, numres=6, cblk=(64, 64))


def test_ht_tiled_precincts(orc):
    _roundtrip(300, 260, 3, 12, 0, "prog=RPCL;prc=128x128;plt", ht=True, numres=4, cblk=(64, 64), tile=(128, 128))


@pytest.mark.parametrize("opts,rate", [("", 0.0), ("rates=40:20:10", 0.0), ("", 12.0), ("prog=CPRL;prc=128x128", 6.0)])
def test_irreversible(orc, opts, rate):
    _roundtrip(256, 200, 3, 12, 1, opts, rate=rate, numres=6, cblk=(64, 64))


def test_cinema_profiles(orc):
    """-x 24 / -w 24 (BASELINE configs[4]): CPRL, 32x32 blocks, 256x256 precincts (128 at the lowest
    resolution), one tile part per component and, for 4K, the two-record POC
    (codestream/Profile.cpp:988-1150)"""
    cs, tabs = _roundtrip(512, 270, 3, 12, 1, "cinema=4k24", numres=6)
    c = CS.Codestream(cs)
    assert c.info.prog_order == 4 and c.info.cblkw_exp == 5 and c.info.tile_parts >= 3
    c.close()
    _roundtrip(512, 270, 3, 12, 1, "cinema=2k24", numres=6)


def test_roi(orc):
    """RGN: the component's bands gain roishift bit planes and the Roi* dequantisers apply"""
    cs, tabs = _roundtrip(160, 120, 3, 8, 0, "roi=0:5")
    assert tabs[0]["blocks"]["roishift"].max() == 5
    _roundtrip(160, 120, 3, 10, 1, "roi=1:3", rate=4.0)


@pytest.mark.parametrize("reduce,layers", [(1, 0), (2, 0), (4, 0), (0, 1), (0, 2), (1, 2)])
@pytest.mark.parametrize("prog", ["LRCP", "RPCL"])
def test_reduce_and_layer_limits(orc, reduce, layers, prog):
    """cp_reduce / cp_layer (the cheap half): dropped resolutions shrink the tile
    to ceil(rect / 2^reduce), dropped layers shorten the blocks' segments"""
    _roundtrip(250, 190, 3, 8, 0, "prog=%s;rates=30:10:1;off=3:5" % prog, tile=(128, 96), reduce=reduce,
               layers=layers)
    _roundtrip(250, 190, 3, 8, 1, "prog=%s;rates=30:10:2" % prog, reduce=reduce, layers=layers, cblk=(64, 64))


def test_jp2_wrapper_and_bad_input(orc):
    img = H.synth_image(64, 48, 1, 8, 1)
    cs = O.ref_codec_encode(img, 8, 0, numres=3)
    box = lambda t, body: (8 + len(body)).to_bytes(4, "big") + t + body
    jp2 = (bytes([0, 0, 0, 12]) + b"jP  " + bytes([0x0D, 0x0A, 0x87, 0x0A]) +
           box(b"ftyp", b"jp2 \0\0\0\0jp2 ") + box(b"jp2h", box(b"ihdr", bytes(14))) +
           box(b"jp2c", cs.tobytes()))
    got, _ = H.decode_codestream_oracle(np.frombuffer(jp2, dtype=np.uint8))
    assert np.array_equal(got[0], img[0])
    for bad in (b"", b"\xff\x4f", b"not a codestream at all", cs.tobytes()[:40]):
        with pytest.raises(Exception):
            CS.Codestream(np.frombuffer(bad, dtype=np.uint8))


def test_truncated_codestream_does_not_crash(orc):
    img = H.synth_image(200, 150, 3, 8, 3)
    cs = O.ref_codec_encode(img, 8, 0, numres=5)
    for frac in (0.3, 0.6, 0.95):
        cut = cs[:int(len(cs) * frac)]
        c = CS.Codestream(cut)
        t = c.parse_tile(0)
        assert t["truncated"]
        c.close()


@pytest.mark.parametrize("seed", [100, 101, 103, 105, 107, 108, 110, 113, 117, 119, 122, 124, 127, 131, 1003, 1017, 1044, 1090])
def test_drawn_combinations_on_the_host_path(orc, seed):
    """the drawn sweep of tests/test_random_gpu.py, host half: stock codestream -> our marker + packet parser
    -> the oracle's block decoders / wavelet / MCT == the stock decoder (layer- and resolution-limited too)"""
    import test_random_gpu as R
    c = R.draw(seed)
    r = c["r"]
    img = H.synth_image(c["w"], c["h"], c["ncomp"], c["prec"], seed=seed)
    layers = ""
    if not c["ht"] and r.integers(0, 2):
        layers = "rates=" + ":".join(str(x) for x in sorted(r.choice([60, 40, 25, 12, 6, 3], size=int(r.integers(2, 4)),
                                                                     replace=False).tolist(), reverse=True))
    cs = R.stock_encode(img, c, R.stock_opts(c, layers))
    if cs.size < 200 or not R.consistent(cs):
        pytest.skip("nothing valid from the stock encoder for this combination")
    reduce = int(r.integers(0, min(c["nr"], 3))) if r.integers(0, 3) == 0 else 0
    maxl = int(r.integers(1, 3)) if layers and r.integers(0, 2) else 0
    if layers and (c["sty"] & 5) and c["prog"] != "LRCP":
        reduce, maxl = 0, 0  # (stock decoder defect)
    co = CS.Codestream(cs)
    rw, rh = co.image_size(reduce)
    co.close()
    try:
        O.ref_set_decode_options(reduce, maxl)
        want = O.ref_codec_decode(cs, c["ncomp"], rw, rh, threads=4)
    finally:
        O.ref_set_decode_options(0, 0)
    got, tabs = H.decode_codestream_oracle(cs, reduce, maxl)
    x0, y0 = c["off"]
    up = lambda v: (v + (1 << reduce) - 1) >> reduce  # noqa: E731
    vw, vh = up(x0 + c["w"]) - up(x0), up(y0 + c["h"]) - up(y0)
    for k in range(c["ncomp"]):
        assert np.array_equal(got[k][:vh, :vw], want[k][:vh, :vw]), (seed, k, reduce, maxl)
