"""GPU parity of the codestream-level door (grkb_decompress_*, include/grk_b200.h section E): the
STOCK reference encoder's codestreams go in as bytes, our own T2 + the CUDA tile decoders produce
the image, and it must equal the STOCK reference decoder's image of the same codestream --
including the BASELINE shapes themselves (configs[1], [2], [3])."""
import os

import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import codestream as CS
from grok_b200 import tile as T
from oracle import oracle as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")]


def _encode(img, prec, irr, opts="", rate=0.0, **kw):
    try:
        O.ref_set_rate(rate)
        O.ref_set_encode_options(opts)
        return O.ref_codec_encode(img, prec, irr, threads=os.cpu_count(), **kw)
    finally:
        O.ref_set_rate(0)
        O.ref_set_encode_options("")


def _ref_decode(cs, ncomp, reduce=0, layers=0):
    c = CS.Codestream(cs)
    w, h = c.image_size(reduce)
    c.close()
    try:
        O.ref_set_decode_options(reduce, layers)
        return O.ref_codec_decode(cs, ncomp, w, h, threads=os.cpu_count())
    finally:
        O.ref_set_decode_options(0, 0)


@pytest.fixture(scope="module")
def dec(gpu):
    d = CS.Decompressor(0, slots=3)
    yield d
    d.close()


CASES = [
    # w, h, ncomp, prec, irr, ht, numres, cblk, tile, opts, rate
    (200, 150, 3, 8, 0, 0, 5, (32, 32), (0, 0), "prog=RPCL;prc=64x64:32x32;rates=40:15:1", 0.0),
    (250, 190, 3, 8, 0, 0, 5, (32, 32), (96, 80), "off=13:7;torg=5:3;prog=PCRL;prc=64x64", 0.0),
    (250, 190, 3, 8, 0, 0, 4, (64, 64), (128, 128), "rates=20:1;tp=R;eph", 0.0),
    (130, 100, 1, 10, 0, 0, 5, (64, 64), (0, 0), "sty=63;rates=30:8:1", 0.0),
    (130, 100, 1, 10, 0, 0, 5, (64, 64), (0, 0), "sty=5", 0.0),
    (200, 150, 3, 8, 0, 0, 5, (32, 32), (0, 0), "poc=0,0,1,3,3,CPRL/3,0,1,5,3,RPCL;prc=64x64", 0.0),
    (320, 200, 3, 12, 0, 1, 6, (64, 64), (0, 0), "", 0.0),
    (300, 260, 3, 12, 0, 1, 4, (64, 64), (128, 128), "prog=RPCL;prc=128x128;plt", 0.0),
    (256, 200, 3, 12, 1, 0, 6, (64, 64), (0, 0), "rates=40:20:10", 0.0),
    (256, 200, 3, 12, 1, 0, 6, (64, 64), (100, 90), "prog=CPRL;prc=128x128", 6.0),
    (512, 270, 3, 12, 1, 0, 6, (32, 32), (0, 0), "cinema=4k24", 0.0),
    (160, 120, 3, 8, 0, 0, 5, (32, 32), (0, 0), "roi=0:5", 0.0),
    (160, 120, 3, 10, 1, 0, 5, (32, 32), (0, 0), "roi=1:3", 4.0),
    (97, 61, 4, 9, 0, 0, 3, (16, 16), (40, 40), "", 0.0),
]


@pytest.mark.parametrize("w,h,ncomp,prec,irr,ht,nr,cblk,tile,opts,rate", CASES)
def test_decompress_matches_stock_decoder(dec, w, h, ncomp, prec, irr, ht, nr, cblk, tile, opts, rate):
    img = H.synth_image(w, h, ncomp, prec, seed=w + 3 * h)
    cs = _encode(img, prec, irr, opts, rate, ht=bool(ht), numres=nr, cblk=cblk, tile=tile)
    want = _ref_decode(cs, ncomp)
    got = dec.decode(cs)
    assert dec.rejected == 0
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c]), "component %d (%s)" % (c, opts)


@pytest.mark.parametrize("reduce,layers", [(1, 0), (2, 0), (3, 0), (0, 1), (1, 2)])
@pytest.mark.parametrize("irr", [0, 1])
def test_reduce_and_layers_match_stock_decoder(dec, reduce, layers, irr):
    """cp_reduce / cp_layer (row N4, reduced-resolution half): the last wavelet level that runs is a
    lower one and takes the fused MCT + DC shift + clamp"""
    img = H.synth_image(250, 190, 3, 8, seed=9)
    cs = _encode(img, 8, irr, "prog=RPCL;rates=30:10:%d;off=3:5" % (2 if irr else 1), tile=(128, 96), numres=5,
                 cblk=(32, 32))
    want = _ref_decode(cs, 3, reduce, layers)
    got = dec.decode(cs, reduce=reduce, max_layers=layers)
    for c in range(3):
        assert np.array_equal(got[c], want[c])


@pytest.mark.parametrize("interleaved", [False, True])
@pytest.mark.parametrize("dtype", [np.uint16, np.uint8])
def test_packed_output_multi_tile(dec, orc, interleaved, dtype):
    """row N3 at the codestream door: every tile lands in its rectangle of one narrow image buffer"""
    prec = 12 if dtype == np.uint16 else 8
    img = H.synth_image(250, 190, 3, prec, seed=2)
    cs = _encode(img, prec, 0, "off=3:5", tile=(128, 96), numres=4, cblk=(32, 32))
    want = orc.orc_pack_samples(_ref_decode(cs, 3), dtype, interleaved)
    out = np.full(want.shape, 0xAB, dtype=dtype)
    dec.submit_packed(cs, out, 3, interleaved)
    dec.wait()
    assert np.array_equal(out, want)


@pytest.mark.parametrize("irr,ht", [(0, 0), (1, 0), (0, 1)])
@pytest.mark.parametrize("tile", [(0, 0), (128, 96)])
@pytest.mark.parametrize("win,reduce", [((40, 30, 170, 140), 0), ((0, 0, 64, 64), 0), ((129, 97, 250, 190), 0),
                                        ((100, 50, 101, 51), 0), ((37, 21, 203, 177), 1), ((60, 60, 200, 150), 2)])
def test_windowed_decode_matches_stock_decoder(dec, irr, ht, tile, win, reduce):
    """row N4, the windowed half: only the tiles and code blocks the window depends on are decoded;
    the samples equal the stock decoder's own windowed decode (grk_decompress_set_window)"""
    prec = 12 if ht else 8
    img = H.synth_image(250, 190, 3, prec, seed=31)
    cs = _encode(img, prec, irr, "" if ht else "rates=30:10:%d" % (2 if irr else 1), tile=tile, numres=5,
                 cblk=(32, 32), ht=bool(ht))
    w = ((win[2] + (1 << reduce) - 1) >> reduce) - ((win[0] + (1 << reduce) - 1) >> reduce)
    h = ((win[3] + (1 << reduce) - 1) >> reduce) - ((win[1] + (1 << reduce) - 1) >> reduce)
    try:
        O.ref_set_decode_window(*win)
        O.ref_set_decode_options(reduce, 0)
        want = O.ref_codec_decode(cs, 3, w, h)
    finally:
        O.ref_set_decode_window()
        O.ref_set_decode_options(0, 0)
    got = dec.decode_window(cs, win, reduce=reduce)
    for k in range(3):
        assert got[k].shape == want[k].shape and np.array_equal(got[k], want[k])


@pytest.mark.parametrize("irr", [0, 1])
@pytest.mark.parametrize("win,reduce", [((40, 30, 170, 140), 0), ((129, 97, 250, 190), 0), ((37, 21, 203, 177), 1)])
def test_windowed_decode_with_image_offset_is_the_crop_of_the_full_decode(dec, irr, win, reduce):
    """With an image origin other than (0, 0) the REFERENCE's windowed decode disagrees with its own
    full decode (and, for lossless streams, with the image): a reference defect.  Ours is, as the
    window of a decode must be, the same region of the full decode -- the stock decoder's."""
    img = H.synth_image(250, 190, 3, 8, seed=31)
    cs = _encode(img, 8, irr, "off=3:5;rates=30:10:%d" % (2 if irr else 1), tile=(128, 96), numres=5, cblk=(32, 32))
    full = _ref_decode(cs, 3, reduce)
    x0, y0 = 3 + win[0], 5 + win[1]
    x1, y1 = 3 + win[2], 5 + win[3]
    up = lambda v: (v + (1 << reduce) - 1) >> reduce
    ox, oy = up(3), up(5)
    got = dec.decode_window(cs, win, reduce=reduce)
    for k in range(3):
        assert np.array_equal(got[k], full[k][up(y0) - oy:up(y1) - oy, up(x0) - ox:up(x1) - ox])


def test_windowed_decode_skips_work(dec, frame16):
    """a 512x512 window of the 4K HT frame: same samples as the full decode's region, a fraction of
    the blocks"""
    img = [p >> 4 for p in frame16]
    cs = _encode(img, 12, 0, ht=True, numres=6, cblk=(64, 64))
    got = dec.decode_window(cs, (1800, 900, 2312, 1412))
    for k in range(3):
        assert np.array_equal(got[k], img[k][900:1412, 1800:2312])
    c = CS.Codestream(cs)
    full = c.parse_tile(0)
    c.close()
    assert dec.stats()["tiles"] == 1 and len(full["blocks"]) == 6708


def test_refused_streams_fail_loudly(dec):
    with pytest.raises(Exception):
        dec.decode(np.frombuffer(b"garbage" * 10, dtype=np.uint8))


# ---------------------------------------------------------------------------- BASELINE shapes

W4K, H4K = 4096, 2160


@pytest.fixture(scope="module")
def frame16():
    return H.synth_image(W4K, H4K, 3, 16, seed=4096)


def test_configs1_mq97_4k_codestream(dec, frame16):
    """BASELINE configs[1] through the codestream door: 26 MB of stock codestream in, samples equal
    to the stock decoder's (bit-exact; the contract allows +-1)"""
    cs = _encode(frame16, 16, 1, numres=6, cblk=(64, 64))
    want = _ref_decode(cs, 3)
    got = dec.decode(cs)
    assert dec.rejected == 0 and dec.stats()["tiles"] == 1
    for c in range(3):
        assert np.array_equal(got[c], want[c])


@pytest.mark.parametrize("irr", [1, 0])
def test_4k_frame_of_32x32_blocks_two_blocks_per_warp(dec, frame16, irr):
    """a 4096x2160 12-bit frame in 32x32 code blocks and 256x256 precincts (the cinema profiles'
    sizes): 26 280 blocks, enough for the MQ decoder to put two blocks on every warp (t1_mq_dec.cu
    picks it by block count); == stock decoder"""
    img = [p >> 4 for p in frame16]
    cs = _encode(img, 12, irr, "prc=256x256", numres=6, cblk=(32, 32))
    want = _ref_decode(cs, 3)
    assert "GRKB_T1_MQ_GROUP" not in os.environ
    got = dec.decode(cs)
    assert dec.rejected == 0
    for c in range(3):
        assert np.array_equal(got[c], want[c])
        assert irr or np.array_equal(got[c], img[c])


def test_precincts_of_more_than_1024_blocks(dec, frame16):
    """The same frame with maximal precincts: 64x34 = 2 176 blocks in a precinct of the finest bands.
    The STOCK DECODER gets such streams wrong (cache/SparseCache.h:40-54: tryGet() of a chunk that does
    not exist yet records its index but keeps the previous chunk's pointer, so block 1024 + i is read
    into block i's object); ours gives the image back, so the stock ENCODER's stream is right."""
    img = [p >> 4 for p in frame16]
    cs = _encode(img, 12, 0, numres=6, cblk=(32, 32))
    got = dec.decode(cs)
    assert dec.rejected == 0
    for c in range(3):
        assert np.array_equal(got[c], img[c])


def test_configs2_ht_4k_codestream(dec, frame16):
    img = [p >> 4 for p in frame16]
    cs = _encode(img, 12, 0, ht=True, numres=6, cblk=(64, 64))
    got = dec.decode(cs)
    assert dec.rejected == 0
    for c in range(3):
        assert np.array_equal(got[c], img[c])
    want = _ref_decode(cs, 3, reduce=2)
    got = dec.decode(cs, reduce=2)
    for c in range(3):
        assert np.array_equal(got[c], want[c])


@pytest.mark.parametrize("world", [2, 3])
def test_tile_subsets_compose_the_full_decode(dec, world):
    """SURVEY 8e at the codestream door: decompressor r takes tiles r, r + N, ... of the same bytes
    (grkb_decompress_submit_tiles); the rectangles are disjoint and together they are the stock
    decoder's image; the rule is grok_b200.shard.shard_units' (the gloo test's)"""
    from grok_b200 import shard
    img = H.synth_image(250, 190, 3, 8, seed=3)
    cs = _encode(img, 8, 0, "off=3:5;torg=1:2", tile=(96, 64), numres=4, cblk=(32, 32))
    want = _ref_decode(cs, 3)
    c = CS.Codestream(cs)
    ntiles = c.num_tiles
    c.close()
    tiles = shard.tile_grid(3, 5, 253, 195, 1, 2, 96, 64)
    assert len(tiles) == ntiles
    full = [np.zeros_like(p) for p in want]
    for r in range(world):
        out = [np.full(p.shape, -7, np.int32) for p in want]
        dec.submit_tiles(cs, out, r, world)
        dec.wait()
        mine = set(shard.shard_units(ntiles, r, world))
        for t, (x0, y0, x1, y1) in enumerate(tiles):
            for k in range(3):
                rect = out[k][y0 - 5:y1 - 5, x0 - 3:x1 - 3]
                if t in mine:
                    assert np.array_equal(rect, want[k][y0 - 5:y1 - 5, x0 - 3:x1 - 3]), (r, t)
                    full[k][y0 - 5:y1 - 5, x0 - 3:x1 - 3] = rect
                else:
                    assert np.all(rect == -7), (r, t)  # not this rank's tile: untouched
    for k in range(3):
        assert np.array_equal(full[k], want[k])


def test_configs3_8k_tiled_ht_codestream(dec, frame8k):
    """BASELINE configs[3]: 8192x8192 16-bit RGB HTJ2K in 64 tiles of 1024x1024, all tiles through
    one decompressor (three tile decoders in flight, one geometry), == stock decoder == image"""
    img = frame8k
    cs = _encode(img, 16, 0, ht=True, numres=6, cblk=(64, 64), tile=(1024, 1024))
    got = dec.decode(cs)
    assert dec.rejected == 0 and dec.stats()["tiles"] == 64
    for c in range(3):
        assert np.array_equal(got[c], img[c])
    # one tile row against the stock decoder's own samples as well
    want = _ref_decode(cs, 3)
    for c in range(3):
        assert np.array_equal(got[c], want[c])
