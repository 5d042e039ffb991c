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
