"""Narrow sample formats at the bus: grkb_decoder_submit_packed /
grkb_encoder_submit_packed against the oracle's restatement of the reference's application-side
conversions (oracle/samples.c: planarToInterleaved, convert.cpp:397-405; the 16-bit narrowing,
convert.cpp:1158-1163; convert_16u32s_C1R, :547-556).  Bit-exact: integer work."""
import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import encode as E
from grok_b200 import tile as T


# ---------------------------------------------------------------- CPU: the restatement itself

@pytest.mark.parametrize("dtype", [np.uint16, np.int16, np.uint8, np.int8])
@pytest.mark.parametrize("interleaved", [False, True])
def test_oracle_pack_is_the_reference_cast(orc, dtype, interleaved):
    rng = np.random.default_rng(7)
    info = np.iinfo(dtype)
    planes = [rng.integers(info.min, info.max + 1, size=(13, 21)).astype(np.int32) for _ in range(3)]
    got = orc.orc_pack_samples(planes, dtype, interleaved, row_pad=3)
    # convert.cpp:1158-1163 is a C cast; :397-405 puts component j of pixel i at N*i+j
    want = np.stack(planes, axis=-1 if interleaved else 0).astype(dtype)
    if interleaved:
        assert np.array_equal(got[:, :21, :], want)
    else:
        assert np.array_equal(got[:, :, :21], want)
    back = orc.orc_unpack_samples(got, 21, interleaved)
    for a, b in zip(planes, back):
        assert np.array_equal(a, b)


def test_oracle_pack_truncates_like_a_cast(orc):
    planes = [np.array([[70000, -1, 65535, 32768]], dtype=np.int32)]
    got = orc.orc_pack_samples(planes, np.uint16)
    assert got[0, 0].tolist() == [70000 & 0xFFFF, 0xFFFF, 65535, 32768]


# ---------------------------------------------------------------- GPU

def _coded_tile(w, h, ncomp, prec, nr, sgnd=False):
    img = H.synth_image(w, h, ncomp, prec, seed=w * 3 + h)
    if sgnd:
        img = [p - (1 << (prec - 1)) for p in img]
    params = T.tile_params(0, 0, w, h, ncomp, nr, 1, 1 if ncomp >= 3 else 0, prec, sgnd=sgnd)
    blocks, geom, qcd = E.enc_block_table(0, 0, w, h, ncomp, nr, prec, 1, ht=True)
    enc = E.TileEncoder(params)
    coded, passes, data = enc.encode(img, blocks)
    enc.close()
    assert not coded["failed"].any()
    dblocks, dsegs = E.decode_table(blocks, coded, passes, geom, qcd, prec, 1, True)
    pool = np.concatenate([data, np.zeros(16, np.uint8)])
    return img, params, blocks, (dblocks, dsegs, pool)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,ncomp,prec,dtype,sgnd,pad", [
    (256, 96, 3, 12, np.uint16, False, 0),    # vector path
    (250, 67, 3, 12, np.uint16, False, 0),    # ragged rows: 500-byte rows, scalar path
    (250, 67, 3, 12, np.uint16, False, 6),    # ragged width, padded rows (vector path + tail)
    (131, 50, 1, 8, np.uint8, False, 0),
    (128, 64, 3, 8, np.uint8, False, 0),
    (200, 64, 3, 10, np.int16, True, 0),      # signed components
    (96, 40, 4, 16, np.uint16, False, 0),     # full 16 bits, extra component
    (3, 1, 1, 8, np.uint8, False, 0),         # narrower than one group of four
    (5, 2, 3, 12, np.uint16, False, 0),
    (256, 96, 3, 12, ">u2", False, 0),        # big-endian samples (PNM byte order), vector path
    (250, 67, 3, 12, ">u2", False, 0),        # and the scalar path
])
@pytest.mark.parametrize("interleaved", [False, True])
def test_decode_packed_matches_oracle(gpu, orc, w, h, ncomp, prec, dtype, sgnd, pad, interleaved):
    img, params, _, (dblocks, dsegs, pool) = _coded_tile(w, h, ncomp, prec, 4 if min(w, h) >= 16 else 1, sgnd)
    dec = T.TileDecoder(params)
    planes = dec.decode(dblocks, dsegs, pool)
    for a, b in zip(planes, img):
        assert np.array_equal(a, b)  # lossless
    out = np.full((h, w + pad, ncomp) if interleaved else (ncomp, h, w + pad), 0x55, dtype=dtype)
    dec.submit_packed(dblocks, dsegs, pool, out, interleaved)
    dec.wait()
    dec.close()
    assert dec.rejected == 0
    # (a big-endian dtype only changes the byte order in memory: values compare the same)
    want = orc.orc_pack_samples(planes, np.dtype(dtype).newbyteorder("="), interleaved, row_pad=pad)
    if interleaved:
        assert np.array_equal(out[:, :w, :], want[:, :w, :])
        assert (out[:, w:, :] == 0x55).all()  # padding untouched
    else:
        assert np.array_equal(out[:, :, :w], want[:, :, :w])
        assert (out[:, :, w:] == 0x55).all()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,ncomp,prec,dtype,sgnd,pad", [
    (256, 96, 3, 12, np.uint16, False, 0),
    (250, 67, 3, 12, np.uint16, False, 2),
    (131, 50, 1, 8, np.uint8, False, 0),
    (200, 64, 3, 10, np.int16, True, 0),
    (256, 96, 3, 12, ">u2", False, 0),
    (250, 67, 3, 12, ">u2", False, 2),
])
@pytest.mark.parametrize("interleaved", [False, True])
def test_encode_packed_equals_encode_of_int32(gpu, orc, w, h, ncomp, prec, dtype, sgnd, pad, interleaved):
    img, params, blocks, _ = _coded_tile(w, h, ncomp, prec, 4 if min(w, h) >= 16 else 1, sgnd)
    enc = E.TileEncoder(params)
    coded, passes, data = enc.encode(img, blocks)
    packed = orc.orc_pack_samples(img, np.dtype(dtype).newbyteorder("="), interleaved, row_pad=pad)
    # widening is the oracle's too
    for a, b in zip(orc.orc_unpack_samples(packed, w, interleaved), img):
        assert np.array_equal(a, b)
    packed = packed.astype(dtype)  # same values, the byte order under test
    n = len(blocks)
    out = (np.zeros(n, T.CODED_DTYPE), np.zeros((n, T.ENC_MAX_PASSES), T.PASS_DTYPE),
           np.zeros(w * h * ncomp * 4 + 4096, np.uint8))
    enc.submit_packed(packed, blocks, out, interleaved=interleaved)
    coded2, passes2, data2 = enc.wait()
    enc.close()
    assert np.array_equal(coded, coded2)
    assert bytes(data) == bytes(data2)
    for i in range(n):
        k = int(coded[i]["numpasses"])
        assert np.array_equal(passes[i][:k], passes2[i][:k])


@pytest.mark.gpu
def test_packed_format_refused_when_too_narrow(gpu, orc):
    img, params, _, (dblocks, dsegs, pool) = _coded_tile(64, 64, 1, 12, 2)
    dec = T.TileDecoder(params)
    out = np.zeros((1, 64, 64), np.uint8)
    with pytest.raises(RuntimeError):
        dec.submit_packed(dblocks, dsegs, pool, out)
    dec.close()
