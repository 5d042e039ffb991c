"""GPU parity at the BASELINE shapes themselves: whole 4096x2160x3 frames
through the tile decoder / encoder against the UNMODIFIED reference codec (oracle/_ref: stock
encoder's codestream decoded by the stock decoder), not against a round trip.  These exercise the
full-size paths the small cases cannot: 6 708-block work stealing, > 16 MB byte pools, 32-bit
offsets, 423-CTA wavelet grids."""
import os

import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import encode as E
from grok_b200 import tile as T
from oracle import oracle as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")]

W4K, H4K = 4096, 2160


@pytest.fixture(scope="module")
def frame16():
    return H.synth_image(W4K, H4K, 3, 16, seed=4096)


def _decode_frame(tile, params):
    dec = T.TileDecoder(params, len(tile["blocks"]), tile["data"].size)
    out = dec.decode(tile["blocks"], tile["segs"], tile["data"])
    rejected = dec.rejected
    dec.close()
    return out, rejected


def test_configs1_mq97_4k_decode_vs_stock_decoder(gpu, frame16):
    """BASELINE configs[1]: 4096x2160 16-bit RGB, 9/7 + ICT, 5 levels, 64x64 blocks, all passes.
    The frame bench.py times, decoded on the GPU, equals what the stock decoder makes of the stock
    encoder's codestream of the same image: bit-exact (the contract allows +-1 LSB)."""
    img, prec, nr = frame16, 16, 6
    cs = O.ref_codec_encode(img, prec, 1, numres=nr, threads=os.cpu_count())
    want = O.ref_codec_decode(cs, 3, W4K, H4K, threads=os.cpu_count())
    coefs = gpu.stages.mct_fdwt(img, 0, 0, nr, 0, 1, [1 << (prec - 1)] * 3)
    tile = H.encode_tile(coefs, prec, False, nr, (6, 6), use_ref=True, threads=os.cpu_count())
    assert len(tile["blocks"]) == 6708 and tile["data"].size > (16 << 20)
    got, rejected = _decode_frame(tile, T.tile_params(0, 0, W4K, H4K, 3, nr, 0, 1, prec))
    assert rejected == 0
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "component %d: max |diff| %d" % (
            c, int(np.abs(got[c].astype(np.int64) - want[c]).max()))


def test_configs2_ht_4k_decode_and_encode_vs_stock_codec(gpu, frame16):
    """BASELINE configs[2]: 4096x2160 12-bit RGB HTJ2K lossless.  Encode: every block's bytes
    from the GPU tile encoder equal the reference HT encoder's on the same coefficients (which is
    what makes the codestream byte-identical once the host's T2 wraps them; the whole-file compare
    through the plugin is in test_plugin_gpu.py).  Decode: GPU decode == stock decoder == image."""
    prec, nr = 12, 6
    img = [p >> 4 for p in frame16]
    cs = O.ref_codec_encode(img, prec, 0, ht=True, numres=nr, threads=os.cpu_count())
    want = O.ref_codec_decode(cs, 3, W4K, H4K, threads=os.cpu_count())
    assert all(np.array_equal(a, b) for a, b in zip(want, img))
    params = T.tile_params(0, 0, W4K, H4K, 3, nr, 1, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, W4K, H4K, 3, nr, prec, 1, ht=True)
    enc = E.TileEncoder(params)
    coded, passes, data = enc.encode(img, blocks)
    enc.close()
    assert not coded["failed"].any()
    coefs = gpu.stages.mct_fdwt(img, 0, 0, nr, 1, 1, [1 << (prec - 1)] * 3)
    ref_tile = H.encode_tile(coefs, prec, True, nr, (6, 6), use_ref=True, threads=os.cpu_count(), ht=True)
    assert len(ref_tile["blocks"]) == len(blocks) == 6708
    # the packed stream of the GPU encoder is the reference's blocks back to back
    assert np.array_equal(coded["data_len"], ref_tile["blocks"]["data_len"])
    assert np.array_equal(data, ref_tile["data"][:data.size])
    dblocks, dsegs = E.decode_table(blocks, coded, passes, geom, qcd, prec, 1, True)
    pool = np.concatenate([data, np.zeros(16, np.uint8)])
    got, rejected = _decode_frame(dict(blocks=dblocks, segs=dsegs, data=pool), params)
    assert rejected == 0
    for c in range(3):
        assert np.array_equal(got[c], want[c])


def test_configs3_tile_of_8k_at_its_origin_vs_oracle(gpu, orc):
    """BASELINE configs[3]: one 1024x1024 tile of the 8192x8192 16-bit HTJ2K image at its own
    place in the grid (tile 3,5: origin 3072,5120 -- every band rectangle is offset), coded by the
    reference's HT block encoder on oracle coefficients, decoded on the GPU == oracle chain == image."""
    prec, nr, x0, y0, tw = 16, 6, 3072, 5120, 1024
    img = H.synth_image(tw, tw, 3, prec, seed=8192)
    coefs = H.forward_oracle(img, prec, True, nr, 1, x0=x0, y0=y0)
    tile = H.encode_tile(coefs, prec, True, nr, (6, 6), use_ref=True, x0=x0, y0=y0, ht=True)
    params = T.tile_params(x0, y0, x0 + tw, y0 + tw, 3, nr, 1, 1, prec)
    got, rejected = _decode_frame(tile, params)
    assert rejected == 0
    want, _ = H.decode_tile_oracle(tile, 3, prec, True, nr, 1, x0=x0, y0=y0, use_ref=True)
    for c in range(3):
        assert np.array_equal(got[c], want[c]) and np.array_equal(got[c], img[c])
