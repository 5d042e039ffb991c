"""CPU: the oracle's whole decode chain (T1 decode -> dequantise -> inverse wavelet -> inverse MCT
+ DC shift) on tiles coded with the oracle's forward chain, against what the STOCK reference
codec decodes from its own codestream of the same image (golden fixture e2e.npz; live through
oracle/_ref when it is present).  Bit-exact, 9/7 included."""
import os

import numpy as np
import pytest

import codec_helpers as H
from oracle import oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def test_oracle_chain_matches_reference_decode_golden(orc):
    e = np.load(os.path.join(G, "e2e.npz"))
    for k, (w, h, prec, irr, nr) in enumerate(e["cases"]):
        w, h, prec, irr, nr = int(w), int(h), int(prec), int(irr), int(nr)
        img = H.synth_image(w, h, 3, prec, 100 + k)
        coefs = H.forward_oracle(img, prec, not irr, nr, 1)
        tile = H.encode_tile(coefs, prec, not irr, nr)
        got, _ = H.decode_tile_oracle(tile, 3, prec, not irr, nr, 1)
        assert np.array_equal(np.stack(got), e["decoded_%d" % k]), k


@pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")
def test_oracle_chain_matches_reference_decode_live(orc):
    for (w, h, prec, irr, nr) in [(200, 150, 8, 0, 4), (256, 135, 12, 1, 6), (77, 131, 16, 1, 2)]:
        img = H.synth_image(w, h, 3, prec, 7)
        cs = O.ref_codec_encode(img, prec, irr, numres=nr)
        ref = O.ref_codec_decode(cs, 3, w, h)
        coefs = H.forward_oracle(img, prec, not irr, nr, 1)
        tile = H.encode_tile(coefs, prec, not irr, nr)
        got, _ = H.decode_tile_oracle(tile, 3, prec, not irr, nr, 1)
        assert all(np.array_equal(a, b) for a, b in zip(ref, got))
        # the coded bytes are the codestream minus its headers
        assert 0 < len(cs) - int(tile["blocks"]["data_len"].sum()) < 4096 + len(tile["blocks"]) * 8


@pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")
def test_ht_lossless_chain_and_qcd_match_reference_codec(orc):
    """HTJ2K lossless (BASELINE config 3 in small): the QCD our restatement derives is the one the
    stock encoder writes; the oracle's HT chain decodes its own tile to the stock decoder's
    output; the coded bytes add up to the codestream minus headers (byte-identical blocks)"""
    from oracle import quant as Q
    for (w, h, prec, nr) in [(200, 150, 12, 4), (130, 67, 8, 6), (64, 64, 16, 1)]:
        img = H.synth_image(w, h, 3, prec, 9)
        cs = O.ref_codec_encode(img, prec, 0, ht=True, numres=nr)
        assert Q.parse_qcd(cs)[1] == Q.qcd_ht_rev(nr, prec)
        ref = O.ref_codec_decode(cs, 3, w, h)
        coefs = H.forward_oracle(img, prec, True, nr, 1)
        tile = H.encode_tile(coefs, prec, True, nr, ht=True)
        got, _ = H.decode_tile_oracle(tile, 3, prec, True, nr, 1)
        assert all(np.array_equal(a, b) for a, b in zip(ref, got))
        assert all(np.array_equal(a, b) for a, b in zip(img, got))
        assert 0 < len(cs) - int(tile["blocks"]["data_len"].sum()) < 4096 + len(tile["blocks"]) * 8


@pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")
def test_part1_qcd_matches_reference_codec(orc):
    from oracle import quant as Q
    img = H.synth_image(150, 100, 3, 12, 2)
    for irr, prec, nr in [(1, 12, 4), (0, 12, 6), (1, 16, 6)]:
        cs = O.ref_codec_encode([p >> (12 - min(prec, 12)) if prec < 12 else p for p in img], prec, irr, numres=nr)
        assert Q.parse_qcd(cs)[1] == Q.qcd_part1(nr, not irr, prec)
