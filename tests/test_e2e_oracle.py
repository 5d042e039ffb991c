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
