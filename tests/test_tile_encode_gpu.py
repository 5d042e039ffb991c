"""GPU parity of the whole-tile ENCODER (grkb_encoder_*: DC shift + MCT + forward wavelet +
quantiser + block coder + packing) against the oracle chain block by block, and encode -> decode
round trips on the GPU (lossless: identical image; 9/7: identical to the oracle's decode of the
oracle's encode)."""
import numpy as np
import pytest

import codec_helpers as H
from grok_b200 import encode as E
from grok_b200 import tile as T

pytestmark = pytest.mark.gpu


def _oracle_tile(img, prec, reversible, nr, mct, ht, sty=0):
    ncomp = len(img)
    coefs = H.forward_oracle(img[:3] if mct else img, prec, reversible, nr, mct)
    if mct and ncomp > 3:
        coefs += H.forward_oracle(img[3:], prec, reversible, nr, 0)
    return H.encode_tile(coefs, prec, reversible, nr, cblksty=sty, ht=ht)


@pytest.mark.parametrize("w,h,ncomp,prec,nr,reversible,mct,ht,sty", [
    (200, 150, 3, 12, 4, 1, 1, 1, 0),     # HTJ2K lossless RGB
    (130, 67, 1, 8, 6, 1, 0, 1, 0),       # HTJ2K lossless grey, odd sizes
    (257, 129, 4, 10, 3, 1, 1, 1, 0),     # RGB + one extra component
    (200, 150, 3, 8, 4, 1, 1, 0, 0),      # J2K lossless RGB
    (131, 97, 1, 16, 5, 1, 0, 0, 0x3F),   # all mode switches
    (192, 128, 3, 12, 4, 0, 1, 0, 0),     # 9/7 + ICT
    (64, 64, 3, 8, 1, 1, 1, 0, 0),        # no wavelet level
])
def test_tile_encode_matches_oracle_chain(gpu, orc, w, h, ncomp, prec, nr, reversible, mct, ht, sty):
    img = H.synth_image(w, h, ncomp, prec, seed=w + h)
    want = _oracle_tile(img, prec, bool(reversible), nr, mct, bool(ht), sty)
    params = T.tile_params(0, 0, w, h, ncomp, nr, reversible, mct, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, w, h, ncomp, nr, prec, reversible, ht=bool(ht),
                                          cblk_sty=sty)
    enc = E.TileEncoder(params)
    coded, passes, data = enc.encode(img, blocks)
    enc.close()
    wb = want["blocks"]
    assert len(wb) == len(blocks)
    assert not coded["failed"].any()
    for i in range(len(blocks)):
        assert (blocks[i]["x"], blocks[i]["y"], blocks[i]["comp"]) == (wb[i]["x"], wb[i]["y"], wb[i]["comp"])
        got = bytes(data[int(coded[i]["data_off"]):int(coded[i]["data_off"]) + int(coded[i]["data_len"])])
        exp = bytes(want["data"][int(wb[i]["data_off"]):int(wb[i]["data_off"]) + int(wb[i]["data_len"])])
        assert got == exp, "block %d" % i
        if not ht:
            assert coded[i]["numbps"] == wb[i]["numbps"]
    # the decoder table derived from the encoder's output equals the oracle's
    dblocks, dsegs = E.decode_table(blocks, coded, passes, geom, qcd, prec, reversible, bool(ht))
    for k in ("numbps", "k_msbs", "numsegs", "data_len") + (() if reversible else ("scale",)):
        assert np.array_equal(dblocks[k], wb[k]), k  # (the reversible dequantisers use no scale)
    assert np.array_equal(dsegs, want["segs"])
    # and the GPU decoder turns it back into the image (lossless) / the oracle's reconstruction
    dec = T.TileDecoder(params)
    pool = np.concatenate([data, np.zeros(16, np.uint8)])
    got = dec.decode(dblocks, dsegs, pool)
    assert dec.rejected == 0
    dec.close()
    ref_img, _ = H.decode_tile_oracle(want, ncomp, prec, bool(reversible), nr, mct)
    for c in range(ncomp):
        assert np.array_equal(got[c], ref_img[c])
        if reversible:
            assert np.array_equal(got[c], img[c])


def test_tile_encode_rate_control_distortions(gpu, orc):
    """with rate control on, every block's total distortion is the last pass's cumulative value"""
    w, h, nr, prec = 160, 120, 4, 8
    img = H.synth_image(w, h, 3, prec, seed=3)
    params = T.tile_params(0, 0, w, h, 3, nr, 0, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, w, h, 3, nr, prec, 0)
    enc = E.TileEncoder(params)
    coded, passes, data = enc.encode(img, blocks, mct_norms=E.MCT_NORMS_IRREV, rate_control=True)
    tm = enc.time_device(2, 1)
    enc.close()
    assert tm["total_ms"] > 0
    for i, c in enumerate(coded):
        n = int(c["numpasses"])
        if n:
            d = passes[i, :n]["distortion"]
            assert np.all(np.diff(d) >= 0) and d[-1] == c["distortion"] and d[-1] > 0
            assert np.all(np.diff(passes[i, :n]["rate"].astype(np.int64)) >= 0)


def test_two_encoders_in_flight(gpu, orc):
    """grkb_encoder_submit / grkb_encoder_wait: two encoders with different frames in flight give
    the same block tables, pass records and bytes as the synchronous call; a second submit on a
    busy encoder and a wait on an idle one are errors"""
    w, h, nr, prec = 200, 136, 4, 10
    params = T.tile_params(0, 0, w, h, 3, nr, 0, 1, prec)
    blocks, geom, qcd = E.enc_block_table(0, 0, w, h, 3, nr, prec, 0, cblk_exp=(5, 5))
    n = len(blocks)
    imgs = [H.synth_image(w, h, 3, prec, seed=40 + k) for k in range(2)]
    ref = E.TileEncoder(params)
    want = []
    for img in imgs:
        c, p, d = ref.encode(img, blocks, mct_norms=E.MCT_NORMS_IRREV, rate_control=True)
        want.append((c.copy(), p.copy(), bytes(d)))
    ref.close()
    encs = [E.TileEncoder(params) for _ in range(2)]
    outs = [(np.zeros(n, T.CODED_DTYPE), np.zeros((n, T.ENC_MAX_PASSES), T.PASS_DTYPE),
             np.zeros(w * h * 3 * 4, np.uint8)) for _ in range(2)]
    for k in range(2):
        encs[k].submit(imgs[k], blocks, outs[k], mct_norms=E.MCT_NORMS_IRREV, rate_control=True)
    with pytest.raises(Exception):
        encs[0].submit(imgs[0], blocks, outs[0])
    for k in range(2):
        c, p, d = encs[k].wait()
        wc, wp, wd = want[k]
        assert np.array_equal(c, wc) and bytes(d) == wd
        for i in range(n):
            m = int(c[i]["numpasses"])
            assert np.array_equal(p[i, :m], wp[i, :m])
    with pytest.raises(Exception):
        encs[0].wait()
    for e in encs:
        e.close()
