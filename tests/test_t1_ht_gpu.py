"""GPU parity: the CUDA HTJ2K cleanup-pass block decoder (through the C-ABI) against the golden
fixtures produced by the reference's encoder + decoder, bit-exact, with both HT dequantisers."""
import os

import numpy as np
import pytest

from grok_b200 import tile as T

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def _run(entries, reversible, pw=1100):
    """entries: [(bytes, k_msbs, w, h, numpasses[, lengths2])] stacked vertically in one plane; with
    lengths2 the block has two codeword segments: cleanup | SigProp [+ MagRef]"""
    blocks = np.zeros(len(entries), dtype=T.CBLK_DTYPE)
    segs, chunks, off, y = [], [], 0, 0
    for i, ent in enumerate(entries):
        data, kmsbs, w, h, npass = ent[:5]
        l2 = ent[5] if len(ent) > 5 else 0
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = (i * 3) % 7, y, w, h
        b["cblk_sty"], b["k_msbs"], b["numbps"] = T.CBLKSTY_HT, kmsbs, 1
        b["numsegs"], b["seg_first"] = ((2 if l2 else 1) if data is not None else 0), len(segs)
        b["data_off"], b["data_len"] = off, (len(data) if data is not None else 0)
        b["scale"] = 2.0 ** -20 * 1.25
        if data is not None:
            if l2:
                segs.append((len(data) - l2, 1))
                segs.append((l2, npass - 1))
            else:
                segs.append((len(data), npass))
            chunks.append(data)
            off += len(data)
        y += h + 1
    data = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    sa = np.array(segs, dtype=T.SEG_DTYPE) if segs else np.zeros(0, dtype=T.SEG_DTYPE)
    base = np.full((y, pw), 5, dtype=np.int32)
    got, rej = T.t1_decode(blocks, sa, data, [base], reversible)
    return blocks, got[0], rej, base


def _dequant(dec, kmsbs, reversible, scale):
    mag = (dec & 0x7FFFFFFF).astype(np.int64)
    neg = (dec >> 31).astype(bool)
    if reversible:
        v = mag >> (31 - (kmsbs + 1))
        return np.where(neg, -v, v).astype(np.int32)
    f = mag.astype(np.float32) * np.float32(scale)
    return np.where(neg, -f, f).astype(np.float32).view(np.int32)


@pytest.mark.parametrize("reversible", [1, 0])
def test_ht_decode_golden(gpu, reversible):
    t = np.load(os.path.join(G, "t1_ht.npz"))
    entries, want = [], []
    for k, (w, h, mm) in enumerate(t["cases"]):
        entries.append((t["bytes_%d" % k].tobytes(), int(mm) - 1, int(w), int(h), 1))
        want.append(t["dec_%d" % k])
    blocks, got, rej, base = _run(entries, reversible)
    assert rej == 0
    exp = base.copy()
    for b, d in zip(blocks, want):
        exp[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]] = _dequant(d, int(b["k_msbs"]), reversible,
                                                                    b["scale"])
    assert np.array_equal(got, exp)
    if reversible:
        for k, b in enumerate(blocks):
            assert np.array_equal(got[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]], t["vals_%d" % k])


@pytest.mark.parametrize("reversible", [1, 0])
def test_ht_refinement_passes_golden(gpu, reversible):
    """SigProp / MagRef: blocks with a second segment against what the reference's
    block decoder returned for the same bytes (tests/golden/make_golden_ht_refine.py)"""
    t = np.load(os.path.join(G, "t1_ht_refine.npz"))
    entries, want = [], []
    for k, (w, h, kmsbs, passes, l2) in enumerate(t["cases"]):
        entries.append((t["bytes_%d" % k].tobytes(), int(kmsbs), int(w), int(h), int(passes), int(l2)))
        want.append(t["dec_%d" % k])
    blocks, got, rej, base = _run(entries, reversible)
    assert rej == 0
    exp = base.copy()
    for b, d in zip(blocks, want):
        exp[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]] = _dequant(d, int(b["k_msbs"]), reversible,
                                                                    b["scale"])
    assert np.array_equal(got, exp)


def test_ht_refinement_passes_match_oracle(gpu, orc):
    """many more shapes (up to 1024 wide / tall), densities and refinement byte patterns, against
    the oracle restatement; mixed with cleanup-only blocks in the same launch"""
    from common import ht_refine_case
    from oracle import oracle as O
    rng = np.random.default_rng(11)
    shapes = [(64, 64), (32, 32), (8, 8), (4, 4), (5, 7), (7, 5), (1, 1), (2, 1), (1, 2), (3, 3), (16, 9),
              (9, 16), (13, 10), (33, 6), (6, 33), (64, 1), (1, 64), (64, 2), (64, 3), (64, 5), (64, 6),
              (64, 7), (12, 11), (17, 13), (100, 4), (4, 100), (40, 21), (1024, 4), (4, 1024), (512, 8),
              (96, 40), (65, 63)]
    entries, want = [], []
    for it in range(4 * len(shapes)):
        w, h = shapes[it % len(shapes)]
        _, kmsbs, data, l2 = ht_refine_case(rng, w, h, O.orc_ht_encode)
        passes = 1 + it % 3
        if passes == 1:
            data, l2 = data[:len(data) - l2], 0
        rc, dec = O.orc_ht_decode(data, kmsbs, w, h, num_passes=passes, lengths2=l2)
        assert rc == 0
        entries.append((data, kmsbs, w, h, passes, l2))
        want.append(dec)
    blocks, got, rej, base = _run(entries, 1)
    assert rej == 0
    for i, (b, d) in enumerate(zip(blocks, want)):
        e = _dequant(d, int(b["k_msbs"]), 1, b["scale"])
        assert np.array_equal(got[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]], e), (i, entries[i][1:])


def test_ht_decode_rejects_like_reference(gpu):
    """T1HT::decompress / ojph_decode_codeblock reject: more than one pass without a second
    segment (lengths2 is always 0 in Grok), a missing_msbs too small for the coded U_q, bad Scup, empty streams; uncoded blocks
    decode to zeros without error"""
    t = np.load(os.path.join(G, "t1_ht.npz"))
    good = t["bytes_7"].tobytes()
    w, h, mm = [int(v) for v in t["cases"][7]]
    bad_scup = good[:-1] + b"\xff"
    entries = [(good, mm - 1, w, h, 2), (good, 0, w, h, 1), (bad_scup, mm - 1, w, h, 1),
               (b"\x00", mm - 1, w, h, 1), (None, mm - 1, w, h, 0), (good, mm - 1, w, h, 1)]
    blocks, got, rej, base = _run(entries, 1)
    assert rej == 4
    for i in range(5):
        b = blocks[i]
        assert not got[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]].any()
    b = blocks[5]
    assert np.array_equal(got[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]], t["vals_7"])


@pytest.mark.parametrize("w,h,ncomp,prec,nr,mct", [(200, 150, 3, 12, 4, 1), (130, 67, 1, 8, 6, 0),
                                                    (64, 64, 3, 16, 1, 1), (257, 129, 4, 10, 3, 1)])
def test_ht_tile_decode_lossless(gpu, orc, w, h, ncomp, prec, nr, mct):
    """whole HT lossless tiles (HT cleanup decode -> /shift -> inverse 5/3 -> inverse RCT + DC shift)
    on the GPU reproduce the image and equal the oracle chain"""
    import codec_helpers as H
    img = H.synth_image(w, h, ncomp, prec, seed=w)
    coefs = H.forward_oracle(img[:3] if mct else img, prec, True, nr, mct)
    if mct and ncomp > 3:
        coefs += H.forward_oracle(img[3:], prec, True, nr, 0)
    tile = H.encode_tile(coefs, prec, True, nr, ht=True)
    want, _ = H.decode_tile_oracle(tile, ncomp, prec, True, nr, mct)
    dec = T.TileDecoder(T.tile_params(0, 0, w, h, ncomp, nr, 1, mct, prec))
    got = dec.decode(tile["blocks"], tile["segs"], tile["data"])
    assert dec.rejected == 0
    dec.close()
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c]) and np.array_equal(got[c], img[c])


def test_two_frames_in_flight(gpu, orc):
    """grkb_decoder_submit / grkb_decoder_wait: two decoders with different frames in flight give
    the same planes as the synchronous call"""
    import codec_helpers as H
    w, h, nr, prec = 320, 200, 4, 12
    tiles, imgs = [], []
    for seed in (1, 2):
        img = H.synth_image(w, h, 3, prec, seed=seed)
        tiles.append(H.encode_tile(H.forward_oracle(img, prec, True, nr, 1), prec, True, nr, ht=True))
        imgs.append(img)
    params = T.tile_params(0, 0, w, h, 3, nr, 1, 1, prec)
    decs = [T.TileDecoder(params), T.TileDecoder(params)]
    outs = [[T.PinnedArray((h, w), np.int32) for _ in range(3)] for _ in range(2)]
    for rep in range(3):
        for i in (0, 1):
            for p in outs[i]:
                p.array[:] = -1
            decs[i].submit(tiles[i]["blocks"], tiles[i]["segs"], tiles[i]["data"],
                           [p.array for p in outs[i]])
        for i in (0, 1):
            got = decs[i].wait()
            assert decs[i].rejected == 0
            for c in range(3):
                assert np.array_equal(got[c], imgs[i][c])
    for d in decs:
        d.close()
