"""GPU parity of the block ENCODERS (through the C-ABI, grkb_t1_encode_host) against the CPU
oracle, which is itself pinned byte-for-byte against the reference's T1::compress_cblk and
ojph_encode_codeblock (tests/test_oracle_vs_ref.py): coded bytes, bit-plane counts, per-pass
rate / length / termination and the double-precision distortion estimates must be identical."""
import numpy as np
import pytest

from grok_b200 import tile as T
from oracle import oracle as O

pytestmark = pytest.mark.gpu

SHAPES = [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9),
          (16, 16), (64, 61), (2, 2), (33, 7), (128, 32), (256, 16), (8, 512)]


def _place(shapes, rng, plane_w=300):
    """stack blocks vertically in one plane at ragged x offsets"""
    blocks = np.zeros(len(shapes), dtype=T.ENC_CBLK_DTYPE)
    y = 0
    for i, (w, h) in enumerate(shapes):
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = int(rng.integers(0, 9)), y, w, h
        y += h + int(rng.integers(0, 3))
    plane_w = max(plane_w, max(int(b["x"]) + int(b["w"]) for b in blocks))
    return blocks, y + 1, plane_w


def _coeffs(rng, h, w):
    bits = int(rng.integers(1, 14))
    mag = rng.integers(0, 1 << bits, size=(h, w))
    mag = np.where(rng.random((h, w)) < 0.3, mag, mag >> max(bits - 3, 0))
    return (mag * rng.choice([-1, 1], size=(h, w))).astype(np.int32)


def _compare_mq(coded, passes, data, i, want):
    c = coded[i]
    assert c["failed"] == 0
    assert c["numbps"] == want["numbps"]
    assert c["numpasses"] == len(want["passes"])
    got = bytes(data[int(c["data_off"]):int(c["data_off"]) + int(c["data_len"])])
    assert got == want["data"]
    for p, (rate, ln, term, dist) in enumerate(want["passes"]):
        g = passes[i, p]
        assert (g["rate"], g["len"], g["term"]) == (rate, ln, term)
        assert g["distortion"] == dist  # bit-exact doubles
    assert c["distortion"] == want["distortion"]


@pytest.mark.parametrize("sty", [0, 0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x09, 0x15, 0x3F])
def test_mq_encode_reversible_matches_oracle(gpu, orc, sty):
    rng = np.random.default_rng(100 + sty)
    blocks, ph, pw = _place(SHAPES, rng)
    planes = [np.zeros((ph, pw), np.int32) for _ in range(3)]
    norms = [1.732, 0.8292, 0.8292]
    want = []
    for b in blocks:
        w, h = int(b["w"]), int(b["h"])
        v = _coeffs(rng, h, w)
        if rng.random() < 0.1:
            v[:] = 0  # empty block: numbps 0, no passes
        comp, orient, level = int(rng.integers(0, 3)), int(rng.integers(0, 4)), int(rng.integers(0, 6))
        b["comp"], b["orient"], b["level"], b["cblk_sty"], b["stepsize"] = comp, orient, level, sty, 1.0
        planes[comp][b["y"]:b["y"] + h, b["x"]:b["x"] + w] = v
        smr = O.to_smr(v.astype(np.int64) << 6)
        want.append(O.orc_t1_encode(smr, orient=orient, cblksty=sty, compno=comp, level=level,
                                    qmfbid=1, stepsize=1.0, mct_norms=norms, rate_control=True))
    coded, passes, data = T.t1_encode(blocks, planes, 1, mct_norms=norms, rate_control=True)
    for i in range(len(blocks)):
        _compare_mq(coded, passes, data, i, want[i])
    # streams are packed back to back in block order
    off = 0
    for c in coded:
        assert int(c["data_off"]) == off
        off += int(c["data_len"])
    assert off == len(data)


@pytest.mark.parametrize("sty", [0, 0x01, 0x3F])
def test_mq_encode_irreversible_matches_oracle(gpu, orc, sty):
    rng = np.random.default_rng(200 + sty)
    blocks, ph, pw = _place(SHAPES, rng)
    planes = [np.zeros((ph, pw), np.float32) for _ in range(3)]
    want = []
    for b in blocks:
        w, h = int(b["w"]), int(b["h"])
        x = (rng.standard_normal((h, w)) * float(rng.choice([3.0, 40.0, 900.0]))).astype(np.float32)
        step = np.float32(rng.choice([0.0078125, 0.37109375, 1.0, 1.9501953125]))
        comp, orient, level = int(rng.integers(0, 3)), int(rng.integers(0, 4)), int(rng.integers(0, 6))
        b["comp"], b["orient"], b["level"], b["cblk_sty"], b["stepsize"] = comp, orient, level, sty, step
        planes[comp][b["y"]:b["y"] + h, b["x"]:b["x"] + w] = x
        # T1Part1::preCompress: lrintf((x / stepsize) * 64), single precision, ties to even
        q = np.rint((x / step) * np.float32(64.0)).astype(np.int64)
        want.append(O.orc_t1_encode(O.to_smr(q), orient=orient, cblksty=sty, compno=comp, level=level,
                                    qmfbid=0, stepsize=float(step), mct_norms=None, rate_control=True))
    coded, passes, data = T.t1_encode(blocks, planes, 0, rate_control=True)
    for i in range(len(blocks)):
        _compare_mq(coded, passes, data, i, want[i])


def test_mq_encode_without_rate_control(gpu, orc):
    rng = np.random.default_rng(7)
    blocks, ph, pw = _place(SHAPES[:6], rng)
    plane = np.zeros((ph, pw), np.int32)
    want = []
    for b in blocks:
        w, h = int(b["w"]), int(b["h"])
        v = _coeffs(rng, h, w)
        b["stepsize"] = 1.0
        plane[b["y"]:b["y"] + h, b["x"]:b["x"] + w] = v
        want.append(O.orc_t1_encode(O.to_smr(v.astype(np.int64) << 6), rate_control=False))
    coded, passes, data = T.t1_encode(blocks, [plane], 1, rate_control=False)
    for i in range(len(blocks)):
        _compare_mq(coded, passes, data, i, want[i])


def test_ht_encode_matches_oracle(gpu, orc):
    rng = np.random.default_rng(31)
    shapes = SHAPES + [(64, 64)] * 6 + [(1024, 4), (4, 1024), (7, 3), (2, 1), (1, 2)]
    blocks, ph, pw = _place(shapes, rng, plane_w=1100)
    plane = np.zeros((ph, pw), np.int32)
    want = []
    for b in blocks:
        w, h = int(b["w"]), int(b["h"])
        kmsbs = int(rng.integers(4, 18))
        bits = int(rng.integers(1, kmsbs + 1))
        v = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w)).astype(np.int32)
        kind = rng.random()
        if kind < 0.1:
            v[:] = 0
        elif kind < 0.4:
            v = np.where(rng.random((h, w)) < 0.8, 0, v).astype(np.int32)
        b["cblk_sty"], b["k_msbs"], b["stepsize"] = T.CBLKSTY_HT, kmsbs, 1.0
        plane[b["y"]:b["y"] + h, b["x"]:b["x"] + w] = v
        want.append(O.orc_ht_encode(O.ht_samples(v, kmsbs), kmsbs))
    coded, passes, data = T.t1_encode(blocks, [plane], 1)
    for i, c in enumerate(coded):
        assert c["failed"] == 0 and c["numbps"] == 1 and c["numpasses"] == 1
        got = bytes(data[int(c["data_off"]):int(c["data_off"]) + int(c["data_len"])])
        assert got == want[i], "block %d (%dx%d)" % (i, blocks[i]["w"], blocks[i]["h"])
        assert passes[i, 0]["rate"] == len(want[i]) and passes[i, 0]["len"] == len(want[i])


def test_mixed_coders_and_bad_arguments(gpu, orc):
    rng = np.random.default_rng(9)
    blocks, ph, pw = _place([(64, 64), (32, 32), (16, 16), (64, 64)], rng)
    plane = np.zeros((ph, pw), np.int32)
    want = []
    for i, b in enumerate(blocks):
        w, h = int(b["w"]), int(b["h"])
        v = _coeffs(rng, h, w) >> 3
        plane[b["y"]:b["y"] + h, b["x"]:b["x"] + w] = v
        b["stepsize"] = 1.0
        if i & 1:
            b["cblk_sty"], b["k_msbs"] = T.CBLKSTY_HT, 14
            want.append(O.orc_ht_encode(O.ht_samples(v, 14), 14))
        else:
            want.append(O.orc_t1_encode(O.to_smr(v.astype(np.int64) << 6), rate_control=False)["data"])
    coded, passes, data = T.t1_encode(blocks, [plane], 1)
    for i, c in enumerate(coded):
        assert bytes(data[int(c["data_off"]):int(c["data_off"]) + int(c["data_len"])]) == want[i]
    # HT on irreversible data is refused (the reference's quantiser for it is broken)
    with pytest.raises(Exception):
        T.t1_encode(blocks, [plane.view(np.float32)], 0)
    bad = blocks.copy()
    bad[0]["x"] = pw
    with pytest.raises(Exception):
        T.t1_encode(bad, [plane], 1)
    with pytest.raises(Exception):
        T.t1_encode(blocks, [plane], 1, cap=10)  # pool too small
