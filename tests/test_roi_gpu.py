"""GPU parity of the ROI dequantisers (roishift != 0): the block decoders' fused
stores against the oracle's restatement of RoiShiftFilter / RoiScaleFilter / RoiShiftHTFilter /
RoiScaleHTFilter (oracle/dequant.c, pinned to filters/PostDecompressFilters.h:7-177 by
tests/test_oracle_vs_ref.py::test_dequant_filters).  Bit-exact, 9/7 included."""
import numpy as np
import pytest

from common import same_bits
from grok_b200 import tile as T
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _mq_blocks(rng, roishift):
    """coded MQ blocks whose magnitudes straddle 2^roishift in the decoder's x2 fixed point"""
    out = []
    for (w, h) in [(64, 64), (32, 32), (17, 5), (1, 1), (64, 3), (13, 40)]:
        bits = max(2, roishift + int(rng.integers(-2, 4)))
        mag = rng.integers(0, 1 << bits, size=(h, w))
        mag = np.where(rng.random((h, w)) < 0.5, mag, mag >> max(bits - 2, 0))
        sgn = rng.integers(0, 2, size=(h, w))
        blk = ((mag.astype(np.int64) << 6) | (sgn.astype(np.int64) << 31)).astype(np.uint32).view(np.int32)
        orient = int(rng.integers(0, 4))
        e = O.orc_t1_encode(blk, orient=orient, cblksty=0, rate_control=False)
        if e["numbps"] == 0:
            continue
        sg = O.segments_from_passes(e["passes"], e["numbps"], 0)
        rc, v = O.orc_t1_decode(e["data"], sg, e["numbps"], w, h, orient=orient)
        assert rc == 0
        out.append((e["data"], sg, e["numbps"], w, h, orient, v))
    return out


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("roishift", [1, 4, 9])
def test_mq_roi_dequant(gpu, orc, reversible, roishift):
    rng = np.random.default_rng(300 + roishift)
    ents = _mq_blocks(rng, roishift)
    blocks = np.zeros(len(ents), dtype=T.CBLK_DTYPE)
    segs, chunks, off, y = [], [], 0, 0
    step = np.float32(0.6181640625)
    for i, (data, sg, numbps, w, h, orient, _v) in enumerate(ents):
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = (i * 5) % 9, y, w, h
        b["orient"], b["numbps"], b["roishift"] = orient, numbps, roishift
        b["numsegs"], b["seg_first"] = len(sg), len(segs)
        b["data_off"], b["data_len"] = off, len(data)
        b["scale"] = step / np.float32(2)  # MQ 9/7: stepsize / 2 (PostDecompressFilters.h:51)
        segs.extend(sg)
        chunks.append(data)
        off += len(data)
        y += h + 1
    pool = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    base = np.full((y, 80), 3, dtype=np.int32)
    got, rej = T.t1_decode(blocks, np.array(segs, dtype=T.SEG_DTYPE), pool, [base], reversible)
    assert rej == 0
    exp = base.copy()
    changed = 0
    for b, e in zip(blocks, ents):
        v = e[6]
        d = O.orc_dequant(0 if reversible else 1, v, roishift=roishift, stepsize=float(step))
        plain = O.orc_dequant(0 if reversible else 1, v, roishift=0, stepsize=float(step))
        changed += int((d.view(np.int32) != plain.view(np.int32)).sum())
        exp[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]] = d.view(np.int32).reshape(v.shape)
    assert changed > 0, "no sample reached the ROI threshold: the case does not test anything"
    assert same_bits(got[0], exp)


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("roishift", [3, 20, 27])
def test_ht_roi_dequant(gpu, orc, reversible, roishift):
    rng = np.random.default_rng(400 + roishift)
    ents = []
    for (w, h) in [(64, 64), (32, 32), (9, 7), (64, 2), (3, 33)]:
        kmsbs = int(rng.integers(4, 16))
        bits = int(rng.integers(1, kmsbs))
        v = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w)).astype(np.int32)
        cup = O.orc_ht_encode(O.ht_samples(v, kmsbs), kmsbs)
        rc, dec = O.orc_ht_decode(cup, kmsbs, w, h)
        assert rc == 0
        ents.append((cup, kmsbs, w, h, dec))
    blocks = np.zeros(len(ents), dtype=T.CBLK_DTYPE)
    segs, chunks, off, y = [], [], 0, 0
    step = np.float32(2.0 ** -18 * 1.375)
    for i, (data, kmsbs, w, h, _d) in enumerate(ents):
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = (i * 3) % 7, y, w, h
        b["cblk_sty"], b["k_msbs"], b["numbps"], b["roishift"] = T.CBLKSTY_HT, kmsbs, 1, roishift
        b["numsegs"], b["seg_first"] = 1, len(segs)
        b["data_off"], b["data_len"] = off, len(data)
        b["scale"] = step
        segs.append((len(data), 1))
        chunks.append(data)
        off += len(data)
        y += h + 1
    pool = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    base = np.full((y, 80), 5, dtype=np.int32)
    got, rej = T.t1_decode(blocks, np.array(segs, dtype=T.SEG_DTYPE), pool, [base], reversible)
    assert rej == 0
    exp = base.copy()
    changed = 0
    for b, e in zip(blocks, ents):
        dec = e[4]
        kind = 2 if reversible else 3
        d = O.orc_dequant(kind, dec, roishift=roishift, k_msbs=int(b["k_msbs"]), stepsize=float(step))
        plain = O.orc_dequant(kind, dec, roishift=0, k_msbs=int(b["k_msbs"]), stepsize=float(step))
        changed += int((d.view(np.int32) != plain.view(np.int32)).sum())
        exp[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]] = d.view(np.int32).reshape(dec.shape)
    if roishift < 27:
        assert changed > 0, "no sample reached the ROI threshold: the case does not test anything"
    assert same_bits(got[0], exp)
