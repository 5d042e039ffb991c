"""GPU parity: the CUDA EBCOT/MQ block decoder (through the C-ABI) against the golden fixtures
from the reference and against the CPU oracle, bit-exact; then whole tiles (T1 + dequantisation +
inverse wavelet + inverse MCT) against the oracle chain."""
import os

import numpy as np
import pytest

import codec_helpers as H
from common import same_bits
from grok_b200 import tile as T
from oracle import oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(params=["32", "16"], autouse=True)
def lanes_per_block(request):
    """every test of this file runs with one code block per warp and with one per half-warp
    (t1_mq_dec.cu: GRKB_T1_MQ_GROUP, read at every launch; without it the launcher picks by block count)"""
    old = os.environ.get("GRKB_T1_MQ_GROUP")
    os.environ["GRKB_T1_MQ_GROUP"] = request.param
    yield request.param
    if old is None:
        del os.environ["GRKB_T1_MQ_GROUP"]
    else:
        os.environ["GRKB_T1_MQ_GROUP"] = old


def _table(entries, plane_w):
    """entries: [(bytes, segs, numbps, w, h, orient, sty)] -> blocks stacked vertically in one plane"""
    blocks = np.zeros(len(entries), dtype=T.CBLK_DTYPE)
    segs, chunks, off, y = [], [], 0, 0
    for i, (data, sg, numbps, w, h, orient, sty) in enumerate(entries):
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = (i * 3) % 5, y, w, h
        b["comp"], b["orient"], b["numbps"], b["cblk_sty"] = 0, orient, numbps, sty
        b["numsegs"], b["seg_first"] = len(sg), len(segs)
        b["data_off"], b["data_len"] = off, len(data)
        b["scale"] = 0.4375
        segs.extend(sg)
        chunks.append(data)
        off += len(data)
        y += h + 1
    data = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    return blocks, np.array(segs, dtype=T.SEG_DTYPE), data, y


def _check(gpu, entries, want, reversible):
    blocks, segs, data, ph = _table(entries, 80)
    base = np.full((ph, 80), 7 if reversible else 0x40E00000, dtype=np.int32)
    got, rej = T.t1_decode(blocks, segs, data, [base], reversible)
    assert rej == 0
    exp = base.copy()
    for b, v in zip(blocks, want):
        dst = exp[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]]
        if reversible:
            dst[...] = (np.abs(v) >> 1) * np.sign(v)
        else:
            dst[...] = (v.astype(np.float32) * np.float32(0.4375)).view(np.int32)
    assert np.array_equal(got[0], exp)


@pytest.mark.parametrize("reversible", [1, 0])
def test_mq_decode_golden(gpu, reversible):
    t = np.load(os.path.join(G, "t1_mq.npz"))
    entries, want = [], []
    for k, (sty, w, h, orient, numbps) in enumerate(t["cases"]):
        passes = [tuple(p) + (0.0,) for p in t["passes_%d" % k].tolist()]
        sg = O.segments_from_passes(passes, int(numbps), int(sty))
        entries.append((t["bytes_%d" % k].tobytes(), sg, int(numbps), int(w), int(h), int(orient), int(sty)))
        want.append(t["dec_%d" % k])
    _check(gpu, entries, want, reversible)


def test_mq_decode_matches_oracle_all_modes(gpu, orc):
    rng = np.random.default_rng(21)
    entries, want = [], []
    for sty in [0, 0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x09, 0x3F]:
        for (w, h) in [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1),
                       (1, 9), (16, 256 // 16), (1024 // 16, 64)]:
            bits = int(rng.integers(1, 13))
            mag = rng.integers(0, 1 << bits, size=(h, w))
            mag = np.where(rng.random((h, w)) < 0.3, mag, mag >> max(bits - 3, 0))
            sgn = rng.integers(0, 2, size=(h, w))
            blk = ((mag.astype(np.int64) << 6) | (sgn.astype(np.int64) << 31)).astype(np.uint32).view(np.int32)
            orient = int(rng.integers(0, 4))
            e = O.orc_t1_encode(blk, orient=orient, cblksty=sty, rate_control=False)
            if e["numbps"] == 0:
                continue
            passes = e["passes"]
            if sty == 0 and rng.random() < 0.4 and len(passes) > 2:
                passes = passes[:int(rng.integers(1, len(passes)))]  # truncated block
            sg = O.segments_from_passes(passes, e["numbps"], sty)
            nb = sum(s[0] for s in sg)
            rc, v = O.orc_t1_decode(e["data"][:nb], sg, e["numbps"], w, h, orient=orient, cblksty=sty)
            assert rc == 0
            entries.append((e["data"][:nb], sg, e["numbps"], w, h, orient, sty))
            want.append(v)
    assert len(entries) > 80
    _check(gpu, entries, want, 1)
    _check(gpu, entries, want, 0)


def test_mq_decode_arbitrary_bytes(gpu, orc):
    """any byte string is a decodable MQ stream: same decisions as the oracle"""
    rng = np.random.default_rng(22)
    entries, want = [], []
    for (w, h) in [(64, 64), (7, 9), (32, 32), (64, 64)]:
        for numbps in (1, 5, 12, 20):
            raw = rng.integers(0, 256, size=int(rng.integers(1, 900)), dtype=np.uint8).tobytes()
            sg = [(len(raw), 3 * numbps - 2)]
            for sty in (0, 0x08):
                rc, v = O.orc_t1_decode(raw, sg, numbps, w, h, orient=3, cblksty=sty)
                entries.append((raw, sg, numbps, w, h, 3, sty))
                want.append(v)
    # an empty block (no segments) decodes to zeros; too many bit-planes is rejected
    entries.append((b"", [], 0, 8, 8, 0, 0))
    want.append(np.zeros((8, 8), dtype=np.int32))
    _check(gpu, entries, want, 1)
    blocks, segs, data, ph = _table([(b"\x12\x34", [(2, 1)], 27, 8, 8, 0, 0)], 80)
    got, rej = T.t1_decode(blocks, segs, data, [np.ones((ph, 80), dtype=np.int32)], 1)
    assert rej == 1 and not got[0][:8, :8].any()


@pytest.mark.parametrize("reversible,mct,ncomp,prec,w,h,nr,keep", [
    (1, 1, 3, 8, 200, 150, 4, None), (0, 1, 3, 12, 256, 135, 6, None), (1, 0, 1, 16, 97, 131, 3, None),
    (0, 0, 2, 10, 130, 70, 2, 7), (1, 1, 3, 12, 64, 64, 1, None), (0, 1, 4, 9, 90, 60, 3, None)])
def test_tile_decode_matches_oracle(gpu, orc, reversible, mct, ncomp, prec, w, h, nr, keep):
    img = H.synth_image(w, h, ncomp, prec, seed=w + h)
    if reversible or mct:
        coefs = H.forward_oracle(img[:3] if mct else img, prec, reversible, nr, mct)
        if mct and ncomp > 3:
            coefs += H.forward_oracle(img[3:], prec, reversible, nr, 0)
    else:
        # non-MCT lossy ingest is defective in the reference (TileProcessor.cpp:514-521): feed the
        # decoder coefficients of a float forward transform instead
        coefs = [O.orc_dwt((p - (1 << (prec - 1))).astype(np.float32), 0, 0, nr, 0, 1) for p in img]
    tile = H.encode_tile(coefs, prec, reversible, nr, keep_passes=keep)
    want, want_coef = H.decode_tile_oracle(tile, ncomp, prec, reversible, nr, mct)
    dec = T.TileDecoder(T.tile_params(0, 0, w, h, ncomp, nr, reversible, mct, prec))
    got = dec.decode(tile["blocks"], tile["segs"], tile["data"])
    assert dec.rejected == 0
    for c in range(ncomp):
        assert np.array_equal(got[c], want[c]), c
    if reversible and keep is None:
        for c in range(ncomp):
            assert np.array_equal(got[c], img[c])  # lossless
    # the resident re-run used for device timing reproduces the same planes
    tm = dec.time_device(2, 1)
    assert tm["launches"] >= 1 and tm["total_ms"] > 0
    again = dec.download()
    for c in range(ncomp):
        assert np.array_equal(again[c], want[c])
    dec.close()


def test_tile_decode_matches_reference_codec_golden(gpu, orc):
    """GPU tile decode == what the stock reference decoder produced for the same image
    (tests/golden/e2e.npz, generated by the reference's public C API)"""
    e = np.load(os.path.join(G, "e2e.npz"))
    for k, (w, h, prec, irr, nr) in enumerate(e["cases"]):
        w, h, prec, irr, nr = int(w), int(h), int(prec), int(irr), int(nr)
        img = H.synth_image(w, h, 3, prec, 100 + k)
        coefs = H.forward_oracle(img, prec, not irr, nr, 1)
        tile = H.encode_tile(coefs, prec, not irr, nr)
        dec = T.TileDecoder(T.tile_params(0, 0, w, h, 3, nr, not irr, 1, prec))
        got = dec.decode(tile["blocks"], tile["segs"], tile["data"])
        dec.close()
        assert np.array_equal(np.stack(got), e["decoded_%d" % k]), k


def test_simt_kernel_same_results(gpu):
    """the one-block-per-lane MQ decoder (t1_mq_dec_simt.cu, GRKB_T1_MQ_KERNEL=simt) passes the
    same parity tests; the kernel choice is read once per process, hence the subprocess"""
    import subprocess
    import sys
    if os.environ.get("GRKB_T1_MQ_KERNEL") == "simt":
        pytest.skip("already running under the SIMT kernel")
    if os.environ.get("GRKB_T1_MQ_GROUP") == "16":
        pytest.skip("once is enough")
    env = dict(os.environ, GRKB_T1_MQ_KERNEL="simt")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu",
                        "-k", "golden or all_modes or arbitrary or tile_decode_matches_oracle"],
                       env=env, capture_output=True, text=True, timeout=600,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
