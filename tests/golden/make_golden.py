"""Generates the committed golden fixtures from the UNMODIFIED reference (oracle/_ref/
libgrok_ref.so, compiled from /root/reference by oracle/Makefile).  Run in the build container:
    python tests/golden/make_golden.py
Outputs (small, committed): layout.json, dwt_mct.npz, t1_mq.npz"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402
from common import ref_defect  # noqa: E402
import test_layout as TL  # noqa: E402

assert O.have_ref(), "build the reference first: make -C oracle ref"

# ---- layout -------------------------------------------------------------------------------
gold = {}
for case in TL.CASES:
    x0, y0, x1, y1, nr, cw, ch, prec = case
    pw = (C.c_uint32 * 33)(*([15] * 33 if prec is None else list(prec) + [15] * (33 - len(prec))))
    out = (TL.RefGeom * 100000)()
    n = O.ref().ref_layout(x0, y0, x1, y1, nr, cw, ch, pw, pw, out, 100000)
    ref = [[getattr(out[i], f[0]) for f in TL.RefGeom._fields_] for i in range(n)]
    ref = [r for r in ref if r[4] and r[5]]
    gold[json.dumps(list(case))] = [len(ref), hashlib.sha256(json.dumps(ref).encode()).hexdigest()]
json.dump(gold, open(os.path.join(HERE, "layout.json"), "w"), indent=1)

# ---- wavelet + MCT ------------------------------------------------------------------------
rng = np.random.default_rng(2024)
d = {}
geoms = [(33, 17, 0, 0, 4), (100, 77, 13, 7, 6), (7, 13, 5, 9, 5), (64, 64, 0, 0, 3), (5, 1, 1, 1, 3),
         (1, 17, 1, 0, 2), (128, 68, 0, 0, 6)]
d["geoms"] = np.array(geoms, dtype=np.int32)
for gi, (w, h, x0, y0, nr) in enumerate(geoms):
    for rev in (1, 0):
        p = (rng.integers(-2000, 2000, size=(h, w)).astype(np.int32) if rev else
             (rng.standard_normal((h, w)) * 500).astype(np.float32))
        d["dwt_in_%d_%d" % (gi, rev)] = p
        for fwd in (0, 1):
            if ref_defect(w, h, x0, y0, nr, rev, fwd):
                continue
            d["dwt_out_%d_%d_%d" % (gi, rev, fwd)] = O.ref_dwt(p, x0, y0, nr, rev, fwd)
for irrev in (0, 1):
    pl = [rng.integers(-32768, 32768, size=(37, 53)).astype(np.int32) for _ in range(3)]
    d["mct_fwd_in_%d" % irrev] = np.stack(pl)
    d["mct_fwd_out_%d" % irrev] = np.stack([a.view(np.int32) for a in O.ref_mct_fwd(pl, irrev)])
    inp = ([(rng.standard_normal((37, 53)) * 20000).astype(np.float32).view(np.int32) for _ in range(3)]
           if irrev else [rng.integers(-40000, 40000, size=(37, 53)).astype(np.int32) for _ in range(3)])
    d["mct_inv_in_%d" % irrev] = np.stack(inp)
    d["mct_inv_out_%d" % irrev] = np.stack(O.ref_mct_inv(inp, irrev, 1, [16] * 3, [0] * 3, [32768] * 3))
    d["dc_inv_out_%d" % irrev] = np.stack(O.ref_mct_inv(inp, irrev, 0, [12] * 3, [1] * 3, [0] * 3))
np.savez_compressed(os.path.join(HERE, "dwt_mct.npz"), **d)

# ---- EBCOT / MQ ----------------------------------------------------------------------------
t = {}
cases = []
k = 0
for sty in (0, 0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x3F, 0x05):
    for (w, h) in [(64, 64), (32, 32), (17, 5), (1, 9), (64, 3)]:
        bits = int(rng.integers(2, 12))
        mag = rng.integers(0, 1 << bits, size=(h, w))
        mag = np.where(rng.random((h, w)) < 0.3, mag, mag >> max(bits - 3, 0))
        sgn = rng.integers(0, 2, size=(h, w))
        blk = ((mag.astype(np.int64) << 6) | (sgn.astype(np.int64) << 31)).astype(np.uint32).view(np.int32)
        orient = int(rng.integers(0, 4))
        e = O.ref_t1_encode(blk, orient=orient, cblksty=sty, compno=0, level=1, qmfbid=1, stepsize=1.0)
        if e["numbps"] == 0:
            continue
        segs = O.segments_from_passes(e["passes"], e["numbps"], sty)
        rc, dec = O.ref_t1_decode(e["data"], segs, e["numbps"], w, h, orient=orient, cblksty=sty)
        assert rc == 0
        t["blk_%d" % k] = blk
        t["bytes_%d" % k] = np.frombuffer(e["data"], dtype=np.uint8)
        t["passes_%d" % k] = np.array([(p[0], p[1], p[2]) for p in e["passes"]], dtype=np.uint32)
        t["dist_%d" % k] = np.array([p[3] for p in e["passes"]], dtype=np.float64)
        t["dec_%d" % k] = dec
        cases.append((sty, w, h, orient, e["numbps"]))
        k += 1
t["cases"] = np.array(cases, dtype=np.int32)
np.savez_compressed(os.path.join(HERE, "t1_mq.npz"), **t)
print("golden fixtures written:", os.listdir(HERE))

# ---- end to end: the stock codec (public C API) on small synthetic images --------------------
import codec_helpers as H  # noqa: E402
e = {}
e2e_cases = [(96, 80, 8, 0, 4), (120, 67, 12, 1, 5), (64, 90, 16, 1, 3), (150, 40, 12, 0, 6)]
e["cases"] = np.array(e2e_cases, dtype=np.int32)
for k, (w, h, prec, irr, nr) in enumerate(e2e_cases):
    img = H.synth_image(w, h, 3, prec, 100 + k)
    cs = O.ref_codec_encode(img, prec, irr, numres=nr)
    e["decoded_%d" % k] = np.stack(O.ref_codec_decode(cs, 3, w, h)).astype(np.int32)
np.savez_compressed(os.path.join(HERE, "e2e.npz"), **e)
print("e2e fixture written")

# ---- HTJ2K block coder: reference encoder output and reference decoder output ---------------
hh = {}
ht_cases = []
k = 0
for (w, h) in [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9),
               (2, 2), (7, 2), (128, 32), (6, 6)]:
    for bits, dens in [(1, 0.05), (3, 0.5), (7, 0.3), (11, 0.9), (12, 0.02)]:
        mm = int(rng.integers(bits + 1, 30 - bits + 1)) if 30 - bits > bits + 1 else bits + 1
        vals = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w))
        vals = np.where(rng.random((h, w)) < dens, vals, 0)
        enc = O.ref_ht_encode(O.ht_samples(vals, mm), mm)
        rc, dec = O.ref_ht_decode(enc, mm - 1, w, h)
        assert rc == 0
        hh["vals_%d" % k] = vals.astype(np.int32)
        hh["bytes_%d" % k] = np.frombuffer(enc, dtype=np.uint8)
        hh["dec_%d" % k] = dec
        ht_cases.append((w, h, mm))
        k += 1
hh["cases"] = np.array(ht_cases, dtype=np.int32)
np.savez_compressed(os.path.join(HERE, "t1_ht.npz"), **hh)
print("HT fixture written", k)
