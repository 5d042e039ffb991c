"""CPU: the oracle against the committed golden fixtures (generated from the unmodified
reference by tests/golden/make_golden.py).  Runs without /root/reference."""
import os

import numpy as np

from common import same_bits
from oracle import oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


def test_dwt_mct_golden(orc):
    d = np.load(os.path.join(G, "dwt_mct.npz"))
    n = 0
    for gi, (w, h, x0, y0, nr) in enumerate(d["geoms"]):
        for rev in (1, 0):
            p = d["dwt_in_%d_%d" % (gi, rev)]
            for fwd in (0, 1):
                key = "dwt_out_%d_%d_%d" % (gi, rev, fwd)
                if key in d:
                    assert same_bits(O.orc_dwt(p, int(x0), int(y0), int(nr), rev, fwd), d[key]), key
                    n += 1
    assert n >= 20
    for irrev in (0, 1):
        got = O.orc_mct_fwd(list(d["mct_fwd_in_%d" % irrev]), irrev)
        assert all(same_bits(a, b) for a, b in zip(got, d["mct_fwd_out_%d" % irrev]))
        inp = list(d["mct_inv_in_%d" % irrev])
        got = O.orc_mct_inv(inp, irrev, 1, [16] * 3, [0] * 3, [32768] * 3)
        assert all(same_bits(a, b) for a, b in zip(got, d["mct_inv_out_%d" % irrev]))
        got = O.orc_mct_inv(inp, irrev, 0, [12] * 3, [1] * 3, [0] * 3)
        assert all(same_bits(a, b) for a, b in zip(got, d["dc_inv_out_%d" % irrev]))


def test_t1_mq_golden(orc):
    t = np.load(os.path.join(G, "t1_mq.npz"))
    for k, (sty, w, h, orient, numbps) in enumerate(t["cases"]):
        blk = t["blk_%d" % k]
        e = O.orc_t1_encode(blk, orient=int(orient), cblksty=int(sty), compno=0, level=1, qmfbid=1,
                            stepsize=1.0)
        assert e["numbps"] == numbps
        assert e["data"] == t["bytes_%d" % k].tobytes()
        assert [list(p[:3]) for p in e["passes"]] == t["passes_%d" % k].tolist()
        assert [p[3] for p in e["passes"]] == t["dist_%d" % k].tolist()
        segs = O.segments_from_passes(e["passes"], e["numbps"], int(sty))
        rc, dec = O.orc_t1_decode(e["data"], segs, int(numbps), int(w), int(h), orient=int(orient),
                                  cblksty=int(sty))
        assert rc == 0 and np.array_equal(dec, t["dec_%d" % k])


def test_t1_ht_golden(orc):
    t = np.load(os.path.join(G, "t1_ht.npz"))
    for k, (w, h, mm) in enumerate(t["cases"]):
        rc, dec = O.orc_ht_decode(t["bytes_%d" % k].tobytes(), int(mm) - 1, int(w), int(h))
        assert rc == 0 and np.array_equal(dec, t["dec_%d" % k]), k
        # lossless: magnitude and sign come back (decoder convention: missing_msbs - 1, T1HT.cpp:128)
        mag = (dec & 0x7FFFFFFF) >> (31 - int(mm))
        val = np.where(dec >> 31, -mag.astype(np.int64), mag.astype(np.int64))
        assert np.array_equal(val, t["vals_%d" % k])
        # and the encoder reproduces the reference's bytes
        assert O.orc_ht_encode(O.ht_samples(t["vals_%d" % k], int(mm)), int(mm)) == t["bytes_%d" % k].tobytes()


def test_t1_ht_refinement_golden(orc):
    """SigProp / MagRef restatement against what the reference's block decoder returned"""
    t = np.load(os.path.join(G, "t1_ht_refine.npz"))
    for k, (w, h, kmsbs, passes, l2) in enumerate(t["cases"]):
        rc, dec = O.orc_ht_decode(t["bytes_%d" % k].tobytes(), int(kmsbs), int(w), int(h),
                                  num_passes=int(passes), lengths2=int(l2))
        assert rc == 0 and np.array_equal(dec, t["dec_%d" % k]), k
