"""Golden fixture for the HT SigProp / MagRef passes, from the UNMODIFIED reference:
cleanup segments come from the reference's HT encoder, the second segment is arbitrary bytes (the
reference has no encoder for these passes; every byte string is a valid raw-bit segment), and the
expected samples are what the reference's ojph_decode_codeblock returns when called directly with
both segment lengths (oracle/ref_harness.cpp ref_ht_decode).  Run in the build container:
    python tests/golden/make_golden_ht_refine.py
Output (small, committed): t1_ht_refine.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle as O  # noqa: E402
from common import ht_refine_case  # noqa: E402

assert O.have_ref(), "build the reference first: make -C oracle ref"
rng = np.random.default_rng(77)
out, cases = {}, []
shapes = [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9), (2, 2),
          (7, 2), (128, 32), (6, 6), (64, 5), (64, 6), (64, 7), (33, 9), (9, 33), (40, 21)]
k = 0
for (w, h) in shapes:
    for passes in (2, 3):
        vals, kmsbs, data, l2 = ht_refine_case(rng, w, h, O.ref_ht_encode)
        rc, dec = O.ref_ht_decode(data, kmsbs, w, h, num_passes=passes, lengths2=l2)
        assert rc == 0
        out["bytes_%d" % k] = np.frombuffer(data, dtype=np.uint8)
        out["dec_%d" % k] = dec
        cases.append((w, h, kmsbs, passes, l2))
        k += 1
out["cases"] = np.array(cases, dtype=np.int32)
np.savez_compressed(os.path.join(HERE, "t1_ht_refine.npz"), **out)
print("HT refinement fixture written", k)
