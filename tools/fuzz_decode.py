#!/usr/bin/env python3
"""Malformed codestreams through the GPU decoder: valid streams of the stock encoder with bytes flipped,
spliced out or cut off.  Nothing to compare with -- the run must neither crash nor hang, and the device must
still decode a valid stream correctly afterwards.   python tools/fuzz_decode.py <first seed> <last seed>"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import codec_helpers as H  # noqa: E402
import test_random_gpu as R  # noqa: E402
from grok_b200 import codestream as CS  # noqa: E402
from grok_b200._lib import GrokB200Error  # noqa: E402

lo, hi = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(lo)
dec = CS.Decompressor(0, slots=2)
decoded = refused = 0
for seed in range(lo, hi):
    if seed in R.STOCK_ABORTS:
        continue
    c = R.draw(seed)
    img = H.synth_image(c["w"], c["h"], c["ncomp"], c["prec"], seed=seed)
    try:
        cs = R.stock_encode(img, c, R.stock_opts(c, "rates=40:10" if not c["ht"] else ""))
    except BaseException:
        continue
    if cs.size < 200 or not R.consistent(cs):
        continue
    good = dec.decode(cs)
    for trial in range(12):
        b = cs.copy()
        kind = trial % 3
        if kind == 0:
            b = b[:int(rng.integers(150, b.size))]
        elif kind == 1:
            for _ in range(int(rng.integers(1, 12))):
                b[int(rng.integers(120, b.size))] = int(rng.integers(0, 256))
        else:
            a, d = sorted(rng.integers(150, b.size, size=2).tolist())
            b = np.concatenate([b[:a], b[d:]])
        print(seed, trial, flush=True)
        try:
            dec.decode(b)
            decoded += 1
        except GrokB200Error:
            refused += 1
    again = dec.decode(cs)
    assert all(np.array_equal(a, b_) for a, b_ in zip(good, again)), "the device decodes differently after malformed input"
dec.close()
print("malformed streams: %d decoded to something, %d refused; valid streams still decode identically" % (decoded, refused))
