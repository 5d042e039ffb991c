#!/usr/bin/env python3
"""MQ decode throughput against the number of frames in flight and the code-block size:
    python tools/bench_mq_inflight.py <cblk 32|64> <frames in flight>     (GRKB_T1_MQ_GROUP / _BUSY to force)
One decompressor per frame in flight, the stock encoder's 4096x2160x3 16-bit 9/7 codestream, wall clock.
The table in DESIGN.md section 4 (one block per warp vs two) comes from this."""
import os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import synth
from oracle import oracle as O
from grok_b200 import codestream as CS, tile as T
cb = int(sys.argv[1]); nfl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
img = synth.synth_image(4096, 2160, 3, 16, 7)
cs = O.ref_codec_encode(img, 16, 1, numres=6, cblk=(cb, cb), threads=os.cpu_count())
pin = T.PinnedArray((cs.size + 16,), np.uint8); pin.array[:cs.size] = cs; csb = pin.array[:cs.size]
outs = [[T.PinnedArray((2160, 4096), np.int32).array for _ in range(3)] for _ in range(nfl)]
dzs = [CS.Decompressor(0, slots=1) for _ in range(nfl)]
def run(n):
    for i in range(n):
        d = dzs[i % nfl]
        if i >= nfl:
            d.wait()
        d.submit(csb, outs[i % nfl])
    for d in dzs:
        d.wait()
run(nfl)
t = time.perf_counter(); n = 4 * nfl; run(n); dt = (time.perf_counter() - t) / n
print("cblk %d inflight %d group %s: %.2f ms/frame %.0f MP/s" % (cb, nfl, os.environ.get("GRKB_T1_MQ_GROUP"), dt * 1e3, 4096 * 2160 / dt / 1e6))
