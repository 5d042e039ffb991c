"""Device-resident timing of the wavelet kernels (development aid; bench.py is the contract).
usage: bench_dwt.py [w h numres [rev fwd ncomp mct]]"""
import ctypes as C
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import grok_b200
lib = grok_b200.lib
lib.grkb_bench_dwt.argtypes = [C.c_uint32] * 3 + [C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_uint32,
                                                 C.c_uint32, C.POINTER(C.c_float)]
PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0


def alg_bytes(w, h, nr, ncomp):
    b = 0
    for lvl in range(nr - 1):
        b += 2 * 4 * ((w + (1 << lvl) - 1) >> lvl) * ((h + (1 << lvl) - 1) >> lvl)
    return b * ncomp


a = [int(x) for x in sys.argv[1:]]
shapes = [(4096, 2160, 6), (2048, 2048, 6), (1024, 1024, 6)] if len(a) < 3 else [tuple(a[:3])]
combos = [(rev, fwd, nc, mct) for rev in (0, 1) for fwd in (0, 1) for (nc, mct) in [(1, 0), (3, 0), (3, 1)]] \
    if len(a) < 7 else [tuple(a[3:7])]
iters = a[7] if len(a) > 7 else 20
for (w, h, nr) in shapes:
    for (rev, fwd, nc, mct) in combos:
        ms = C.c_float()
        rc = lib.grkb_bench_dwt(w, h, nr, rev, fwd, nc, mct, iters, 5, C.byref(ms))
        if rc:
            print("ERR", lib.grkb_last_error()); continue
        gbs = alg_bytes(w, h, nr, nc) / (ms.value * 1e-3) / 1e9
        print(json.dumps(dict(w=w, h=h, rev=rev, fwd=fwd, ncomp=nc, mct=mct,
                              ms=round(ms.value, 4), gbs=round(gbs, 1),
                              frac=round(gbs / PEAK, 3))))
