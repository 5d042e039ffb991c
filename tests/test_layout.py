"""CPU: the host-side code-block layout (grkb_tile_layout) against the reference's own
TileComponent geometry (through oracle/_ref) and against committed golden fixtures."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from grok_b200 import tile as T
from oracle import oracle as O

CASES = [
    (0, 0, 4096, 2160, 6, 6, 6, None),
    (0, 0, 2048, 2048, 6, 6, 6, None),
    (13, 7, 113, 84, 4, 4, 5, None),
    (1, 1, 18, 2, 3, 6, 6, None),
    (0, 0, 1024, 1024, 6, 5, 5, [7, 8, 8, 8, 8, 8]),   # DCI-like: 128 at res 0, 256 above
    (1024, 2048, 2048, 3072, 6, 6, 6, None),           # tile (1,2) of an 8192^2 image
    (5, 9, 12, 22, 5, 2, 3, [3, 4, 4, 5, 5]),
    (0, 0, 1, 1, 2, 6, 6, None),
]


def _ours(c):
    x0, y0, x1, y1, nr, cw, ch, prec = c
    g = T.layout(x0, y0, x1, y1, nr, cw, ch, prec, prec)
    return [[int(v) for v in (b["x"], b["y"], b["cx0"], b["cy0"], b["w"], b["h"], b["res"],
                              b["band"], b["orient"], b["precinct"], b["cblkno"])] for b in g]


class RefGeom(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("x", "y", "cx0", "cy0", "w", "h", "res", "band", "orient", "precinct", "cblkno")]


@pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")
@pytest.mark.parametrize("case", CASES)
def test_layout_matches_reference(case):
    x0, y0, x1, y1, nr, cw, ch, prec = case
    pw = (C.c_uint32 * 33)(*([15] * 33 if prec is None else list(prec) + [15] * (33 - len(prec))))
    cap = 100000
    out = (RefGeom * cap)()
    n = O.ref().ref_layout(x0, y0, x1, y1, nr, cw, ch, pw, pw, out, cap)
    assert n >= 0
    ref = [[getattr(out[i], f[0]) for f in RefGeom._fields_] for i in range(n)]
    ours = _ours(case)
    # the reference enumerates zero-area blocks of empty precincts differently from band to
    # band; only blocks with samples matter to the coder
    ref = [r for r in ref if r[4] and r[5]]
    ours = [r for r in ours if r[4] and r[5]]
    assert ours == ref


def test_layout_golden():
    path = os.path.join(os.path.dirname(__file__), "golden", "layout.json")
    gold = json.load(open(path))
    for key, want in gold.items():
        case = json.loads(key)
        ours = [r for r in _ours(tuple(case)) if r[4] and r[5]]
        import hashlib
        digest = hashlib.sha256(json.dumps(ours).encode()).hexdigest()
        assert [len(ours), digest] == want, key
