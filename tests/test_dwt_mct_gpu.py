"""GPU parity: CUDA wavelet / MCT kernels (through the C-ABI) against the CPU oracle.
Bit-exact for 5/3, RCT and all integer work; 9/7 and ICT are compared bit-for-bit as well
(tolerance 0 ULP: the kernels use the reference's operation order with no FMA contraction)."""
import numpy as np
import pytest

from common import GEOMS, plane, same_bits

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("forward", [0, 1])
def test_dwt_matches_oracle(gpu, orc, reversible, forward):
    rng = np.random.default_rng(7)
    for (w, h, x0, y0, nr) in GEOMS:
        p = plane(rng, w, h, reversible)
        want = orc.orc_dwt(p, x0, y0, nr, reversible, forward)
        got = gpu.stages.dwt(p, x0, y0, nr, reversible, forward)
        assert same_bits(want, got), (w, h, x0, y0, nr, reversible, forward)


@pytest.mark.parametrize("irrev", [0, 1])
def test_mct_matches_oracle(gpu, orc, irrev):
    rng = np.random.default_rng(3)
    for n in [(1, 1), (3, 5), (37, 53), (64, 64), (211, 97)]:
        pl = [rng.integers(0, 65536, size=n).astype(np.int32) for _ in range(3)]
        sh = [32768, 32768, 32768]
        pre = [p - 32768 for p in pl]
        want = orc.orc_mct_fwd(pre, irrev)
        got = gpu.stages.mct(pl, 3 if irrev else 2, shift=sh)
        assert all(same_bits(a, b) for a, b in zip(want, got))
        if irrev:
            inp = [(rng.standard_normal(n) * 20000).astype(np.float32).view(np.int32)
                   for _ in range(3)]
        else:
            inp = [rng.integers(-40000, 40000, size=n).astype(np.int32) for _ in range(3)]
        want = orc.orc_mct_inv(inp, irrev, 1, [16] * 3, [0] * 3, sh)
        got = gpu.stages.mct(inp, 1 if irrev else 0, shift=sh, lo=[0] * 3, hi=[65535] * 3)
        assert all(same_bits(a, b) for a, b in zip(want, got))
        want = orc.orc_mct_inv(inp[:1], irrev, 0, [12], [1], [0])
        got = gpu.stages.dc_shift(inp[0], 1 if irrev else 0, 0, -2048, 2047)
        assert same_bits(want[0], got)


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("mct_on", [1, 0])
def test_fused_idwt_mct(gpu, orc, reversible, mct_on):
    """inverse wavelet with the component transform fused into the last level == oracle stages"""
    rng = np.random.default_rng(11)
    for (w, h, x0, y0, nr) in GEOMS + [(40, 30, 0, 0, 1)]:
        pl = [plane(rng, w, h, reversible, amp=30000) for _ in range(3)]
        sh, lo, hi = [32768] * 3, [0] * 3, [65535] * 3
        stage = [orc.orc_dwt(p, x0, y0, nr, reversible, 0) for p in pl]
        want = orc.orc_mct_inv([s.view(np.int32) for s in stage], 0 if reversible else 1, mct_on,
                               [16] * 3, [0] * 3, sh)
        got = gpu.stages.idwt_mct(pl, x0, y0, nr, reversible, mct_on, sh, lo, hi)
        assert all(same_bits(a, b) for a, b in zip(want, got)), (w, h, x0, y0, nr)


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("mct_on", [1, 0])
def test_fused_mct_fdwt(gpu, orc, reversible, mct_on):
    rng = np.random.default_rng(12)
    for (w, h, x0, y0, nr) in GEOMS + [(40, 30, 0, 0, 1)]:
        if not reversible and not mct_on:
            continue  # reference defect: (x-shift)*2048 misread as float (TileProcessor.cpp:514-521)
        pl = [rng.integers(0, 4096, size=(h, w)).astype(np.int32) for _ in range(3)]
        sh = [2048] * 3
        pre = [p - 2048 for p in pl]
        if mct_on:
            pre = orc.orc_mct_fwd(pre, 0 if reversible else 1)
        pre = [p if reversible else p.view(np.float32) for p in pre]
        want = [orc.orc_dwt(p, x0, y0, nr, reversible, 1) for p in pre]
        got = gpu.stages.mct_fdwt(pl, x0, y0, nr, reversible, mct_on, sh)
        assert all(same_bits(a, b) for a, b in zip(want, got)), (w, h, x0, y0, nr)


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("forward", [0, 1])
def test_full_size_vs_oracle(gpu, orc, reversible, forward):
    """BASELINE shape B (4096x2160, 5 levels), one component: both directions against the oracle
    (423-CTA strip grids, every level on the bulk-copy path), bit-exact, 9/7 included."""
    rng = np.random.default_rng(4096 + reversible * 2 + forward)
    w, h, nr = 4096, 2160, 6
    p = plane(rng, w, h, reversible, amp=30000)
    want = orc.orc_dwt(p, 0, 0, nr, reversible, forward)
    got = gpu.stages.dwt(p, 0, 0, nr, reversible, forward)
    assert same_bits(want, got)


@pytest.mark.parametrize("reversible", [1, 0])
def test_full_size_fused_idwt_mct_vs_oracle(gpu, orc, reversible):
    """the decoder back end of shape B as the frame runs it: three components per launch, inverse
    RCT / ICT + DC shift + clamp fused into the last level"""
    rng = np.random.default_rng(2160 + reversible)
    w, h, nr, prec = 4096, 2160, 6, 12
    pl = [plane(rng, w, h, reversible, amp=3000) for _ in range(3)]
    sh, lo, hi = [1 << (prec - 1)] * 3, [0] * 3, [(1 << prec) - 1] * 3
    stage = [orc.orc_dwt(p, 0, 0, nr, reversible, 0) for p in pl]
    want = orc.orc_mct_inv([s.view(np.int32) for s in stage], 0 if reversible else 1, 1, [prec] * 3,
                           [0] * 3, sh)
    got = gpu.stages.idwt_mct(pl, 0, 0, nr, reversible, 1, sh, lo, hi)
    assert all(np.array_equal(a, b) for a, b in zip(want, got))


@pytest.mark.parametrize("nc", [1, 3, 4])
def test_custom_mct_matches_oracle(gpu, orc, nc):
    """row M7: the array-based MCT of Part 2, both directions, bit-exact"""
    rng = np.random.default_rng(170 + nc)
    m = (rng.standard_normal((nc, nc)) * 0.7).astype(np.float32)
    for n in (1, 5, 4099):
        ints = [rng.integers(-40000, 40000, size=n).astype(np.int32) for _ in range(nc)]
        for a, b in zip(orc.orc_mct_custom(ints, m, 1), gpu.stages.mct_custom(ints, m, 1)):
            assert np.array_equal(a, b)
        flts = [(rng.standard_normal(n) * 3000).astype(np.float32) for _ in range(nc)]
        for a, b in zip(orc.orc_mct_custom(flts, m, 0), gpu.stages.mct_custom(flts, m, 0)):
            assert same_bits(a, b)


@pytest.mark.parametrize("seed", range(24))
def test_dwt_random_geometries(gpu, orc, seed):
    """drawn tile-component rectangles, half of them of the kind the strip kernels take (widths that are
    multiples of 8 times a power of two at x0 = 0), both filters, both directions, fused and stage by stage"""
    rng = np.random.default_rng(1000 + seed)
    nr = int(rng.integers(1, 7))
    if seed % 2:
        w = int(rng.integers(1, 40)) * 8 << int(rng.integers(0, 4))
        x0 = 0
    else:
        w, x0 = int(rng.integers(1, 700)), int(rng.integers(0, 40))
    h, y0 = int(rng.integers(1, 400)), int(rng.integers(0, 40))
    if min(h >> (nr - 1), 3) == 2 or (w >> (nr - 1)) == 0 or (h >> (nr - 1)) == 0:
        nr = 1 + int(rng.integers(0, 2))  # (stay clear of the reference's degenerate-geometry defects, common.py)
    for reversible in (1, 0):
        p = plane(rng, w, h, reversible)
        for forward in (0, 1):
            want = orc.orc_dwt(p, x0, y0, nr, reversible, forward)
            got = gpu.stages.dwt(p, x0, y0, nr, reversible, forward)
            assert same_bits(want, got), (w, h, x0, y0, nr, reversible, forward)
        pl = [plane(rng, w, h, reversible, amp=30000) for _ in range(3)]
        sh, lo, hi = [32768] * 3, [0] * 3, [65535] * 3
        stage = [orc.orc_dwt(q, x0, y0, nr, reversible, 0) for q in pl]
        want = orc.orc_mct_inv([s_.view(np.int32) for s_ in stage], 0 if reversible else 1, 1, [16] * 3, [0] * 3, sh)
        got = gpu.stages.idwt_mct(pl, x0, y0, nr, reversible, 1, sh, lo, hi)
        assert all(same_bits(a, b) for a, b in zip(want, got)), (w, h, x0, y0, nr, reversible, "fused inverse")
