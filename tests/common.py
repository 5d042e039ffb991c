"""shared test data: geometries and synthetic planes"""
import numpy as np

# (w, h, x0, y0, numres): odd sizes, width/height 1..5, odd origins (parity 1), ragged strides
GEOMS = [
    (64, 64, 0, 0, 3), (33, 17, 0, 0, 4), (1, 5, 0, 0, 3), (5, 1, 1, 1, 3), (3, 2, 0, 0, 2),
    (100, 77, 13, 7, 6), (7, 13, 5, 9, 5), (256, 135, 0, 0, 6), (17, 1, 0, 1, 3), (1, 17, 1, 0, 2),
    (2, 2, 0, 0, 2), (4, 4, 1, 1, 3), (5, 5, 3, 3, 2), (250, 31, 0, 0, 3), (241, 40, 1, 0, 3),
    (481, 66, 2, 3, 4), (512, 270, 0, 0, 6), (37, 300, 0, 1, 5), (8, 8, 0, 0, 4), (9, 300, 1, 1, 6),
    # widths that are multiples of 8 at even x0: the levels the strip kernel (idwt_strip.cu) takes,
    # several strips wide, odd/even heights and origins, chunk boundaries inside the plane
    (1024, 333, 0, 0, 5), (960, 270, 0, 1, 4), (2048, 67, 0, 0, 3), (128, 515, 0, 3, 2),
    (1920, 130, 0, 0, 3), (64, 2, 0, 0, 2), (72, 1031, 0, 1, 2),
]
# Degenerate geometries on which the REFERENCE itself misbehaves (documented in DESIGN.md):
#  * inverse 5/3, a level of height exactly 2 at odd y0 writes dest[1] instead of dest[stride]
#    (WaveletReverse.cpp:639-649);
#  * forward, a resolution of zero width and non-zero height reads/writes out of range
#    (WaveletFwd.cpp:909-957).
REF_BROKEN = [(2, 2, 1, 1, 2), (3, 3, 1, 0, 6), (1, 17, 1, 0, 3)]


def ref_defect(w, h, x0, y0, nr, reversible, forward):
    """True when the REFERENCE itself misbehaves on this geometry (see REF_BROKEN above), so the
    oracle -- which implements the transform the standard defines -- cannot be pinned on it."""
    for lvl in range(nr - 1):
        rx0, rx1 = -((-x0) >> lvl), -((-(x0 + w)) >> lvl)
        ry0, ry1 = -((-y0) >> lvl), -((-(y0 + h)) >> lvl)
        rw, rh = rx1 - rx0, ry1 - ry0
        if not forward and reversible and rh == 2 and (ry0 & 1) and rw > 0:
            return True
        if forward and (rw == 0) != (rh == 0):
            return True
    return False


def plane(rng, w, h, reversible, amp=2000):
    if reversible:
        return rng.integers(-amp, amp, size=(h, w)).astype(np.int32)
    return (rng.standard_normal((h, w)) * amp / 4).astype(np.float32)


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.int32),
                          np.ascontiguousarray(b).view(np.int32))


def ht_refine_case(rng, w, h, ht_encode):
    """One HT block with a refinement segment: (samples, k_msbs, cleanup bytes + second segment,
    length of the second segment).  The cleanup segment comes from `ht_encode` (oracle or reference
    encoder, they are byte-identical); the second segment is arbitrary bytes -- random, 0xFF-heavy
    or drawn from the bytes that trigger the stuffing rules -- sometimes shorter than the passes
    need, so the zero fill at the end of both readers is exercised too."""
    from oracle import oracle as O
    kmsbs = int(rng.integers(3, 20))
    bits = int(rng.integers(1, kmsbs))
    v = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w)).astype(np.int32)
    dens = rng.choice([0.02, 0.1, 0.3, 0.7, 1.0])
    v = np.where(rng.random((h, w)) < dens, v, 0).astype(np.int32)
    cup = ht_encode(O.ht_samples(v, kmsbs), kmsbs)
    kind = rng.random()
    l2 = int(rng.integers(1, 1 + max(2, w * h // 4))) if kind < 0.8 else int(rng.integers(1, 6))
    if kind < 0.5:
        seg = rng.integers(0, 256, size=l2, dtype=np.uint8)
    elif kind < 0.7:
        seg = np.where(rng.random(l2) < 0.5, 0xFF, rng.integers(0, 256, size=l2)).astype(np.uint8)
    else:
        seg = rng.choice(np.array([0x7F, 0xFF, 0x8F, 0x90, 0x00, 0x80], dtype=np.uint8), size=l2)
    return v, kmsbs, bytes(cup) + bytes(seg), l2
