"""Deterministic synthetic frames (numpy only: importable without the product library)."""
import numpy as np


def synth_image(w, h, ncomp, prec, seed):
    """deterministic 'natural-like' planes: sum of random low-frequency cosines + noise"""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    out = []
    for c in range(ncomp):
        f = np.zeros((h, w), dtype=np.float32)
        for k in range(6):
            fx, fy = rng.uniform(0.5, 6.0, size=2) * (k + 1)
            ph = rng.uniform(0, 2 * np.pi)
            f += np.cos(2 * np.pi * (fx * xx / w + fy * yy / h) + ph).astype(np.float32) / (k + 1)
        f = (f - f.min()) / (f.max() - f.min())
        f = f * ((1 << prec) - 1) + rng.normal(0, (1 << prec) / 1024.0, size=(h, w))
        out.append(np.clip(np.rint(f), 0, (1 << prec) - 1).astype(np.int32))
    return out
