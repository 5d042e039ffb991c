import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    """the CPU oracle (test infrastructure)"""
    from oracle import oracle as O
    O.build_oracle()
    return O


@pytest.fixture(scope="session")
def gpu():
    import grok_b200
    if grok_b200.lib.grkb_device_count() < 1:
        pytest.fail("GPU test selected but no CUDA device is visible (no CPU fallback exists)")
    return grok_b200


@pytest.fixture(scope="session")
def frame8k():
    """BASELINE configs[3] input: 8192x8192x3 16-bit.  Built from one 2048x2048 synthetic field,
    repeated with a different shift and offset per 2048-block (a full-size synth costs 25 s)."""
    import numpy as np
    from synth import synth_image
    base = synth_image(2048, 2048, 3, 16, seed=8192)
    out = []
    for c, p in enumerate(base):
        rows = []
        for j in range(4):
            rows.append(np.concatenate([np.roll(p, (97 * (i + 4 * j + c), 31 * (i + j)), axis=(0, 1)) ^ (i * 257 + j)
                                        for i in range(4)], axis=1))
        out.append(np.ascontiguousarray(np.concatenate(rows, axis=0)).astype(np.int32))
    return out
