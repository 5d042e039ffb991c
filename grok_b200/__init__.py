"""grok_b200 -- B200 (sm_100a) accelerator for Grok's JPEG 2000 tile-processing hot path.

The product is the C-ABI shared library grok_b200/libgrok_b200.so (CUDA kernels + host glue,
sources under grok_b200/csrc) and the Grok plugin built on it.  This Python package is only a
thin ctypes binding of that C-ABI, used by the tests and by bench.py; it contains no codec
arithmetic and has no CPU fallback: if the library is missing, importing `lib` raises.
"""
from ._lib import lib, LIB_PATH, GrokB200Error  # noqa: F401
from . import stages  # noqa: F401
from . import tile  # noqa: F401
