"""The drop-in boundary without a GPU: the C-ABI library loads, exports every function
include/grk_b200.h declares, its PODs have the documented sizes, compute entry points fail loudly
(no CPU fallback) when there is no device, and the plugin exports the symbols Grok's loader
looks up (plugin/plugin_interface.h, plugin/minpf_plugin_manager.cpp)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import grok_b200
from grok_b200 import tile as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "grk_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(grkb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_function_is_exported():
    names = _declared()
    assert len(names) >= 25
    lib = C.CDLL(os.path.join(ROOT, "grok_b200", "libgrok_b200.so"))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_pod_sizes():
    assert T.CBLK_DTYPE.itemsize == 36 and T.SEG_DTYPE.itemsize == 8
    assert T.ENC_CBLK_DTYPE.itemsize == 32 and T.PASS_DTYPE.itemsize == 16
    assert T.CODED_DTYPE.itemsize == 24 and T.GEOM_DTYPE.itemsize == 32


def test_version_and_error_text():
    assert grok_b200.lib.grkb_version() >= 1
    grok_b200.lib.grkb_last_error.restype = C.c_char_p
    assert isinstance(grok_b200.lib.grkb_last_error(), bytes)


def test_no_cpu_fallback():
    """without a CUDA device every compute entry point must fail, never compute on the host"""
    if grok_b200.lib.grkb_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    a = np.arange(64 * 64, dtype=np.int32).reshape(64, 64)
    before = a.copy()
    rc = grok_b200.lib.grkb_dwt_host(a.ctypes.data, 64, 0, 0, 64, 64, 3, 1, 0)
    assert rc != 0 and np.array_equal(a, before)
    with pytest.raises(Exception):
        T.TileDecoder(T.tile_params(0, 0, 64, 64, 1, 3, 1, 0, 8))
    blocks = np.zeros(1, dtype=T.ENC_CBLK_DTYPE)
    blocks[0]["w"] = blocks[0]["h"] = 8
    blocks[0]["stepsize"] = 1.0
    with pytest.raises(Exception):
        T.t1_encode(blocks, [a], 1)


PLUGIN_SYMBOLS = [
    "minpf_post_load_plugin", "plugin_init", "plugin_encode", "plugin_batch_encode",
    "plugin_is_batch_complete", "plugin_stop_batch_encode", "plugin_decompress",
    "plugin_init_batch_decompress", "plugin_batch_decompress", "plugin_stop_batch_decompress",
    "plugin_get_debug_state", "plugin_debug_mqc_next_cxd", "plugin_debug_mqc_next_plane",
    "create", "destroy", "exit_func",
]


def test_plugin_exports_the_symbols_the_host_looks_up():
    path = os.path.join(ROOT, "grok_b200", "plugin", "_build", "libgrokj2k_plugin.so")
    if not os.path.exists(path):
        pytest.skip("plugin not built (needs a Grok source tree at build time)")
    lib = C.CDLL(path)
    missing = [n for n in PLUGIN_SYMBOLS if not hasattr(lib, n)]
    assert not missing, missing
    lib.plugin_get_debug_state.restype = C.c_uint32
    assert lib.plugin_get_debug_state() == 0  # GRK_PLUGIN_STATE_NO_DEBUG
