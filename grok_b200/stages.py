"""numpy-facing wrappers of the stage-level C-ABI (include/grk_b200.h, section A).

Argument meaning follows the reference stage each call replaces; see the header for the
file:line of every one.  All work happens in libgrok_b200.so on the GPU.
"""
import ctypes as C

import numpy as np

from ._lib import lib, check


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _i32(v):
    return np.ascontiguousarray(np.array(v, dtype=np.int32))


def clamp_range(prec, sgnd):
    """clamp range of a component (point_transform/mct.cpp:556-571)"""
    if sgnd:
        return -(1 << (prec - 1)), (1 << (prec - 1)) - 1
    return 0, (1 << prec) - 1


def dwt(plane, x0, y0, numres, reversible, forward):
    """whole-tile wavelet transform of one tile-component; returns a new array (Mallat layout)"""
    a = np.ascontiguousarray(plane).copy()
    assert a.dtype in (np.int32, np.float32) and a.ndim == 2
    h, w = a.shape
    check(lib.grkb_dwt_host(_p(a), w, x0, y0, x0 + w, y0 + h, numres, int(reversible),
                            int(forward)), "grkb_dwt_host")
    return a


def mct(planes, mode, shift=(0, 0, 0), lo=(0, 0, 0), hi=(0, 0, 0)):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    s, l, h = _i32(shift), _i32(lo), _i32(hi)
    check(lib.grkb_mct_host(_p(a[0]), _p(a[1]), _p(a[2]), a[0].size, mode, _p(s), _p(l), _p(h)),
          "grkb_mct_host")
    return a


def dc_shift(plane, mode, shift, lo=0, hi=0):
    a = np.ascontiguousarray(plane).copy()
    check(lib.grkb_dc_shift_host(_p(a), a.size, mode, shift, lo, hi), "grkb_dc_shift_host")
    return a


def idwt_mct(planes, x0, y0, numres, reversible, mct_on, shift, lo, hi):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    h, w = a[0].shape
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    s, l, hh = _i32(shift), _i32(lo), _i32(hi)
    check(lib.grkb_idwt_mct_host(ptrs, len(a), w, x0, y0, x0 + w, y0 + h, numres, int(reversible),
                                 int(mct_on), _p(s), _p(l), _p(hh)), "grkb_idwt_mct_host")
    return [p.view(np.int32) for p in a]


def mct_fdwt(planes, x0, y0, numres, reversible, mct_on, shift):
    a = [np.ascontiguousarray(p, dtype=np.int32).copy() for p in planes]
    h, w = a[0].shape
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    s = _i32(shift)
    check(lib.grkb_mct_fdwt_host(ptrs, len(a), w, x0, y0, x0 + w, y0 + h, numres, int(reversible),
                                 int(mct_on), _p(s)), "grkb_mct_fdwt_host")
    return a if reversible else [p.view(np.float32) for p in a]


def mct_custom(planes, matrix, forward):
    """array-based MCT (row M7): planes float32 (inverse) or int32 (forward), matrix (nc, nc) float32"""
    a = [np.ascontiguousarray(p).copy() for p in planes]
    m = np.ascontiguousarray(matrix, dtype=np.float32)
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    check(lib.grkb_mct_custom_host(ptrs, len(a), a[0].size, _p(m), int(forward)), "grkb_mct_custom_host")
    return a
