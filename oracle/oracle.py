"""ctypes doors to the CPU oracle (liboracle.so) and the compiled reference (_ref/libgrok_ref.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the cpu_baseline /
--impl reference legs of bench.py.  The product (grok_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE = None
_REF = None


class Rect(C.Structure):
    _fields_ = [("x0", C.c_uint32), ("y0", C.c_uint32), ("x1", C.c_uint32), ("y1", C.c_uint32)]


class OrcPass(C.Structure):
    _fields_ = [("rate", C.c_uint32), ("len", C.c_uint32), ("term", C.c_uint8),
                ("distortiondec", C.c_double)]


class OrcEncResult(C.Structure):
    _fields_ = [("numbps", C.c_uint32), ("numPassesTotal", C.c_uint32),
                ("passes", OrcPass * 100), ("num_bytes", C.c_uint32)]


class RefPass(C.Structure):
    _fields_ = [("rate", C.c_uint32), ("len", C.c_uint32), ("term", C.c_uint32),
                ("distortiondec", C.c_double)]


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])


def oracle():
    global _ORACLE
    if _ORACLE is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        _ORACLE = C.CDLL(path)
        _ORACLE.orc_band_rect.restype = Rect
        _ORACLE.orc_band_rect.argtypes = [Rect, C.c_uint32, C.c_uint32]
        _ORACLE.orc_stride.restype = C.c_uint32
        if hasattr(_ORACLE, "orc_t1_encode"):
            _ORACLE.orc_t1_encode.restype = C.c_double
    return _ORACLE


def have_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libgrok_ref.so"))


def ref():
    global _REF
    if _REF is None:
        _REF = C.CDLL(os.path.join(HERE, "_ref", "libgrok_ref.so"))
        _REF.ref_t1_encode.restype = C.c_double
        _REF.ref_stride.restype = C.c_uint32
    return _REF


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


# ---------------------------------------------------------------- wavelets

def _dwt(lib_is_ref, plane, x0, y0, numres, qmfbid, forward):
    """plane: 2-D int32 (5/3) or float32 (9/7) array; returns transformed copy (Mallat layout)."""
    h, w = plane.shape
    a = np.ascontiguousarray(plane).copy()
    if lib_is_ref:
        rc = ref().ref_dwt(_p(a), C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x0 + w),
                           C.c_uint32(y0 + h), C.c_uint32(numres), C.c_uint32(qmfbid),
                           C.c_int(forward))
        assert rc == 0
    else:
        fn = {(1, 0): "orc_idwt53", (0, 0): "orc_idwt97", (1, 1): "orc_fdwt53",
              (0, 1): "orc_fdwt97"}[(qmfbid, forward)]
        getattr(oracle(), fn)(_p(a), C.c_uint32(w), Rect(x0, y0, x0 + w, y0 + h),
                              C.c_uint32(numres))
    return a


def orc_dwt(plane, x0, y0, numres, qmfbid, forward):
    return _dwt(False, plane, x0, y0, numres, qmfbid, forward)


def ref_dwt(plane, x0, y0, numres, qmfbid, forward):
    return _dwt(True, plane, x0, y0, numres, qmfbid, forward)


# ---------------------------------------------------------------- MCT

def orc_mct_fwd(planes, irrev):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    fn = oracle().orc_ict_fwd if irrev else oracle().orc_rct_fwd
    fn(_p(a[0]), _p(a[1]), _p(a[2]), C.c_uint64(a[0].size))
    return a


def ref_mct_fwd(planes, irrev):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    ref().ref_mct_fwd(_p(a[0]), _p(a[1]), _p(a[2]), C.c_uint64(a[0].size), C.c_int(irrev))
    return a


def _range(prec, sgnd):
    if sgnd:
        return -(1 << (prec - 1)), (1 << (prec - 1)) - 1
    return 0, (1 << prec) - 1


def orc_mct_inv(planes, irrev, mct_on, prec, sgnd, shift):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    lo = np.array([_range(p, s)[0] for p, s in zip(prec, sgnd)], dtype=np.int32)
    hi = np.array([_range(p, s)[1] for p, s in zip(prec, sgnd)], dtype=np.int32)
    sh = np.array(shift, dtype=np.int32)
    n = C.c_uint64(a[0].size)
    if mct_on:
        fn = oracle().orc_ict_inv if irrev else oracle().orc_rct_inv
        fn(_p(a[0]), _p(a[1]), _p(a[2]), n, _p(sh), _p(lo), _p(hi))
    else:
        fn = oracle().orc_dc_shift_inv_irrev if irrev else oracle().orc_dc_shift_inv_rev
        for c, p in enumerate(a):
            fn(_p(p), n, C.c_int32(int(sh[c])), C.c_int32(int(lo[c])), C.c_int32(int(hi[c])))
    return a


def ref_mct_inv(planes, irrev, mct_on, prec, sgnd, shift):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    h, w = a[0].shape
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    rc = ref().ref_mct_inv(ptrs, C.c_uint32(len(a)), C.c_uint32(w), C.c_uint32(h), C.c_int(irrev),
                           C.c_int(mct_on), _p(np.array(prec, dtype=np.uint8)),
                           _p(np.array(sgnd, dtype=np.uint8)), _p(np.array(shift, dtype=np.int32)))
    assert rc == 0
    return a
