"""Loads grok_b200/libgrok_b200.so and declares the C-ABI of include/grk_b200.h."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgrok_b200.so")


class GrokB200Error(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise GrokB200Error(
        "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU fallback)" % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_u32, _i32, _u64, _vp, _int = C.c_uint32, C.c_int32, C.c_uint64, C.c_void_p, C.c_int
_pi32 = C.POINTER(C.c_int32)

lib.grkb_version.restype = _int
lib.grkb_device_count.restype = _int
lib.grkb_set_device.argtypes = [_int]
lib.grkb_last_error.restype = C.c_char_p
lib.grkb_dwt_host.argtypes = [_vp, _u32, _u32, _u32, _u32, _u32, _u32, _int, _int]
lib.grkb_mct_host.argtypes = [_vp, _vp, _vp, _u64, _int, _vp, _vp, _vp]
lib.grkb_dc_shift_host.argtypes = [_vp, _u64, _int, _i32, _i32, _i32]
lib.grkb_idwt_mct_host.argtypes = [_vp, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _int, _int,
                                   _vp, _vp, _vp]
lib.grkb_mct_fdwt_host.argtypes = [_vp, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _int, _int, _vp]


def check(rc, what):
    if rc != 0:
        raise GrokB200Error("%s failed (%d): %s" % (what, rc, lib.grkb_last_error().decode()))
