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


# ---- section B: tile entry points ------------------------------------------------------------

class Cblk(C.Structure):
    """grkb_cblk"""
    _fields_ = [("x", _u32), ("y", _u32), ("w", C.c_uint16), ("h", C.c_uint16),
                ("comp", C.c_uint8), ("orient", C.c_uint8), ("numbps", C.c_uint8),
                ("cblk_sty", C.c_uint8), ("roishift", C.c_uint8), ("k_msbs", C.c_uint8),
                ("numsegs", C.c_uint16), ("seg_first", _u32), ("data_off", _u32),
                ("data_len", _u32), ("scale", C.c_float)]


class Seg(C.Structure):
    """grkb_seg"""
    _fields_ = [("len", _u32), ("numpasses", _u32)]


class TileParams(C.Structure):
    """grkb_tile_params"""
    _fields_ = [("x0", _u32), ("y0", _u32), ("x1", _u32), ("y1", _u32), ("numcomps", _u32),
                ("numres", _u32), ("reversible", _i32), ("mct", _i32), ("shift", _i32 * 4),
                ("lo", _i32 * 4), ("hi", _i32 * 4)]


class LayoutParams(C.Structure):
    """grkb_layout_params"""
    _fields_ = [("x0", _u32), ("y0", _u32), ("x1", _u32), ("y1", _u32), ("numres", _u32),
                ("cblkw_exp", _u32), ("cblkh_exp", _u32), ("prec_w_exp", _u32 * 33),
                ("prec_h_exp", _u32 * 33)]


class CblkGeom(C.Structure):
    """grkb_cblk_geom"""
    _fields_ = [("x", _u32), ("y", _u32), ("cx0", _u32), ("cy0", _u32), ("w", C.c_uint16),
                ("h", C.c_uint16), ("res", C.c_uint8), ("band", C.c_uint8), ("orient", C.c_uint8),
                ("level", C.c_uint8), ("precinct", _u32), ("cblkno", _u32)]


lib.grkb_tile_layout.argtypes = [C.POINTER(LayoutParams), C.POINTER(CblkGeom), _u32,
                                 C.POINTER(_u32)]
lib.grkb_decoder_create.argtypes = [C.POINTER(TileParams), _u32, C.c_size_t, C.POINTER(_vp)]
lib.grkb_decoder_destroy.argtypes = [_vp]
lib.grkb_decoder_destroy.restype = None
lib.grkb_decoder_decode.argtypes = [_vp, _vp, _u32, _vp, _u32, _vp, C.c_size_t, _vp, _u32,
                                    C.POINTER(_u32)]
lib.grkb_decoder_submit.argtypes = [_vp, _vp, _u32, _vp, _u32, _vp, C.c_size_t, _vp, _u32]
lib.grkb_decoder_wait.argtypes = [_vp, C.POINTER(_u32)]
lib.grkb_decoder_submit_packed.argtypes = [_vp, _vp, _u32, _vp, _u32, _vp, C.c_size_t, _vp, _u32, C.c_size_t]
lib.grkb_samples_size.argtypes = [_u32, _u32, _u32, C.c_size_t]
lib.grkb_samples_size.restype = C.c_size_t
lib.grkb_encoder_submit_packed.argtypes = [_vp, _vp, _u32, _int, C.c_size_t, _vp, _u32, _vp, _u32, _int, _vp, _vp,
                                           _vp, C.c_size_t]
lib.grkb_decoder_time_device.argtypes = [_vp, _u32, _u32, C.POINTER(C.c_float), C.POINTER(_u32)]
lib.grkb_decoder_download.argtypes = [_vp, _vp, _u32]
lib.grkb_t1_decode_host.argtypes = [_vp, _u32, _vp, _u32, _vp, C.c_size_t, _vp, _u32, _u32, _u32,
                                    _int, C.POINTER(_u32)]
lib.grkb_t1_encode_host.argtypes = [_vp, _u32, _vp, _u32, _u32, _u32, _int, _vp, _u32, _int, _vp, _vp,
                                    _vp, C.c_size_t, C.POINTER(C.c_size_t)]
lib.grkb_host_alloc.argtypes = [C.c_size_t]
lib.grkb_host_alloc.restype = _vp
lib.grkb_host_free.argtypes = [_vp]
lib.grkb_host_free.restype = None
lib.grkb_host_register.argtypes = [_vp, C.c_size_t]
lib.grkb_host_unregister.argtypes = [_vp]
lib.grkb_bench_link.argtypes = [C.c_size_t, _u32, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                C.POINTER(C.c_double)]
lib.grkb_mct_custom_host.argtypes = [_vp, _u32, _u64, _vp, _int]
# section D: the allocation-free door (device pointers in and out, built-in scratch)
lib.grkb_idwt_device_noalloc.argtypes = [_vp, _u32, _vp, _u32, _u32, _u32, _u32, _u32, _u32, _int, _vp]
NOALLOC_SCRATCH_BYTES = 8 << 20
