"""numpy-facing wrappers of the tile-level C-ABI (include/grk_b200.h, section B): code-block
layout, descriptor tables and the whole-tile decoder.  No codec arithmetic here."""
import ctypes as C

import numpy as np

from ._lib import lib, check, Cblk, Seg, TileParams, LayoutParams, CblkGeom, GrokB200Error

CBLKSTY_HT = 0x40

CBLK_DTYPE = np.dtype([("x", "<u4"), ("y", "<u4"), ("w", "<u2"), ("h", "<u2"), ("comp", "u1"),
                       ("orient", "u1"), ("numbps", "u1"), ("cblk_sty", "u1"), ("roishift", "u1"),
                       ("k_msbs", "u1"), ("numsegs", "<u2"), ("seg_first", "<u4"),
                       ("data_off", "<u4"), ("data_len", "<u4"), ("scale", "<f4")])
assert CBLK_DTYPE.itemsize == C.sizeof(Cblk)
SEG_DTYPE = np.dtype([("len", "<u4"), ("numpasses", "<u4")])
GEOM_DTYPE = np.dtype([("x", "<u4"), ("y", "<u4"), ("cx0", "<u4"), ("cy0", "<u4"), ("w", "<u2"),
                       ("h", "<u2"), ("res", "u1"), ("band", "u1"), ("orient", "u1"),
                       ("level", "u1"), ("precinct", "<u4"), ("cblkno", "<u4")])
assert GEOM_DTYPE.itemsize == C.sizeof(CblkGeom)


def layout(x0, y0, x1, y1, numres, cblkw_exp=6, cblkh_exp=6, prec_w_exp=None, prec_h_exp=None):
    """code blocks of one tile-component in the reference's walking order (structured array)"""
    p = LayoutParams()
    p.x0, p.y0, p.x1, p.y1, p.numres = x0, y0, x1, y1, numres
    p.cblkw_exp, p.cblkh_exp = cblkw_exp, cblkh_exp
    for r in range(33):
        p.prec_w_exp[r] = 15 if prec_w_exp is None else (prec_w_exp[r] if r < len(prec_w_exp) else 15)
        p.prec_h_exp[r] = 15 if prec_h_exp is None else (prec_h_exp[r] if r < len(prec_h_exp) else 15)
    n = C.c_uint32()
    check(lib.grkb_tile_layout(C.byref(p), None, 0, C.byref(n)), "grkb_tile_layout")
    out = np.zeros(n.value, dtype=GEOM_DTYPE)
    if n.value:
        check(lib.grkb_tile_layout(C.byref(p), out.ctypes.data_as(C.POINTER(CblkGeom)), n.value,
                                   C.byref(n)), "grkb_tile_layout")
    return out


def clamp_range(prec, sgnd):
    if sgnd:
        return -(1 << (prec - 1)), (1 << (prec - 1)) - 1
    return 0, (1 << prec) - 1


def tile_params(x0, y0, x1, y1, numcomps, numres, reversible, mct, prec, sgnd=0):
    p = TileParams()
    p.x0, p.y0, p.x1, p.y1 = x0, y0, x1, y1
    p.numcomps, p.numres, p.reversible, p.mct = numcomps, numres, int(reversible), int(mct)
    lo, hi = clamp_range(prec, sgnd)
    for c in range(numcomps):
        p.shift[c] = 0 if sgnd else 1 << (prec - 1)
        p.lo[c], p.hi[c] = lo, hi
    return p


class PinnedArray:
    """numpy view of page-locked host memory from grkb_host_alloc"""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = lib.grkb_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise GrokB200Error("grkb_host_alloc failed: " + lib.grkb_last_error().decode())
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib.grkb_host_free(self.ptr)
            self.ptr = None


SAMPLES_32, SAMPLES_16, SAMPLES_16_INTERLEAVED, SAMPLES_8, SAMPLES_8_INTERLEAVED = range(5)
SAMPLES_BIG_ENDIAN = 0x10


def sample_format(a, interleaved, numcomps, H):
    """(format, row_bytes) of a packed sample array for the grkb_*_submit_packed calls."""
    if a.ndim != 3 or a.dtype.itemsize not in (1, 2) or a.dtype.kind not in "iu":
        raise ValueError("packed samples: 3-D array of 8- or 16-bit integers expected")
    if interleaved:
        ok = a.shape[0] == H and a.shape[2] == numcomps and a.strides[2] == a.dtype.itemsize and \
            a.strides[1] == numcomps * a.dtype.itemsize
        row = a.strides[0]
    else:
        ok = a.shape[0] == numcomps and a.shape[1] == H and a.strides[2] == a.dtype.itemsize and \
            a.strides[0] == a.strides[1] * H
        row = a.strides[1]
    if not ok:
        raise ValueError("packed samples: unexpected shape/strides %r %r" % (a.shape, a.strides))
    wide = a.dtype.itemsize == 2
    fmt = (SAMPLES_16_INTERLEAVED if interleaved else SAMPLES_16) if wide else \
        (SAMPLES_8_INTERLEAVED if interleaved else SAMPLES_8)
    if wide and a.dtype.byteorder == ">":
        fmt |= SAMPLES_BIG_ENDIAN  # as PNM files store them
    return fmt, row


class TileDecoder:
    """grkb_decoder: device-resident state for one tile geometry"""

    def __init__(self, params, max_blocks=0, max_bytes=0):
        self.params = params
        self.h = C.c_void_p()
        check(lib.grkb_decoder_create(C.byref(params), max_blocks, max_bytes, C.byref(self.h)),
              "grkb_decoder_create")
        self.W, self.H = params.x1 - params.x0, params.y1 - params.y0

    def close(self):
        if self.h:
            lib.grkb_decoder_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def decode(self, blocks, segs, data, out=None, download=True):
        """blocks: CBLK_DTYPE array, segs: SEG_DTYPE array, data: uint8 array (byte pool).
        out: optional list of int32 (H, stride) arrays to receive the planes."""
        blocks = np.ascontiguousarray(blocks, dtype=CBLK_DTYPE)
        segs = np.ascontiguousarray(segs, dtype=SEG_DTYPE)
        nc = self.params.numcomps
        ptrs = None
        stride = 0
        if download:
            if out is None:
                out = [np.empty((self.H, self.W), dtype=np.int32) for _ in range(nc)]
            stride = out[0].strides[0] // 4
            ptrs = (C.c_void_p * nc)(*[o.ctypes.data for o in out])
        st = C.c_uint32()
        check(lib.grkb_decoder_decode(self.h, blocks.ctypes.data, len(blocks), segs.ctypes.data,
                                      len(segs), data.ctypes.data, data.size, ptrs, stride,
                                      C.byref(st)), "grkb_decoder_decode")
        self.rejected = st.value
        return out

    def submit(self, blocks, segs, data, out):
        """asynchronous half of decode(): enqueue and return; the arrays must stay alive and
        untouched until wait().  One submit at a time per decoder."""
        blocks = np.ascontiguousarray(blocks, dtype=CBLK_DTYPE)
        segs = np.ascontiguousarray(segs, dtype=SEG_DTYPE)
        nc = self.params.numcomps
        stride = out[0].strides[0] // 4
        ptrs = (C.c_void_p * nc)(*[o.ctypes.data for o in out])
        self._inflight = (blocks, segs, data, out, ptrs)
        check(lib.grkb_decoder_submit(self.h, blocks.ctypes.data, len(blocks), segs.ctypes.data,
                                      len(segs), data.ctypes.data, data.size, ptrs, stride),
              "grkb_decoder_submit")

    def submit_packed(self, blocks, segs, data, out, interleaved=False):
        """grkb_decoder_submit_packed (row N3): like submit(), the result narrowed on the device to
        out's dtype (uint16 / int16 / uint8 / int8).  out: (numcomps, H, >=W) planar, or
        (H, >=W, numcomps) with interleaved=True; rows contiguous."""
        blocks = np.ascontiguousarray(blocks, dtype=CBLK_DTYPE)
        segs = np.ascontiguousarray(segs, dtype=SEG_DTYPE)
        fmt, row_bytes = sample_format(out, interleaved, self.params.numcomps, self.H)
        self._inflight = (blocks, segs, data, out, None)
        check(lib.grkb_decoder_submit_packed(self.h, blocks.ctypes.data, len(blocks), segs.ctypes.data,
                                             len(segs), data.ctypes.data, data.size, out.ctypes.data,
                                             fmt, row_bytes), "grkb_decoder_submit_packed")

    def wait(self):
        st = C.c_uint32()
        check(lib.grkb_decoder_wait(self.h, C.byref(st)), "grkb_decoder_wait")
        self.rejected = st.value
        out = self._inflight[3] if getattr(self, "_inflight", None) else None
        self._inflight = None
        return out

    def time_device(self, iters, warmup):
        ms = (C.c_float * 4)()
        n = C.c_uint32()
        check(lib.grkb_decoder_time_device(self.h, iters, warmup, ms, C.byref(n)),
              "grkb_decoder_time_device")
        return dict(total_ms=ms[0], t1_ms=ms[1], dwt_ms=ms[2], dwt_last_ms=ms[3], launches=n.value)

    def download(self):
        nc = self.params.numcomps
        out = [np.empty((self.H, self.W), dtype=np.int32) for _ in range(nc)]
        ptrs = (C.c_void_p * nc)(*[o.ctypes.data for o in out])
        check(lib.grkb_decoder_download(self.h, ptrs, self.W), "grkb_decoder_download")
        return out


def t1_decode(blocks, segs, data, planes, reversible):
    """block decoder alone: writes the dequantised blocks into copies of `planes` (list of 2-D
    int32/float32 arrays of identical shape) and returns them with the reject count"""
    blocks = np.ascontiguousarray(blocks, dtype=CBLK_DTYPE)
    segs = np.ascontiguousarray(segs, dtype=SEG_DTYPE)
    a = [np.ascontiguousarray(p).copy() for p in planes]
    h, w = a[0].shape
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    st = C.c_uint32()
    check(lib.grkb_t1_decode_host(blocks.ctypes.data, len(blocks), segs.ctypes.data, len(segs),
                                  data.ctypes.data, data.size, ptrs, len(a), w, h, int(reversible),
                                  C.byref(st)), "grkb_t1_decode_host")
    return a, st.value


# ---------------------------------------------------------------------------------- encode side
ENC_CBLK_DTYPE = np.dtype([("x", "<u4"), ("y", "<u4"), ("w", "<u2"), ("h", "<u2"), ("comp", "u1"),
                           ("orient", "u1"), ("level", "u1"), ("cblk_sty", "u1"), ("k_msbs", "u1"),
                           ("reserved", "u1", (3,)), ("stepsize", "<f4"),
                           ("reserved2", "<u4", (2,))])
assert ENC_CBLK_DTYPE.itemsize == 32
PASS_DTYPE = np.dtype([("distortion", "<f8"), ("rate", "<u4"), ("len", "<u2"), ("term", "u1"),
                       ("reserved", "u1")])
assert PASS_DTYPE.itemsize == 16
CODED_DTYPE = np.dtype([("data_off", "<u8"), ("data_len", "<u4"), ("numbps", "u1"),
                        ("numpasses", "u1"), ("failed", "<u2"), ("distortion", "<f8")])
assert CODED_DTYPE.itemsize == 24
ENC_MAX_PASSES = 100


def t1_encode(blocks, planes, reversible, mct_norms=None, rate_control=False, cap=None):
    """block encoders alone (grkb_t1_encode_host): blocks is an ENC_CBLK_DTYPE array, planes a list
    of 2-D int32 (reversible) / float32 coefficient arrays of identical shape.
    Returns (coded CODED_DTYPE array, passes PASS_DTYPE array [nblocks, 100], bytes uint8 array)."""
    blocks = np.ascontiguousarray(blocks, dtype=ENC_CBLK_DTYPE)
    a = [np.ascontiguousarray(p) for p in planes]
    h, stride = a[0].shape
    for p in a:
        if p.shape != (h, stride) or p.dtype.itemsize != 4:
            raise ValueError("planes must be 32-bit and of identical shape")
    n = len(blocks)
    coded = np.zeros(n, dtype=CODED_DTYPE)
    passes = np.zeros((n, ENC_MAX_PASSES), dtype=PASS_DTYPE)
    if cap is None:
        cap = int(sum(int(b["w"]) * int(b["h"]) for b in blocks)) * 4 + 20000 * max(n, 1)
    out = np.zeros(max(cap, 1), dtype=np.uint8)
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    norms = None if mct_norms is None else np.ascontiguousarray(mct_norms, dtype=np.float64)
    used = C.c_size_t()
    check(lib.grkb_t1_encode_host(blocks.ctypes.data, n, ptrs, len(a), stride, h, int(reversible),
                                  None if norms is None else norms.ctypes.data,
                                  0 if norms is None else len(norms), int(rate_control),
                                  coded.ctypes.data, passes.ctypes.data, out.ctypes.data, cap,
                                  C.byref(used)), "grkb_t1_encode_host")
    return coded, passes, out[:used.value]
