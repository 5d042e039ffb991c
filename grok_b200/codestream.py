"""numpy-facing wrappers of the codestream-level C-ABI (include/grk_b200.h, section E): the
codestream index, per-tile T2 (descriptor tables from packet headers) and the whole-image
decompressor.  No codec arithmetic here."""
import ctypes as C

import numpy as np

from ._lib import lib, check, TileParams, GrokB200Error
from . import tile as T

_u32, _vp = C.c_uint32, C.c_void_p


class CodestreamInfo(C.Structure):
    """grkb_codestream_info"""
    _fields_ = [("x0", _u32), ("y0", _u32), ("x1", _u32), ("y1", _u32), ("tile_x0", _u32),
                ("tile_y0", _u32), ("tile_w", _u32), ("tile_h", _u32), ("tiles_x", _u32),
                ("tiles_y", _u32), ("numcomps", _u32), ("prec", C.c_uint8 * 4), ("sgnd", C.c_uint8 * 4),
                ("numres", _u32), ("numlayers", _u32), ("prog_order", _u32), ("reversible", _u32),
                ("mct", _u32), ("cblkw_exp", _u32), ("cblkh_exp", _u32), ("cblk_sty", _u32),
                ("csty", _u32), ("ht", _u32), ("truncated", _u32), ("rsiz", _u32), ("tile_parts", _u32),
                ("features", _u32)]


class TileTables(C.Structure):
    """grkb_tile_tables"""
    _fields_ = [("params", TileParams), ("blocks", _vp), ("nblocks", _u32), ("segs", _vp),
                ("nsegs", _u32), ("bytes", _vp), ("nbytes", C.c_size_t), ("zero_copy", _u32),
                ("packets", _u32), ("truncated", _u32)]


lib.grkb_codestream_open.argtypes = [_vp, C.c_size_t, C.POINTER(_vp)]
lib.grkb_codestream_close.argtypes = [_vp]
lib.grkb_codestream_close.restype = None
lib.grkb_codestream_get_info.argtypes = [_vp, C.POINTER(CodestreamInfo)]
lib.grkb_codestream_parse_tile.argtypes = [_vp, _u32, _u32, _u32, C.POINTER(TileTables)]
lib.grkb_decompressor_create.argtypes = [C.c_int, _u32, C.POINTER(_vp)]
lib.grkb_decompressor_destroy.argtypes = [_vp]
lib.grkb_decompressor_destroy.restype = None
lib.grkb_decompress_submit.argtypes = [_vp, _vp, C.c_size_t, _u32, _u32, _vp, _u32]
lib.grkb_decompress_submit_window.argtypes = [_vp, _vp, C.c_size_t, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _u32]
lib.grkb_decompress_submit_tiles.argtypes = [_vp, _vp, C.c_size_t, _u32, _u32, _u32, _u32, _vp, _u32]
lib.grkb_decompress_submit_packed.argtypes = [_vp, _vp, C.c_size_t, _u32, _u32, _vp, _u32, C.c_size_t]
lib.grkb_decompress_wait.argtypes = [_vp, C.POINTER(_u32)]
lib.grkb_decompress.argtypes = [_vp, _vp, C.c_size_t, _u32, _u32, _vp, _u32, C.POINTER(_u32)]
lib.grkb_decompress_stats.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(_u32)]


def _ceildivpow2(a, b):
    return (a + (1 << b) - 1) >> b


class Codestream:
    """grkb_codestream over a uint8 array (kept alive here)"""

    def __init__(self, data):
        self.data = np.ascontiguousarray(data, dtype=np.uint8)
        self.h = _vp()
        check(lib.grkb_codestream_open(self.data.ctypes.data, self.data.size, C.byref(self.h)),
              "grkb_codestream_open")
        self.info = CodestreamInfo()
        check(lib.grkb_codestream_get_info(self.h, C.byref(self.info)), "grkb_codestream_get_info")

    def close(self):
        if self.h:
            lib.grkb_codestream_close(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_tiles(self):
        return self.info.tiles_x * self.info.tiles_y

    def image_size(self, reduce=0):
        """(width, height) of the image the decompressor delivers, sized as the reference sizes its
        grk_image: ceil(width / 2^reduce) (codestream/markers/SIZMarker.cpp:54-57)"""
        i = self.info
        return _ceildivpow2(i.x1 - i.x0, reduce), _ceildivpow2(i.y1 - i.y0, reduce)

    def parse_tile(self, tileno, reduce=0, max_layers=0):
        """T2 of one tile -> dict(params, blocks, segs, data, zero_copy, packets, truncated); the
        arrays are copies (the C tables live until the next parse of that tile)"""
        t = TileTables()
        check(lib.grkb_codestream_parse_tile(self.h, tileno, reduce, max_layers, C.byref(t)),
              "grkb_codestream_parse_tile")

        def arr(ptr, n, dt):
            if not n:
                return np.zeros(0, dtype=dt)
            buf = (C.c_uint8 * (n * dt.itemsize)).from_address(ptr)
            return np.frombuffer(buf, dtype=dt, count=n).copy()

        p = TileParams()
        C.memmove(C.byref(p), C.byref(t.params), C.sizeof(TileParams))
        data = arr(t.bytes, t.nbytes, np.dtype(np.uint8))
        return dict(params=p, blocks=arr(t.blocks, t.nblocks, T.CBLK_DTYPE), segs=arr(t.segs, t.nsegs, T.SEG_DTYPE),
                    data=np.concatenate([data, np.zeros(16, np.uint8)]), zero_copy=bool(t.zero_copy),
                    packets=t.packets, truncated=bool(t.truncated),
                    shape=(p.y1 - p.y0, p.x1 - p.x0))


class Decompressor:
    """grkb_decompressor: whole codestreams -> host planes on one device"""

    def __init__(self, device=0, slots=2):
        self.h = _vp()
        check(lib.grkb_decompressor_create(device, slots, C.byref(self.h)), "grkb_decompressor_create")

    def close(self):
        if self.h:
            lib.grkb_decompressor_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def submit(self, data, out, reduce=0, max_layers=0):
        """data: uint8 array (whole codestream or JP2 file); out: list of int32 (H, stride) planes"""
        nc = len(out)
        ptrs = (_vp * nc)(*[o.ctypes.data for o in out])
        self._keep = (data, out, ptrs)
        check(lib.grkb_decompress_submit(self.h, data.ctypes.data, data.size, reduce, max_layers, ptrs,
                                         out[0].strides[0] // 4), "grkb_decompress_submit")

    def submit_tiles(self, data, out, first_tile, tile_step, reduce=0, max_layers=0):
        """tile sharding: only tiles first_tile, first_tile + tile_step, ... (rank, world of a multi-GPU decode)"""
        nc = len(out)
        ptrs = (_vp * nc)(*[o.ctypes.data for o in out])
        self._keep = (data, out, ptrs)
        check(lib.grkb_decompress_submit_tiles(self.h, data.ctypes.data, data.size, reduce, max_layers, first_tile,
                                               tile_step, ptrs, out[0].strides[0] // 4), "grkb_decompress_submit_tiles")

    def decode_window(self, data, window, reduce=0, max_layers=0):
        """windowed decode: window = (x0, y0, x1, y1) relative to the image origin; returns int32 planes
        of the window at the (reduced) resolution"""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        cs = Codestream(data)
        i = cs.info
        nc = i.numcomps
        x0, y0 = min(i.x0 + window[0], i.x1), min(i.y0 + window[1], i.y1)
        x1, y1 = min(i.x0 + window[2], i.x1), min(i.y0 + window[3], i.y1)
        cs.close()
        w = _ceildivpow2(x1, reduce) - _ceildivpow2(x0, reduce)
        h = _ceildivpow2(y1, reduce) - _ceildivpow2(y0, reduce)
        out = [np.zeros((h, w), dtype=np.int32) for _ in range(nc)]
        ptrs = (_vp * nc)(*[o.ctypes.data for o in out])
        self._keep = (data, out, ptrs)
        check(lib.grkb_decompress_submit_window(self.h, data.ctypes.data, data.size, reduce, max_layers,
                                                window[0], window[1], window[2], window[3], ptrs, w),
              "grkb_decompress_submit_window")
        return self.wait()

    def submit_packed(self, data, out, numcomps, interleaved=False, reduce=0, max_layers=0):
        H = out.shape[0] if interleaved else out.shape[1]
        fmt, row_bytes = T.sample_format(out, interleaved, numcomps, H)
        self._keep = (data, out, None)
        check(lib.grkb_decompress_submit_packed(self.h, data.ctypes.data, data.size, reduce, max_layers,
                                                out.ctypes.data, fmt, row_bytes), "grkb_decompress_submit_packed")

    def wait(self):
        rej = _u32()
        check(lib.grkb_decompress_wait(self.h, C.byref(rej)), "grkb_decompress_wait")
        self.rejected = rej.value
        keep, self._keep = self._keep, None
        return keep[1]

    def decode(self, data, reduce=0, max_layers=0):
        """synchronous: returns a list of int32 planes of the (reduced) image"""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        cs = Codestream(data)
        w, h = cs.image_size(reduce)
        nc = cs.info.numcomps
        cs.close()
        out = [np.zeros((h, w), dtype=np.int32) for _ in range(nc)]
        self.submit(data, out, reduce, max_layers)
        return self.wait()

    def stats(self):
        ms, n = C.c_double(), _u32()
        check(lib.grkb_decompress_stats(self.h, C.byref(ms), C.byref(n)), "grkb_decompress_stats")
        return dict(t2_ms=ms.value, tiles=n.value)


# ---------------------------------------------------------------------------------- encode side

class CompressParams(C.Structure):
    """grkb_compress_params"""
    _fields_ = [("width", _u32), ("height", _u32), ("numcomps", _u32), ("prec", _u32), ("sgnd", _u32),
                ("x0", _u32), ("y0", _u32), ("tile_w", _u32), ("tile_h", _u32), ("tile_x0", _u32),
                ("tile_y0", _u32), ("numres", _u32), ("cblkw_exp", _u32), ("cblkh_exp", _u32),
                ("cblk_sty", _u32), ("reversible", _u32), ("mct", _u32), ("prog_order", _u32), ("csty", _u32),
                ("numgbits", _u32), ("prec_w_exp", C.c_uint8 * 33), ("prec_h_exp", C.c_uint8 * 33),
                ("comment", C.c_char * 64), ("rate", C.c_double), ("numlayers", _u32),
                ("layer_rates", C.c_double * 100)]


lib.grkb_compressor_create.argtypes = [C.c_int, C.POINTER(CompressParams), C.POINTER(_vp)]
lib.grkb_compressor_destroy.argtypes = [_vp]
lib.grkb_compressor_destroy.restype = None
lib.grkb_compress.argtypes = [_vp, _vp, _u32, _vp, C.c_size_t, C.POINTER(C.c_size_t)]
lib.grkb_compress_stats.argtypes = [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
lib.grkb_compress_allocation.argtypes = [_vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]

PROGS = {"LRCP": 0, "RLCP": 1, "RPCL": 2, "PCRL": 3, "CPRL": 4}


class Compressor:
    """grkb_compressor: images -> codestreams on one device (every pass kept, or one / several rate-controlled layers)"""

    def __init__(self, width, height, numcomps, prec, reversible=True, ht=False, numres=6, cblk=(64, 64),
                 tile=(0, 0), origin=(0, 0), tile_origin=None, prog="LRCP", precincts=None, sop=False, eph=False,
                 cblk_sty=0, mct=None, sgnd=False, comment="", rate=0.0, rates=None, device=0):
        """rate: compression ratio of a single layer (0: every pass).  rates: one ratio per quality layer
        (decreasing, as -r a,b,c of grk_compress; 0: every remaining pass)"""
        p = CompressParams()
        p.width, p.height, p.numcomps, p.prec, p.sgnd = width, height, numcomps, prec, int(sgnd)
        p.x0, p.y0 = origin
        p.tile_w, p.tile_h = tile
        p.tile_x0, p.tile_y0 = tile_origin if tile_origin is not None else (0, 0)
        p.numres = numres
        p.cblkw_exp, p.cblkh_exp = cblk[0].bit_length() - 1, cblk[1].bit_length() - 1
        p.cblk_sty = T.CBLKSTY_HT if ht else cblk_sty
        p.reversible = int(reversible)
        p.mct = int(numcomps >= 3 if mct is None else mct)
        p.prog_order = PROGS[prog]
        p.csty = (2 if sop else 0) | (4 if eph else 0)
        for r in range(33):
            p.prec_w_exp[r] = p.prec_h_exp[r] = 15
        if precincts:
            # as grk_compress -c: sizes from the HIGHEST resolution down, the last one halved for
            # every resolution below (CodeStreamCompress.cpp:542-596)
            for r in range(numres):
                k = numres - 1 - r
                if k < len(precincts):
                    w, h = precincts[k]
                else:
                    w, h = precincts[-1][0] >> (k - len(precincts) + 1), precincts[-1][1] >> (k - len(precincts) + 1)
                p.prec_w_exp[r] = max(w, 1).bit_length() - 1
                p.prec_h_exp[r] = max(h, 1).bit_length() - 1
        p.comment = comment.encode()
        p.rate = float(rate)
        if rates:
            p.numlayers = len(rates)
            for i, r in enumerate(rates):
                p.layer_rates[i] = float(r)
            if len(rates) == 1:
                p.rate = float(rates[0])
        self.params = p
        self.h = _vp()
        check(lib.grkb_compressor_create(device, C.byref(p), C.byref(self.h)), "grkb_compressor_create")

    def close(self):
        if self.h:
            lib.grkb_compressor_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode(self, planes):
        a = [np.ascontiguousarray(x, dtype=np.int32) for x in planes]
        ptrs = (_vp * len(a))(*[x.ctypes.data for x in a])
        cap = a[0].size * len(a) * 4 + (1 << 20)
        out = np.zeros(cap, dtype=np.uint8)
        used = C.c_size_t()
        check(lib.grkb_compress(self.h, ptrs, a[0].strides[0] // 4, out.ctypes.data, cap, C.byref(used)),
              "grkb_compress")
        return out[:used.value].copy()

    def stats(self):
        a, b = C.c_double(), C.c_double()
        check(lib.grkb_compress_stats(self.h, C.byref(a), C.byref(b)), "grkb_compress_stats")
        d, h, n = C.c_uint32(), C.c_uint32(), C.c_uint32()
        check(lib.grkb_compress_allocation(self.h, C.byref(d), C.byref(h), C.byref(n)), "grkb_compress_allocation")
        return dict(device_ms=a.value, t2_ms=b.value, tiles_allocated_on_device=d.value,
                    tiles_allocated_on_host=h.value, bisection_steps=n.value)
