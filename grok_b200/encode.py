"""numpy-facing wrapper of the whole-tile encoder (include/grk_b200.h: grkb_encoder_*) and the
host-side construction of its block table -- what T1CompressScheduler::scheduleCompress
(src/lib/jp2/t1/T1CompressScheduler.cpp:31-154) fills into CompressBlockExec.  No codec
arithmetic here: geometry comes from grkb_tile_layout, step sizes from grkb_band_quant."""
import ctypes as C

import numpy as np

from ._lib import lib, check, TileParams
from . import quant as Q
from . import tile as T

lib.grkb_encoder_create.argtypes = [C.POINTER(TileParams), C.POINTER(C.c_void_p)]
lib.grkb_encoder_destroy.argtypes = [C.c_void_p]
lib.grkb_encoder_destroy.restype = None
lib.grkb_encoder_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                    C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
lib.grkb_encoder_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                    C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_size_t]
lib.grkb_encoder_wait.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
lib.grkb_encoder_time_device.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_float)]

# weights of the colour components in the distortion estimate (mct::get_norms_rev / _irrev,
# src/lib/jp2/point_transform/mct.cpp: L2 norms of the inverse RCT / ICT basis vectors)
MCT_NORMS_REV = (1.732, 0.8292, 0.8292)
MCT_NORMS_IRREV = (1.732, 1.805, 1.573)


def enc_block_table(x0, y0, x1, y1, numcomps, numres, prec, reversible, ht=False, cblk_exp=(6, 6),
                    cblk_sty=0, numgbits=None, qcd=None):
    """(ENC_CBLK_DTYPE array, geometry array, qcd): every non-empty code block of every component,
    component-major in the reference's walking order."""
    if numgbits is None:
        numgbits = 1 if ht else 2  # grk_compress.cpp:1123-1124 sets one guard bit for HT
    if qcd is None:
        qcd = Q.qcd_generate(numres, prec, reversible, ht)
    geom = T.layout(x0, y0, x1, y1, numres, cblk_exp[0], cblk_exp[1])
    geom = geom[(geom["w"] > 0) & (geom["h"] > 0)]
    n = len(geom)
    blocks = np.zeros(n * numcomps, dtype=T.ENC_CBLK_DTYPE)
    band_cache = {}
    for c in range(numcomps):
        b = blocks[c * n:(c + 1) * n]
        for k in ("x", "y", "w", "h", "orient", "level"):
            b[k] = geom[k]
        b["comp"] = c
        b["cblk_sty"] = (T.CBLKSTY_HT if ht else 0) | cblk_sty
        for i, g in enumerate(geom):
            key = (int(g["res"]), int(g["band"]), int(g["orient"]))
            if key not in band_cache:
                band_cache[key] = Q.band_quant(qcd, key[0], key[1], key[2], prec, reversible, True,
                                               ht=ht, numgbits=numgbits)
            b[i]["stepsize"], b[i]["k_msbs"] = band_cache[key]
    return blocks, geom, qcd


class TileEncoder:
    """grkb_encoder: device-resident state for one tile geometry"""

    def __init__(self, params):
        self.params = params
        self.h = C.c_void_p()
        check(lib.grkb_encoder_create(C.byref(params), C.byref(self.h)), "grkb_encoder_create")
        self.W, self.H = params.x1 - params.x0, params.y1 - params.y0

    def close(self):
        if self.h:
            lib.grkb_encoder_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode(self, image, blocks, mct_norms=None, rate_control=False, cap=None, out=None):
        """image: list of int32 (H, stride) planes; blocks: ENC_CBLK_DTYPE array.
        Returns (coded, passes[nblocks, 100], bytes) like tile.t1_encode."""
        blocks = np.ascontiguousarray(blocks, dtype=T.ENC_CBLK_DTYPE)
        a = [np.ascontiguousarray(p, dtype=np.int32) for p in image]
        stride = a[0].strides[0] // 4
        n = len(blocks)
        if out is None:
            coded = np.zeros(n, dtype=T.CODED_DTYPE)
            passes = np.zeros((n, T.ENC_MAX_PASSES), dtype=T.PASS_DTYPE)
            if cap is None:
                cap = self.W * self.H * len(a) * 4 + 4096
            buf = np.zeros(max(cap, 1), dtype=np.uint8)
        else:
            coded, passes, buf = out
            cap = buf.size
        ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
        norms = None if mct_norms is None else np.ascontiguousarray(mct_norms, dtype=np.float64)
        used = C.c_size_t()
        check(lib.grkb_encoder_encode(self.h, ptrs, stride, blocks.ctypes.data, n,
                                      None if norms is None else norms.ctypes.data,
                                      0 if norms is None else len(norms), int(rate_control),
                                      coded.ctypes.data, passes.ctypes.data, buf.ctypes.data, cap,
                                      C.byref(used)), "grkb_encoder_encode")
        return coded, passes, buf[:used.value]

    def submit(self, image, blocks, out, mct_norms=None, rate_control=False):
        """grkb_encoder_submit: enqueue one tile without waiting.  image planes, blocks and the
        (coded, passes, bytes) arrays of `out` should be page-locked (tile.PinnedArray) and are
        kept referenced here until wait()."""
        blocks = np.ascontiguousarray(blocks, dtype=T.ENC_CBLK_DTYPE)
        a = [np.ascontiguousarray(p, dtype=np.int32) for p in image]
        stride = a[0].strides[0] // 4
        coded, passes, buf = out
        ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
        norms = None if mct_norms is None else np.ascontiguousarray(mct_norms, dtype=np.float64)
        self._inflight = (a, blocks, norms, ptrs, out)
        check(lib.grkb_encoder_submit(self.h, ptrs, stride, blocks.ctypes.data, len(blocks),
                                      None if norms is None else norms.ctypes.data,
                                      0 if norms is None else len(norms), int(rate_control),
                                      coded.ctypes.data, passes.ctypes.data, buf.ctypes.data, buf.size),
              "grkb_encoder_submit")

    def submit_packed(self, image, blocks, out, mct_norms=None, rate_control=False, interleaved=False):
        """grkb_encoder_submit_packed (row N3): the image as one 8- or 16-bit array, widened on the
        device (signed dtypes sign-extend): (numcomps, H, >=W) planar, or (H, >=W, numcomps) with
        interleaved=True."""
        blocks = np.ascontiguousarray(blocks, dtype=T.ENC_CBLK_DTYPE)
        fmt, row_bytes = T.sample_format(image, interleaved, self.params.numcomps,
                                         self.params.y1 - self.params.y0)
        coded, passes, buf = out
        norms = None if mct_norms is None else np.ascontiguousarray(mct_norms, dtype=np.float64)
        self._inflight = (image, blocks, norms, None, out)
        check(lib.grkb_encoder_submit_packed(self.h, image.ctypes.data, fmt, int(image.dtype.kind == "i"),
                                             row_bytes, blocks.ctypes.data, len(blocks),
                                             None if norms is None else norms.ctypes.data,
                                             0 if norms is None else len(norms), int(rate_control),
                                             coded.ctypes.data, passes.ctypes.data, buf.ctypes.data, buf.size),
              "grkb_encoder_submit_packed")

    def wait(self):
        """grkb_encoder_wait -> (coded, passes, bytes) of the tile submitted last"""
        used = C.c_size_t()
        check(lib.grkb_encoder_wait(self.h, C.byref(used)), "grkb_encoder_wait")
        coded, passes, buf = self._inflight[4]
        self._inflight = None
        return coded, passes, buf[:used.value]

    def time_device(self, iters, warmup):
        ms = (C.c_float * 3)()
        check(lib.grkb_encoder_time_device(self.h, iters, warmup, ms), "grkb_encoder_time_device")
        return dict(total_ms=ms[0], dwt_ms=ms[1], t1_ms=ms[2])


def decode_table(blocks, coded, passes, geom_per_comp, qcd, prec, reversible, ht, numgbits=None):
    """Decoder descriptor table (CBLK_DTYPE, SEG_DTYPE) for the streams an encode produced --
    what T2 would hand the decoder after a lossless single-layer round trip."""
    if numgbits is None:
        numgbits = 1 if ht else 2
    n = len(blocks)
    out = np.zeros(n, dtype=T.CBLK_DTYPE)
    segs = []
    ng = len(geom_per_comp)
    for i in range(n):
        e, c, g = blocks[i], coded[i], geom_per_comp[i % ng]
        b = out[i]
        for k in ("x", "y", "w", "h", "comp", "orient"):
            b[k] = e[k]
        b["cblk_sty"] = e["cblk_sty"]
        b["numbps"] = c["numbps"]
        step_dec, band_numbps = Q.band_quant(qcd, int(g["res"]), int(g["band"]), int(g["orient"]),
                                             prec, reversible, False, ht=ht, numgbits=numgbits)
        b["scale"] = step_dec if ht else np.float32(step_dec) / np.float32(2)
        b["k_msbs"] = max(0, band_numbps - int(c["numbps"]))
        b["data_off"], b["data_len"] = int(c["data_off"]), int(c["data_len"])
        b["seg_first"] = len(segs)
        if ht:
            if c["data_len"]:
                segs.append((int(c["data_len"]), 1))
        else:
            ln = cnt = 0
            for p in passes[i, :int(c["numpasses"])]:
                ln += int(p["len"])
                cnt += 1
                if p["term"]:
                    segs.append((ln, cnt))
                    ln = cnt = 0
            if cnt:
                segs.append((ln, cnt))
        b["numsegs"] = len(segs) - int(b["seg_first"])
    sa = np.array(segs, dtype=T.SEG_DTYPE) if segs else np.zeros(0, dtype=T.SEG_DTYPE)
    return out, sa
