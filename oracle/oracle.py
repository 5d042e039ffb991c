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


# ---------------------------------------------------------------- EBCOT / MQ block coder

# ---------------------------------------------------------------- narrow samples (row N3)

def orc_pack_samples(planes, dtype, interleaved=False, row_pad=0):
    """int32 planes -> one array of `dtype` (8 or 16 bits): (nc, H, W+row_pad) or (H, W+row_pad, nc)"""
    a = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    nc, (H, W) = len(a), a[0].shape
    dt = np.dtype(dtype)
    out = np.zeros((H, W + row_pad, nc) if interleaved else (nc, H, W + row_pad), dtype=dt)
    row = out.strides[0] if interleaved else out.strides[1]
    ptrs = (C.c_void_p * nc)(*[p.ctypes.data for p in a])
    oracle().orc_pack_samples(ptrs, C.c_uint32(nc), C.c_uint32(W), C.c_uint32(W), C.c_uint32(H),
                              C.c_uint32(dt.itemsize), C.c_int(int(interleaved)), _p(out), C.c_uint64(row))
    return out


def orc_unpack_samples(packed, W, interleaved=False):
    """inverse of orc_pack_samples (signed dtypes sign-extend) -> list of int32 planes"""
    packed = np.ascontiguousarray(packed)
    if interleaved:
        H, nc = packed.shape[0], packed.shape[2]
        row = packed.strides[0]
    else:
        nc, H = packed.shape[0], packed.shape[1]
        row = packed.strides[1]
    out = [np.zeros((H, W), dtype=np.int32) for _ in range(nc)]
    ptrs = (C.c_void_p * nc)(*[p.ctypes.data for p in out])
    oracle().orc_unpack_samples(_p(packed), C.c_uint64(row), C.c_uint32(packed.dtype.itemsize),
                                C.c_int(int(interleaved)), C.c_int(int(packed.dtype.kind == "i")), ptrs,
                                C.c_uint32(nc), C.c_uint32(W), C.c_uint32(W), C.c_uint32(H))
    return out


def to_smr(x):
    """two's complement int32 array -> sign-magnitude (t1_common.h:84)"""
    x = np.asarray(x, dtype=np.int64)
    return np.where(x >= 0, x, (-x) | 0x80000000).astype(np.uint32).view(np.int32)


def _t1_encode(is_ref, samples, orient=0, cblksty=0, compno=0, level=0, qmfbid=1, stepsize=1.0,
               mct_norms=None, rate_control=True):
    """samples: 2-D int32 sign-magnitude array with 6 fractional bits (T1Part1::preCompress).
    returns dict(numbps, passes=[(rate,len,term,dist)], data=bytes, distortion)"""
    s = np.ascontiguousarray(samples, dtype=np.int32)
    h, w = s.shape
    mx = int((s.view(np.uint32) & 0x7FFFFFFF).max()) if s.size else 0
    out = np.zeros(w * h * 4 + 128, dtype=np.uint8)
    norms = None if mct_norms is None else np.ascontiguousarray(mct_norms, dtype=np.float64)
    nn = 0 if norms is None else len(norms)
    args = [_p(s), C.c_uint32(w), C.c_uint32(h), C.c_uint32(mx), C.c_uint32(orient),
            C.c_uint32(compno), C.c_uint32(level), C.c_uint32(qmfbid), C.c_double(stepsize),
            C.c_uint32(cblksty), None if norms is None else _p(norms), C.c_uint32(nn),
            C.c_int(1 if rate_control else 0), _p(out)]
    if is_ref:
        numbps, npasses, nbytes = C.c_uint32(), C.c_uint32(), C.c_uint32()
        passes = (RefPass * 100)()
        d = ref().ref_t1_encode(*args, C.byref(numbps), C.byref(npasses), passes, C.byref(nbytes))
        pl = [(passes[i].rate, passes[i].len, int(passes[i].term), passes[i].distortiondec)
              for i in range(npasses.value)]
        return dict(numbps=numbps.value, passes=pl, data=out[:nbytes.value].tobytes(), distortion=d)
    res = OrcEncResult()
    d = oracle().orc_t1_encode(*args, C.byref(res))
    pl = [(res.passes[i].rate, res.passes[i].len, int(res.passes[i].term),
           res.passes[i].distortiondec) for i in range(res.numPassesTotal)]
    return dict(numbps=res.numbps, passes=pl, data=out[:res.num_bytes].tobytes(), distortion=d)


def orc_t1_encode(samples, **kw):
    return _t1_encode(False, samples, **kw)


def ref_t1_encode(samples, **kw):
    return _t1_encode(True, samples, **kw)


def segments_from_passes(passes, numbps, cblksty):
    """group coding passes into codeword segments the way T2 does (a segment ends at every
    terminated pass); returns [(len, numpasses)]"""
    segs, n, ln = [], 0, 0
    for (rate, plen, term, _d) in passes:
        n += 1
        ln += plen
        if term:
            segs.append((ln, n))
            n, ln = 0, 0
    if n:
        segs.append((ln, n))
    return segs


def _t1_decode(is_ref, data, segs, numbps, w, h, orient=0, cblksty=0):
    """returns (rc, int32 h x w array in the decoder's x2 fixed point)"""
    buf = np.frombuffer(bytes(data) + b"\0" * 8, dtype=np.uint8).copy()
    sl = np.array([s[0] for s in segs], dtype=np.uint32)
    sp = np.array([s[1] for s in segs], dtype=np.uint32)
    out = np.zeros((h, w), dtype=np.int32)
    fn = ref().ref_t1_decode if is_ref else oracle().orc_t1_decode
    rc = fn(_p(buf), _p(sl), _p(sp), C.c_uint32(len(segs)), C.c_uint32(numbps), C.c_uint32(w),
            C.c_uint32(h), C.c_uint32(orient), C.c_uint32(cblksty), _p(out))
    return rc, out


def orc_t1_decode(data, segs, numbps, w, h, **kw):
    return _t1_decode(False, data, segs, numbps, w, h, **kw)


def ref_t1_decode(data, segs, numbps, w, h, **kw):
    return _t1_decode(True, data, segs, numbps, w, h, **kw)


# ---------------------------------------------------------------- whole codec (reference C API)

def ref_codec_encode(planes, prec, irreversible, ht=False, numres=6, cblk=(64, 64), tile=(0, 0),
                     threads=0):
    """stock CPU encoder on a memory stream -> codestream bytes (uint8 array)"""
    a = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    h, w = a[0].shape
    cap = w * h * len(a) * 4 + (1 << 20)
    out = np.zeros(cap, dtype=np.uint8)
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    fn = ref().ref_codec_encode
    fn.restype = C.c_long
    n = fn(ptrs, C.c_uint32(len(a)), C.c_uint32(w), C.c_uint32(h), C.c_uint32(prec),
           C.c_int(int(irreversible)), C.c_int(int(ht)), C.c_uint32(numres), C.c_uint32(cblk[0]),
           C.c_uint32(cblk[1]), C.c_uint32(tile[0]), C.c_uint32(tile[1]), C.c_uint32(threads),
           _p(out), C.c_size_t(cap))
    if n < 0:
        raise RuntimeError("reference encoder failed")
    return out[:n].copy()


def ref_codec_decode(cs, ncomp, w, h, threads=0, want_planes=True):
    """stock CPU decoder on a memory stream -> list of int32 planes (or None)"""
    cs = np.ascontiguousarray(cs, dtype=np.uint8)
    planes = [np.zeros((h, w), dtype=np.int32) for _ in range(ncomp)] if want_planes else None
    ptrs = (C.c_void_p * ncomp)(*[p.ctypes.data for p in planes]) if want_planes else None
    rc = ref().ref_codec_decode(_p(cs), C.c_size_t(cs.size), C.c_uint32(threads), ptrs,
                                C.c_uint32(ncomp), C.c_uint32(w), C.c_uint32(h))
    if rc != 0:
        raise RuntimeError("reference decoder failed (%d)" % rc)
    return planes


# ---------------------------------------------------------------- reference host + our plugin

PLUGIN_DIR = os.path.join(os.path.dirname(HERE), "grok_b200", "plugin", "_build")


def ref_set_rate(ratio):
    """compression ratio of the single layer for the following reference encodes (0: lossless /
    all passes, no rate control)"""
    ref().ref_set_rate(C.c_double(ratio))


def have_plugin():
    return have_ref() and os.path.exists(os.path.join(PLUGIN_DIR, "libgrokj2k_plugin.so"))


def ref_plugin_open(device=0, threads=0):
    """grk_plugin_load + grk_plugin_init of the reference host on grok_b200/plugin/_build"""
    rc = ref().ref_plugin_open(PLUGIN_DIR.encode(), C.c_int(device), C.c_uint32(threads))
    if rc != 0:
        raise RuntimeError("reference host could not load/initialise the plugin (%d)" % rc)


def write_pnm(path, planes, prec):
    """binary PGM/PPM (big-endian 16 bit above 8 bits) of 1 or 3 int planes"""
    a = np.stack([np.asarray(p) for p in planes], axis=-1)
    h, w, nc = a.shape
    assert nc in (1, 3)
    with open(path, "wb") as f:
        f.write(b"P%d\n%d %d\n%d\n" % (6 if nc == 3 else 5, w, h, (1 << prec) - 1))
        f.write(a.astype(">u2" if prec > 8 else "u1").tobytes())


def ref_plugin_encode(pnm_path, w, h, ncomp, irreversible, ht=False, numres=6, cblk=(64, 64),
                      threads=0):
    """stock host (rate allocation, T2, markers) + our plugin (everything before T2)"""
    cap = w * h * ncomp * 4 + (1 << 20)
    out = np.zeros(cap, dtype=np.uint8)
    fn = ref().ref_plugin_encode
    fn.restype = C.c_long
    n = fn(pnm_path.encode(), C.c_int(int(irreversible)), C.c_int(int(ht)), C.c_uint32(numres),
           C.c_uint32(cblk[0]), C.c_uint32(cblk[1]), C.c_uint32(threads), _p(out), C.c_size_t(cap))
    if n < 0:
        raise RuntimeError("plugin-driven encode failed (%d)" % n)
    return out[:n].copy()


def ref_plugin_batch_encode(in_dir, out_dir, irreversible, ht=False, numres=6, cblk=(64, 64), threads=0):
    """grk_plugin_batch_compress over a directory of binary PNM frames: stock host callback (rate
    allocation, T2, file write) + our plugin; returns the number of frames written"""
    fn = ref().ref_plugin_batch_encode
    n = fn(in_dir.encode(), out_dir.encode(), C.c_int(int(irreversible)), C.c_int(int(ht)),
           C.c_uint32(numres), C.c_uint32(cblk[0]), C.c_uint32(cblk[1]), C.c_uint32(threads))
    if n < 0:
        raise RuntimeError("plugin-driven batch encode failed (%d)" % n)
    return n


def ref_plugin_batch_decode(in_dir, out_dir):
    """grk_plugin_init_batch_decompress + grk_plugin_batch_decompress over a directory of
    codestreams; returns the number of frames the host callback wrote (read_plane_dump)"""
    n = ref().ref_plugin_batch_decode(in_dir.encode(), out_dir.encode())
    if n < 0:
        raise RuntimeError("plugin-driven batch decode failed (%d)" % n)
    return n


def read_plane_dump(path):
    raw = np.fromfile(path, dtype=np.int32)
    nc, w, h = int(raw[0]), int(raw[1]), int(raw[2])
    return [raw[3 + c * w * h:3 + (c + 1) * w * h].reshape(h, w) for c in range(nc)]


def ref_plugin_decode(j2k_path, ncomp, w, h):
    """stock host (markers, T2) + our plugin (T1, dequantisation, wavelet, MCT)"""
    planes = [np.zeros((h, w), dtype=np.int32) for _ in range(ncomp)]
    ptrs = (C.c_void_p * ncomp)(*[p.ctypes.data for p in planes])
    rc = ref().ref_plugin_decode(j2k_path.encode(), ptrs, C.c_uint32(ncomp), C.c_uint32(w),
                                 C.c_uint32(h))
    if rc != 0:
        raise RuntimeError("plugin-driven decode failed (%d)" % rc)
    return planes


# ---------------------------------------------------------------- HTJ2K block coder

def ht_samples(values, missing_msbs):
    """int array -> sign | magnitude << (30 - missing_msbs), as T1HT::preCompress leaves them
    (t1/t1_ht/T1HT.cpp:71-107, reversible branch)"""
    v = np.asarray(values, dtype=np.int64)
    mag = np.abs(v) << (30 - missing_msbs)
    return (mag | np.where(v < 0, 0x80000000, 0)).astype(np.uint32)


def ref_ht_encode(samples, missing_msbs):
    s = np.ascontiguousarray(samples, dtype=np.uint32)
    h, w = s.shape
    out = np.zeros(16384 + 3072 + 64, dtype=np.uint8)
    n = ref().ref_ht_encode(_p(s), C.c_uint32(missing_msbs), C.c_uint32(w), C.c_uint32(h), _p(out))
    if n < 0:
        raise RuntimeError("reference HT encoder failed")
    return out[:n].tobytes()


def orc_ht_encode(samples, missing_msbs):
    s = np.ascontiguousarray(samples, dtype=np.uint32)
    h, w = s.shape
    out = np.zeros(16384 + 3072 + 64, dtype=np.uint8)
    n = oracle().orc_ht_encode(_p(s), C.c_uint32(missing_msbs), C.c_uint32(w), C.c_uint32(h), _p(out))
    if n < 0:
        raise RuntimeError("oracle HT encoder failed")
    return out[:n].tobytes()


def _ht_decode(is_ref, data, missing_msbs, w, h, num_passes=1, lengths2=0):
    raw = bytes(data)
    l1 = len(raw) - lengths2
    out = np.zeros((h, w), dtype=np.uint32)
    if is_ref:
        buf = np.frombuffer(raw + b"\0" * 8, dtype=np.uint8).copy()
        rc = ref().ref_ht_decode(_p(buf), C.c_uint32(missing_msbs), C.c_uint32(num_passes),
                                 C.c_uint32(l1), C.c_uint32(lengths2), C.c_uint32(w), C.c_uint32(h),
                                 _p(out))
    else:
        buf = np.frombuffer(b"\0" * 8 + raw + b"\0" * 8, dtype=np.uint8).copy()
        rc = oracle().orc_ht_decode(C.c_void_p(buf.ctypes.data + 8), C.c_uint32(missing_msbs),
                                    C.c_uint32(num_passes), C.c_uint32(l1), C.c_uint32(lengths2),
                                    C.c_uint32(w), C.c_uint32(h), _p(out))
    return rc, out


def orc_ht_decode(data, missing_msbs, w, h, **kw):
    return _ht_decode(False, data, missing_msbs, w, h, **kw)


def ref_ht_decode(data, missing_msbs, w, h, **kw):
    return _ht_decode(True, data, missing_msbs, w, h, **kw)


# ---------------------------------------------------------------- dequantising copy (row Q4)

def _dequant(is_ref, kind, src, roishift=0, k_msbs=0, stepsize=1.0):
    """kind 0 MQ 5/3, 1 MQ 9/7, 2 HT 5/3, 3 HT 9/7 (filters/PostDecompressFilters.h); src: int32 /
    uint32 array as the block decoder leaves it.  Returns int32 (kinds 0, 2) or float32."""
    s = np.ascontiguousarray(src).view(np.int32).copy()
    d = np.zeros_like(s)
    fn = ref().ref_dequant if is_ref else oracle().orc_dequant
    fn(C.c_int(kind), _p(s), _p(d), C.c_uint32(s.size), C.c_uint32(roishift), C.c_uint32(k_msbs),
       C.c_float(stepsize))
    return d if kind in (0, 2) else d.view(np.float32)


def orc_dequant(kind, src, **kw):
    return _dequant(False, kind, src, **kw)


def ref_dequant(kind, src, **kw):
    return _dequant(True, kind, src, **kw)


# ---------------------------------------------------------------- application-side sample loops

_APPS = None


def have_ref_apps():
    return os.path.exists(os.path.join(HERE, "_ref", "libgrok_ref_apps.so"))


def ref_apps():
    """src/bin/common/convert.cpp of the reference, compiled in place (oracle/Makefile)"""
    global _APPS
    if _APPS is None:
        ref()  # libgrok_ref.so first: it provides grk_image_new
        _APPS = C.CDLL(os.path.join(HERE, "_ref", "libgrok_ref_apps.so"))
    return _APPS


def ref_planar_to_interleaved(planes, adjust=0):
    a = [np.ascontiguousarray(p, dtype=np.int32).ravel() for p in planes]
    n = a[0].size
    out = np.zeros(n * len(a), dtype=np.int32)
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    rc = ref_apps().ref_planar_to_interleaved(ptrs, C.c_uint32(len(a)), _p(out), C.c_size_t(n),
                                              C.c_int32(adjust))
    assert rc == 0
    return out.reshape(n, len(a))


def ref_interleaved_to_planar(inter, numcomps):
    src = np.ascontiguousarray(inter, dtype=np.int32).ravel()
    n = src.size // numcomps
    out = [np.zeros(n, dtype=np.int32) for _ in range(numcomps)]
    ptrs = (C.c_void_p * numcomps)(*[p.ctypes.data for p in out])
    rc = ref_apps().ref_interleaved_to_planar(_p(src), ptrs, C.c_uint32(numcomps), C.c_size_t(n))
    assert rc == 0
    return out


def ref_narrow(samples, bytes_per_sample):
    """convert_tif_32sto16u / cvtFrom32_LUT[8]: int32 -> uint16 (host order) / uint8"""
    s = np.ascontiguousarray(samples, dtype=np.int32).ravel()
    out = np.zeros(s.size, dtype=np.uint16 if bytes_per_sample == 2 else np.uint8)
    fn = ref_apps().ref_32s_to_16u if bytes_per_sample == 2 else ref_apps().ref_32s_to_8u
    fn(_p(s), _p(out), C.c_size_t(s.size))
    return out


def ref_widen(raw, bytes_per_sample):
    """convert_16u32s_C1R (big-endian file bytes) / cvtTo32_LUT[8]: -> int32"""
    b = np.ascontiguousarray(raw, dtype=np.uint8).ravel()
    n = b.size // bytes_per_sample
    out = np.zeros(n, dtype=np.int32)
    fn = ref_apps().ref_16u_to_32s if bytes_per_sample == 2 else ref_apps().ref_8u_to_32s
    fn(_p(b), _p(out), C.c_size_t(n))
    return out


# ---------------------------------------------------------------- encoder / decoder switches

def ref_set_encode_options(opts=""):
    """further grk_compress switches for the next reference encodes ('prog=RPCL;rates=20:10:1;
    prc=128x128;sop;eph;tp=R;poc=...;cinema=4k24;roi=0:5;off=3:5' ...; see ref_harness.cpp)"""
    ref().ref_set_encode_options(opts.encode())


def ref_set_decode_options(reduce=0, layers=0):
    """cp_reduce / cp_layer of the next ref_codec_decode calls"""
    ref().ref_set_decode_options(C.c_uint32(reduce), C.c_uint32(layers))


# ---------------------------------------------------------------- array-based MCT (row M7)

def _mct_custom(is_ref, planes, matrix, forward):
    a = [np.ascontiguousarray(p).copy() for p in planes]
    m = np.ascontiguousarray(matrix, dtype=np.float32).copy()
    ptrs = (C.c_void_p * len(a))(*[p.ctypes.data for p in a])
    if is_ref:
        rc = ref().ref_mct_custom(ptrs, C.c_uint32(len(a)), C.c_uint64(a[0].size), _p(m), C.c_int(int(forward)))
        assert rc == 0
    else:
        fn = oracle().orc_mct_custom_fwd if forward else oracle().orc_mct_custom_inv
        fn(ptrs, C.c_uint32(len(a)), C.c_uint64(a[0].size), _p(m))
    return a


def orc_mct_custom(planes, matrix, forward):
    return _mct_custom(False, planes, matrix, forward)


def ref_mct_custom(planes, matrix, forward):
    return _mct_custom(True, planes, matrix, forward)


def ref_set_decode_window(x0=0, y0=0, x1=0, y1=0):
    """decode window of the next ref_codec_decode calls, relative to the image origin (0,0,0,0: all)"""
    ref().ref_set_decode_window(C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1))
