"""Test-side construction of coded tiles with the CPU oracle (or the compiled reference):
image -> DC shift + MCT -> forward wavelet -> quantise -> EBCOT encode per block, giving the
descriptor table, segment table and byte pool the tile decoder consumes, plus the oracle's
decode of the same tile.  Used by the GPU parity tests and by bench.py's input generator."""
import concurrent.futures as cf
import os

import numpy as np

from oracle import oracle as O
from oracle import quant as Q
from grok_b200 import tile as T


from synth import synth_image  # noqa: E402,F401  (numpy only: bench.py's reference arm imports it alone)


def forward_oracle(planes, prec, reversible, numres, mct, x0=0, y0=0):
    """DC shift + forward MCT + forward wavelet on the CPU oracle -> coefficient planes"""
    sh = 1 << (prec - 1)
    pre = [p - sh for p in planes]
    if mct:
        pre = O.orc_mct_fwd(pre, 0 if reversible else 1)
    pre = [p if reversible else p.view(np.float32) for p in pre]
    return [O.orc_dwt(p, x0, y0, numres, 1 if reversible else 0, 1) for p in pre]


def quantise_block(coef, reversible, stepsize):
    """T1Part1::preCompress (t1/t1_part1/T1Part1.cpp:37-88): sign-magnitude, 6 fractional bits"""
    if reversible:
        v = coef.astype(np.int64) * 64
    else:
        # lrintf((x / stepsize) * 64): float32 arithmetic, round to nearest even
        q = (coef.astype(np.float32) / np.float32(stepsize)) * np.float32(64)
        v = np.rint(q).astype(np.int64)
    return O.to_smr(v)


def encode_tile(coefs, prec, reversible, numres, cblk_exp=(6, 6), cblksty=0, use_ref=False,
                x0=0, y0=0, threads=None, keep_passes=None, ht=False):
    """coefs: list of coefficient planes (Mallat).  Returns dict(blocks, segs, data, geom, qcd)."""
    h, w = coefs[0].shape
    geom = T.layout(x0, y0, x0 + w, y0 + h, numres, cblk_exp[0], cblk_exp[1])
    if ht:
        return _encode_tile_ht(coefs, prec, numres, geom, use_ref, threads)
    qcd = Q.qcd_part1(numres, reversible, prec)
    enc = O.ref_t1_encode if use_ref else O.orc_t1_encode
    jobs = []
    for c, plane in enumerate(coefs):
        for g in geom:
            if g["w"] == 0 or g["h"] == 0:
                continue
            jobs.append((c, g))

    def run(job):
        c, g = job
        blk = coefs[c][g["y"]:g["y"] + g["h"], g["x"]:g["x"] + g["w"]]
        step_enc, band_numbps = Q.band_params(qcd, int(g["res"]), int(g["band"]), int(g["orient"]),
                                              prec, reversible, True)
        smr = quantise_block(blk, reversible, step_enc)
        e = enc(smr, orient=int(g["orient"]), cblksty=cblksty, level=int(g["level"]),
                qmfbid=1 if reversible else 0, stepsize=float(step_enc), rate_control=False)
        return e, band_numbps

    threads = threads or min(32, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(threads) as ex:
        results = list(ex.map(run, jobs, chunksize=16))

    blocks = np.zeros(len(jobs), dtype=T.CBLK_DTYPE)
    segs, chunks, off = [], [], 0
    for i, ((c, g), (e, band_numbps)) in enumerate(zip(jobs, results)):
        passes = e["passes"]
        if keep_passes is not None and len(passes) > keep_passes and cblksty == 0:
            passes = passes[:keep_passes]
        sg = O.segments_from_passes(passes, e["numbps"], cblksty) if e["numbps"] else []
        nbytes = sum(s[0] for s in sg)
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = g["x"], g["y"], g["w"], g["h"]
        b["comp"], b["orient"], b["numbps"], b["cblk_sty"] = c, g["orient"], e["numbps"], cblksty
        b["numsegs"], b["seg_first"] = len(sg), len(segs)
        b["data_off"], b["data_len"] = off, nbytes
        step_dec, _ = Q.band_params(qcd, int(g["res"]), int(g["band"]), int(g["orient"]), prec,
                                    reversible, False)
        b["scale"] = np.float32(step_dec) / np.float32(2)
        b["k_msbs"] = max(0, band_numbps - e["numbps"])
        segs.extend(sg)
        chunks.append(e["data"][:nbytes])
        off += nbytes
    data = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    segarr = np.array(segs, dtype=T.SEG_DTYPE) if segs else np.zeros(0, dtype=T.SEG_DTYPE)
    return dict(blocks=blocks, segs=segarr, data=data, geom=geom, qcd=qcd, shape=(h, w))


def _encode_tile_ht(coefs, prec, numres, geom, use_ref, threads):
    """HT lossless: T1HT::preCompress + ojph_encode_codeblock per block (t1/t1_ht/T1HT.cpp:56-131);
    the block reports numbps = 1, so the decoder sees k_msbs = band numbps - 1"""
    h, w = coefs[0].shape
    qcd = Q.qcd_ht_rev(numres, prec)
    enc = O.ref_ht_encode if use_ref else O.orc_ht_encode
    jobs = [(c, g) for c in range(len(coefs)) for g in geom if g["w"] and g["h"]]

    def run(job):
        c, g = job
        expn = qcd[(0 if g["res"] == 0 else 3 * int(g["res"]) - 2) + int(g["band"])][0]
        band_numbps = expn  # expn + numgbits - 1 with numgbits = 1 (grk_compress.cpp:1123-1124)
        blk = coefs[c][g["y"]:g["y"] + g["h"], g["x"]:g["x"] + g["w"]]
        return enc(O.ht_samples(blk, band_numbps), band_numbps), band_numbps

    threads = threads or min(32, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(threads) as ex:
        results = list(ex.map(run, jobs, chunksize=16))
    blocks = np.zeros(len(jobs), dtype=T.CBLK_DTYPE)
    segs, chunks, off = [], [], 0
    for i, ((c, g), (data, band_numbps)) in enumerate(zip(jobs, results)):
        b = blocks[i]
        b["x"], b["y"], b["w"], b["h"] = g["x"], g["y"], g["w"], g["h"]
        b["comp"], b["orient"], b["numbps"], b["cblk_sty"] = c, g["orient"], 1, T.CBLKSTY_HT
        b["numsegs"], b["seg_first"] = 1, len(segs)
        b["data_off"], b["data_len"] = off, len(data)
        b["k_msbs"] = band_numbps - 1
        b["scale"] = 1.0
        segs.append((len(data), 1))
        chunks.append(data)
        off += len(data)
    data = np.frombuffer(b"".join(chunks) + b"\0" * 16, dtype=np.uint8).copy()
    return dict(blocks=blocks, segs=np.array(segs, dtype=T.SEG_DTYPE), data=data, geom=geom, qcd=qcd,
                shape=(h, w))


def decode_blocks_oracle(tile, ncomp, reversible, use_ref=False, threads=None):
    """oracle T1 decode + dequantisation of every block -> coefficient planes (int32 / float32)"""
    h, w = tile["shape"]
    planes = [np.zeros((h, w), dtype=np.int32 if reversible else np.float32) for _ in range(ncomp)]
    dec = O.ref_t1_decode if use_ref else O.orc_t1_decode
    data = tile["data"]

    def run(b):
        if b["numsegs"] == 0:
            return None
        sg = [(int(s["len"]), int(s["numpasses"])) for s in
              tile["segs"][b["seg_first"]:b["seg_first"] + b["numsegs"]]]
        raw = data[b["data_off"]:b["data_off"] + b["data_len"]].tobytes()
        if b["cblk_sty"] & T.CBLKSTY_HT:
            rc, v = (O.ref_ht_decode if use_ref else O.orc_ht_decode)(raw, int(b["k_msbs"]), int(b["w"]),
                                                                        int(b["h"]))
            assert rc == 0
            # ShiftHTFilter / ScaleHTFilter and their ROI forms (PostDecompressFilters.h:91-177)
            d = O.orc_dequant(2 if reversible else 3, v, roishift=int(b["roishift"]), k_msbs=int(b["k_msbs"]),
                              stepsize=float(b["scale"]))
            return ("ht", d.reshape(v.shape))
        rc, v = dec(raw, sg, int(b["numbps"]), int(b["w"]), int(b["h"]), orient=int(b["orient"]),
                    cblksty=int(b["cblk_sty"]))
        assert rc == 0
        return v

    threads = threads or min(32, os.cpu_count() or 1)
    with cf.ThreadPoolExecutor(threads) as ex:
        vals = list(ex.map(run, tile["blocks"], chunksize=16))
    for b, v in zip(tile["blocks"], vals):
        if v is None:
            continue
        dst = planes[b["comp"]][b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]]
        if isinstance(v, tuple):
            dst[...] = v[1]
        elif int(b["roishift"]):
            # RoiShiftFilter / RoiScaleFilter (PostDecompressFilters.h:7-66); b.scale is stepsize / 2
            dst[...] = O.orc_dequant(0 if reversible else 1, v, roishift=int(b["roishift"]),
                                     stepsize=float(np.float32(b["scale"]) * np.float32(2))).reshape(v.shape)
        elif reversible:
            # ShiftFilter: C division by 2, truncation toward zero (PostDecompressFilters.h:42)
            dst[...] = (np.abs(v) >> 1) * np.sign(v)
        else:
            # ScaleFilter: (float)v * (stepsize/2) (PostDecompressFilters.h:78-85)
            dst[...] = v.astype(np.float32) * np.float32(b["scale"])
    return planes


def decode_tile_oracle(tile, ncomp, prec, reversible, numres, mct, x0=0, y0=0, use_ref=False, sgnd=0):
    coefs = decode_blocks_oracle(tile, ncomp, reversible, use_ref)
    stage = [O.orc_dwt(p, x0, y0, numres, 1 if reversible else 0, 0) for p in coefs]
    sh = 0 if sgnd else 1 << (prec - 1)
    st = [s.view(np.int32) for s in stage]
    irrev = 0 if reversible else 1
    if mct:
        out = O.orc_mct_inv(st[:3], irrev, 1, [prec] * 3, [sgnd] * 3, [sh] * 3)
        if ncomp > 3:
            out += O.orc_mct_inv(st[3:], irrev, 0, [prec] * (ncomp - 3), [sgnd] * (ncomp - 3),
                                 [sh] * (ncomp - 3))
    else:
        out = O.orc_mct_inv(st, irrev, 0, [prec] * ncomp, [sgnd] * ncomp, [sh] * ncomp)
    return out, coefs


def decode_codestream_oracle(cs, reduce=0, max_layers=0):
    """Whole codestream through the product's own T2 (grok_b200.codestream: host C++, no GPU) and
    the ORACLE's block decoders, wavelet and component transform, tile by tile -> int32 planes of
    the (reduced) image.  What comes out equals the stock decoder's planes only if every table T2
    produced (bytes, segments, passes, bit planes, step sizes, geometry) is right."""
    from grok_b200 import codestream as CS
    c = CS.Codestream(cs)
    i = c.info
    W, Hh = c.image_size(reduce)
    ix0, iy0 = (i.x0 + (1 << reduce) - 1) >> reduce, (i.y0 + (1 << reduce) - 1) >> reduce
    out = [np.zeros((Hh, W), dtype=np.int32) for _ in range(i.numcomps)]
    info = []
    for t in range(c.num_tiles):
        tab = c.parse_tile(t, reduce, max_layers)
        p = tab["params"]
        info.append(tab)
        if p.x1 <= p.x0 or p.y1 <= p.y0:
            continue
        planes, _ = decode_tile_oracle(tab, i.numcomps, int(i.prec[0]), bool(p.reversible), p.numres, p.mct,
                                       x0=p.x0, y0=p.y0, sgnd=int(i.sgnd[0]))
        for k in range(i.numcomps):
            out[k][p.y0 - iy0:p.y1 - iy0, p.x0 - ix0:p.x1 - ix0] = planes[k]
    c.close()
    return out, info
