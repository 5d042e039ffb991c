"""CPU: pins the C restatement (oracle/liboracle.so) against the compiled, unmodified reference
(oracle/_ref/libgrok_ref.so, built by oracle/Makefile from /root/reference).  Skipped when the
compiled reference is absent."""
import os
import re

import numpy as np
import pytest

from common import GEOMS, plane, same_bits, ref_defect
from oracle import oracle as O

pytestmark = pytest.mark.skipif(not O.have_ref(), reason="compiled reference not present")

REF_T1 = "/root/reference/src/lib/jp2/t1/t1_part1"


@pytest.mark.parametrize("reversible", [1, 0])
@pytest.mark.parametrize("forward", [0, 1])
def test_dwt(orc, reversible, forward):
    rng = np.random.default_rng(5)
    for (w, h, x0, y0, nr) in GEOMS:
        if ref_defect(w, h, x0, y0, nr, reversible, forward):
            continue
        p = plane(rng, w, h, reversible)
        assert same_bits(O.ref_dwt(p, x0, y0, nr, reversible, forward),
                         O.orc_dwt(p, x0, y0, nr, reversible, forward)), (w, h, x0, y0, nr)


@pytest.mark.parametrize("irrev", [0, 1])
def test_mct(orc, irrev):
    rng = np.random.default_rng(6)
    for shape in [(1, 1), (5, 3), (64, 64), (97, 211)]:
        pl = [rng.integers(-32768, 32768, size=shape).astype(np.int32) for _ in range(3)]
        for a, b in zip(O.ref_mct_fwd(pl, irrev), O.orc_mct_fwd(pl, irrev)):
            assert same_bits(a, b)
        if irrev:
            inp = [(rng.standard_normal(shape) * 20000).astype(np.float32).view(np.int32)
                   for _ in range(3)]
        else:
            inp = [rng.integers(-40000, 40000, size=shape).astype(np.int32) for _ in range(3)]
        for mct_on in (1, 0):
            for prec, sgnd, sh in [(16, 0, 32768), (12, 1, 0), (8, 0, 128)]:
                r = O.ref_mct_inv(inp, irrev, mct_on, [prec] * 3, [sgnd] * 3, [sh] * 3)
                o = O.orc_mct_inv(inp, irrev, mct_on, [prec] * 3, [sgnd] * 3, [sh] * 3)
                assert all(same_bits(a, b) for a, b in zip(r, o))


def _block(rng, w, h, bits, density):
    """sign-magnitude samples with 6 fractional bits, sparse like wavelet coefficients"""
    mag = rng.integers(0, 1 << bits, size=(h, w))
    mag = np.where(rng.random((h, w)) < density, mag, mag >> max(bits - 3, 0))
    sgn = rng.integers(0, 2, size=(h, w))
    return ((mag.astype(np.int64) << 6) | (sgn.astype(np.int64) << 31)).astype(np.uint32).view(np.int32)


STYLES = [0, 0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x01 | 0x04, 0x01 | 0x10, 0x02 | 0x08 | 0x20, 0x3F]
SHAPES = [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9)]


def test_t1_encode_matches_reference(orc):
    rng = np.random.default_rng(8)
    n = 0
    for sty in STYLES:
        for (w, h) in SHAPES:
            for bits in (1, 4, 11):
                blk = _block(rng, w, h, bits, 0.3)
                orient = int(rng.integers(0, 4))
                kw = dict(orient=orient, cblksty=sty, compno=1, level=2, qmfbid=int(rng.integers(0, 2)),
                          stepsize=0.37, mct_norms=[1.732, 1.805, 1.573])
                r = O.ref_t1_encode(blk, **kw)
                o = O.orc_t1_encode(blk, **kw)
                assert r["numbps"] == o["numbps"]
                assert r["data"] == o["data"], (sty, w, h, bits)
                assert [p[:3] for p in r["passes"]] == [p[:3] for p in o["passes"]]
                assert [p[3] for p in r["passes"]] == [p[3] for p in o["passes"]]
                assert r["distortion"] == o["distortion"]
                n += 1
    assert n > 300


def test_t1_decode_matches_reference(orc):
    rng = np.random.default_rng(9)
    for sty in STYLES:
        for (w, h) in SHAPES:
            for bits in (2, 9):
                blk = _block(rng, w, h, bits, 0.25)
                orient = int(rng.integers(0, 4))
                enc = O.ref_t1_encode(blk, orient=orient, cblksty=sty)
                if enc["numbps"] == 0:
                    continue
                segs = O.segments_from_passes(enc["passes"], enc["numbps"], sty)
                rc_r, r = O.ref_t1_decode(enc["data"], segs, enc["numbps"], w, h, orient=orient, cblksty=sty)
                rc_o, o = O.orc_t1_decode(enc["data"], segs, enc["numbps"], w, h, orient=orient, cblksty=sty)
                assert rc_r == rc_o == 0
                assert np.array_equal(r, o), (sty, w, h, bits)
                # lossless round trip: the x2 fixed point halves back to the coded magnitudes
                mag = (blk.view(np.uint32) & 0x7FFFFFFF) >> 6
                sgn = blk.view(np.uint32) >> 31
                full = np.where(sgn == 1, -mag.astype(np.int64), mag.astype(np.int64))
                assert np.array_equal(np.trunc(o / 2).astype(np.int64), full)
                # truncated stream (fewer passes): still identical to the reference
                if len(enc["passes"]) > 3 and sty == 0:
                    keep = len(enc["passes"]) // 2
                    ln = enc["passes"][keep - 1][0]
                    _, r2 = O.ref_t1_decode(enc["data"][:ln], [(ln, keep)], enc["numbps"], w, h, orient=orient)
                    _, o2 = O.orc_t1_decode(enc["data"][:ln], [(ln, keep)], enc["numbps"], w, h, orient=orient)
                    assert np.array_equal(r2, o2)


def test_t1_decode_garbage_matches_reference(orc):
    """arbitrary bytes are a valid MQ stream: both decoders must make the same decisions"""
    rng = np.random.default_rng(10)
    for (w, h) in [(64, 64), (7, 9), (32, 32)]:
        for numbps in (1, 5, 12):
            data = rng.integers(0, 256, size=int(rng.integers(1, 700)), dtype=np.uint8).tobytes()
            npass = 3 * numbps - 2
            for sty in (0, 0x08, 0x02):
                _, r = O.ref_t1_decode(data, [(len(data), npass)], numbps, w, h, orient=3, cblksty=sty)
                _, o = O.orc_t1_decode(data, [(len(data), npass)], numbps, w, h, orient=3, cblksty=sty)
                assert np.array_equal(r, o)


@pytest.mark.skipif(not os.path.isdir(REF_T1), reason="reference sources not present")
def test_tables_match_reference_sources(orc):
    """the oracle derives its tables from ISO 15444-1 rules; compare with the reference's literals"""
    import ctypes as C
    lib = O.oracle()
    txt = open(os.path.join(REF_T1, "t1_luts.h")).read()

    def arr(name):
        m = re.search(name + r"\[[^\]]*\]\s*=\s*\{([^}]*)\}", txt)
        return [int(x, 0) for x in m.group(1).replace("\n", " ").split(",") if x.strip()]
    for name, fn, n, ty in [("lut_ctxno_zc", "orc_lut_ctxno_zc", 2048, C.c_uint8),
                            ("lut_ctxno_sc", "orc_lut_ctxno_sc", 256, C.c_uint8),
                            ("lut_spb", "orc_lut_spb", 256, C.c_uint8)]:
        f = getattr(lib, fn)
        f.restype = C.POINTER(ty)
        got = [f()[i] for i in range(n)]
        assert got == arr(name), name
    lib.orc_lut_nmsedec.restype = C.POINTER(C.c_int16)
    for which, name in enumerate(["lut_nmsedec_sig", "lut_nmsedec_sig0", "lut_nmsedec_ref", "lut_nmsedec_ref0"]):
        p = lib.orc_lut_nmsedec(which)
        assert [p[i] for i in range(128)] == arr(name), name
    # MQ state table: mqc_enc.cpp:34-82 lists 94 states {qe, mps, &states[nmps], &states[nlps]}
    src = open(os.path.join(REF_T1, "mqc_enc.cpp")).read()
    rows = re.findall(r"\{(0x[0-9a-fA-F]+),\s*([01]),\s*&mqc_states\[(\d+)\],\s*&mqc_states\[(\d+)\]\}", src)
    assert len(rows) == 94
    lib.orc_mq_table.restype = C.POINTER(C.c_uint16)
    t = lib.orc_mq_table()
    for i in range(47):
        qe, nmps, nlps, sw = t[4 * i], t[4 * i + 1], t[4 * i + 2], t[4 * i + 3]
        for mps in (0, 1):
            r = rows[2 * i + mps]
            assert int(r[0], 16) == qe and int(r[1]) == mps
            assert int(r[2]) == 2 * nmps + mps
            assert int(r[3]) == 2 * nlps + (mps ^ sw)


def test_ht_decode_matches_reference(orc):
    rng = np.random.default_rng(31)
    n = 0
    for (w, h) in [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9),
                   (2, 2), (7, 2), (128, 32), (6, 6), (1024, 4), (4, 1024)]:
        for bits in (1, 2, 5, 9, 12):
            for dens in (0.02, 0.3, 0.9):
                for mm in (bits + 1, 14, 20):
                    if 30 - mm < bits:
                        continue
                    vals = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w))
                    vals = np.where(rng.random((h, w)) < dens, vals, 0)
                    enc = O.ref_ht_encode(O.ht_samples(vals, mm), mm)
                    for mmd in (mm, mm - 1):
                        rr, r = O.ref_ht_decode(enc, mmd, w, h)
                        ro, o = O.orc_ht_decode(enc, mmd, w, h)
                        assert rr == ro and (rr != 0 or np.array_equal(r, o)), (w, h, bits, dens, mm, mmd)
                        n += 1
                    # a too-large missing_msbs makes U_q exceed its bound: both reject
                    rr, _ = O.ref_ht_decode(enc, 29, w, h)
                    ro, _ = O.orc_ht_decode(enc, 29, w, h)
                    assert rr == ro
    assert n > 1000


def test_ht_refinement_passes_match_reference(orc):
    """SigProp / MagRef: not reachable through Grok's host, so the pin is the
    reference's ojph_decode_codeblock called directly with a second segment of arbitrary bytes"""
    from common import ht_refine_case
    rng = np.random.default_rng(5)
    shapes = [(64, 64), (32, 32), (8, 8), (4, 4), (5, 7), (7, 5), (1, 1), (2, 1), (1, 2), (3, 3), (16, 9),
              (9, 16), (13, 10), (33, 6), (6, 33), (64, 1), (1, 64), (64, 2), (64, 3), (64, 5), (64, 6),
              (64, 7), (12, 11), (17, 13), (100, 4), (4, 100), (40, 21), (1024, 4), (4, 1024)]
    changed = 0
    for it in range(580):
        w, h = shapes[it % len(shapes)]
        _, kmsbs, data, l2 = ht_refine_case(rng, w, h, O.orc_ht_encode)
        for passes in (2, 3):
            rr, r = O.ref_ht_decode(data, kmsbs, w, h, num_passes=passes, lengths2=l2)
            ro, o = O.orc_ht_decode(data, kmsbs, w, h, num_passes=passes, lengths2=l2)
            assert rr == ro == 0 and np.array_equal(r, o), (it, w, h, kmsbs, passes, l2)
        _, base = O.orc_ht_decode(data[:len(data) - l2], kmsbs, w, h)
        changed += int(not np.array_equal(base, o))
    assert changed > 400  # the passes do something
    # a second pass without a second segment is rejected by both (ojph_block_decoder.cpp:1014-1020)
    rr, _ = O.ref_ht_decode(data[:len(data) - l2], kmsbs, w, h, num_passes=2, lengths2=0)
    ro, _ = O.orc_ht_decode(data[:len(data) - l2], kmsbs, w, h, num_passes=2, lengths2=0)
    assert rr != 0 and ro != 0


def test_ht_encode_matches_reference(orc):
    rng = np.random.default_rng(32)
    n = 0
    for (w, h) in [(64, 64), (32, 32), (1, 1), (3, 5), (17, 4), (64, 3), (5, 64), (13, 13), (4, 1), (1, 9),
                   (2, 2), (7, 2), (128, 32), (6, 6), (1024, 4), (4, 1024), (2, 1), (1, 2)]:
        for bits in (1, 2, 5, 9, 12):
            for dens in (0.0, 0.02, 0.3, 0.9, 1.0):
                for mm in (bits + 1, 14, 20):
                    if 30 - mm < bits:
                        continue
                    vals = rng.integers(-(1 << bits) + 1, 1 << bits, size=(h, w))
                    vals = np.where(rng.random((h, w)) < dens, vals, 0)
                    smp = O.ht_samples(vals, mm)
                    assert O.ref_ht_encode(smp, mm) == O.orc_ht_encode(smp, mm), (w, h, bits, dens, mm)
                    n += 1
    assert n > 1000


# ---------------------------------------------------------------- dequantising copy (row Q4, ROI)

@pytest.mark.parametrize("kind", [0, 1, 2, 3])
@pytest.mark.parametrize("roishift", [0, 1, 5, 9, 14])
def test_dequant_filters(orc, kind, roishift):
    """oracle/dequant.c == the reference's own filter classes (PostDecompressFilters.h:7-177),
    ROI forms included; magnitudes on both sides of 2^roishift, both signs, zero"""
    rng = np.random.default_rng(900 + kind * 31 + roishift)
    n = 4096
    if kind < 2:
        bits = rng.integers(1, 26, size=n)
        v = (rng.integers(0, 1 << 26, size=n) >> (26 - bits)).astype(np.int64)
        v = np.where(rng.random(n) < 0.5, -v, v).astype(np.int32)
        edge = np.array([0, 1, -1, (1 << roishift) - 1, 1 << roishift, -(1 << roishift),
                         -(1 << roishift) + 1, (1 << roishift) + 1], dtype=np.int32)
        v[:edge.size] = edge
        k_msbs = 0
    else:
        k_msbs = int(rng.integers(0, 28))
        mag = rng.integers(0, 1 << 31, size=n, dtype=np.int64)
        mag >>= rng.integers(0, 31, size=n)
        sign = np.where(rng.random(n) < 0.5, 0x80000000, 0)
        v = (mag | sign).astype(np.uint32).view(np.int32)
    for step in (1.0, 0.0009765625, 3.1415927):
        r = O.ref_dequant(kind, v, roishift=roishift, k_msbs=k_msbs, stepsize=step)
        o = O.orc_dequant(kind, v, roishift=roishift, k_msbs=k_msbs, stepsize=step)
        assert same_bits(r, o), (kind, roishift, step)


# ---------------------------------------------------------------- narrow samples (row N3)

needs_apps = pytest.mark.skipif(not O.have_ref_apps(), reason="reference convert.cpp not compiled")


@needs_apps
@pytest.mark.parametrize("numcomps", [1, 3, 4])
@pytest.mark.parametrize("bytes_per_sample", [2, 1])
def test_samples_pack_vs_reference_convert(orc, numcomps, bytes_per_sample):
    """oracle/samples.c == planarToInterleaved<N> (convert.cpp:397-405) followed by
    convert_tif_32sto16u (:1158-1163) / cvtFrom32_LUT[8], including values beyond the narrow range
    (the reference's loops are C casts: they wrap)"""
    rng = np.random.default_rng(41 + numcomps)
    H, W = 19, 37
    dt = np.uint16 if bytes_per_sample == 2 else np.uint8
    lim = 1 << (8 * bytes_per_sample)
    planes = [rng.integers(-lim, 2 * lim, size=(H, W)).astype(np.int32) for _ in range(numcomps)]
    inter = O.ref_planar_to_interleaved(planes)                       # (H*W, numcomps) int32
    want_inter = O.ref_narrow(inter, bytes_per_sample).reshape(H, W, numcomps)
    got_inter = O.orc_pack_samples(planes, dt, interleaved=True)
    assert np.array_equal(got_inter, want_inter)
    want_planar = np.stack([O.ref_narrow(p, bytes_per_sample).reshape(H, W) for p in planes])
    assert np.array_equal(O.orc_pack_samples(planes, dt, interleaved=False), want_planar)


@needs_apps
@pytest.mark.parametrize("numcomps", [1, 3])
@pytest.mark.parametrize("bytes_per_sample", [2, 1])
def test_samples_unpack_vs_reference_convert(orc, numcomps, bytes_per_sample):
    """oracle/samples.c == convert_16u32s_C1R (convert.cpp:547-556, big-endian file bytes) /
    cvtTo32_LUT[8] followed by interleavedToPlanar<N> (:360-384)"""
    rng = np.random.default_rng(43 + numcomps)
    H, W = 11, 29
    dt = np.uint16 if bytes_per_sample == 2 else np.uint8
    packed = rng.integers(0, 1 << (8 * bytes_per_sample), size=(H, W, numcomps)).astype(dt)
    file_bytes = packed.astype(">u2").tobytes() if bytes_per_sample == 2 else packed.tobytes()
    wide = O.ref_widen(np.frombuffer(file_bytes, dtype=np.uint8), bytes_per_sample)
    want = [p.reshape(H, W) for p in O.ref_interleaved_to_planar(wide, numcomps)]
    got = O.orc_unpack_samples(packed, W, interleaved=True)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)


# ---------------------------------------------------------------- array-based MCT (row M7)

@pytest.mark.parametrize("nc", [1, 2, 3, 4])
def test_custom_mct(orc, nc):
    """oracle/mct.c == mct::compress_custom / decompress_custom (point_transform/mct.cpp:622-686)"""
    rng = np.random.default_rng(70 + nc)
    m = (rng.standard_normal((nc, nc)) * 0.7).astype(np.float32)
    ints = [rng.integers(-40000, 40000, size=777).astype(np.int32) for _ in range(nc)]
    for a, b in zip(O.ref_mct_custom(ints, m, 1), O.orc_mct_custom(ints, m, 1)):
        assert np.array_equal(a, b)
    flts = [(rng.standard_normal(777) * 3000).astype(np.float32) for _ in range(nc)]
    for a, b in zip(O.ref_mct_custom(flts, m, 0), O.orc_mct_custom(flts, m, 0)):
        assert same_bits(a, b)
