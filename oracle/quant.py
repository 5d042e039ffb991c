"""CPU restatement of the reference's quantisation parameters (TEST INFRASTRUCTURE ONLY).

Restates  param_qcd::generate (Part-1 branch)   src/lib/jp2/codestream/HTParams.cpp:194-252
          Quantizer::setBandStepSizeAndBps        src/lib/jp2/codestream/Quantizer.cpp:26-66
          T1::getnorm tables                      src/lib/jp2/t1/t1_part1/T1.cpp:239-277
"""
import math

NORMS_97 = [
    [1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9],
    [2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0],
    [2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0],
    [2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2]]


def norm97(level, orient):
    level = min(level, 9 if orient == 0 else 8)
    return NORMS_97[orient][level]


def floorlog2(v):
    return v.bit_length() - 1


def qcd_part1(numres, reversible, prec):
    """[(expn, mant)] per band in codestream order (LL, then HL LH HH per resolution)"""
    out = []
    for bandno in range(3 * numres - 2):
        resno = 0 if bandno == 0 else (bandno - 1) // 3 + 1
        orient = 0 if bandno == 0 else (bandno - 1) % 3 + 1
        level = numres - 1 - resno
        gain = 0 if not reversible else (0 if orient == 0 else (1 if orient in (1, 2) else 2))
        stepsize = 1.0 if reversible else (1 << gain) / norm97(level, orient)
        step = int(math.floor(stepsize * 8192.0))
        p = floorlog2(step) - 13
        n = 11 - floorlog2(step)
        mant = ((step >> -n) if n < 0 else (step << n)) & 0x7FF
        expn = prec + gain - p
        out.append((expn, 0 if reversible else mant))
    return out


def band_params(qcd, resno, bandindex, orient, prec, reversible, compress, numgbits=2, roishift=0):
    """(stepsize as float32 value, numbps) of one band"""
    import numpy as np
    log2_gain = 0 if (not compress and not reversible) else (0 if orient == 0 else 2 if orient == 3 else 1)
    numbps = prec + log2_gain
    expn, mant = qcd[(0 if resno == 0 else 3 * resno - 2) + bandindex]
    stepsize = np.float32((1.0 + mant / 2048.0) * math.pow(2.0, numbps - expn))
    band_numbps = roishift + max(0, expn + numgbits - 1)
    return stepsize, band_numbps


# BIBO gains of the 5/3 analysis filter bank per decomposition depth (OpenJPH tables used by the
# reference's HT QCD generator, codestream/HTParams.cpp:134-146), first entries
BIBO_53_L = [1.0000, 1.5000, 1.6250, 1.6875, 1.6963, 1.7067, 1.7116, 1.7129, 1.7141, 1.7145, 1.7151]
BIBO_53_H = [2.0000, 2.5000, 2.7500, 2.8047, 2.8198, 2.8410, 2.8558, 2.8601, 2.8628, 2.8656, 2.8662]


def qcd_ht_rev(numres, prec):
    """param_qcd::set_rev_quant (HTParams.cpp:253-272); the colour-transform bit is never added
    because tcp->mct is still 0 when the reference calls it (CodeStreamCompress.cpp:382-396)"""
    import numpy as np
    nd = numres - 1
    f32 = np.float32

    def X(a, b):
        return int(math.ceil(math.log(float(f32(a) * f32(b) * f32(1.1))) / math.log(2.0)))
    out = [(prec + X(BIBO_53_L[nd], BIBO_53_L[nd]), 0)]
    for d in range(nd - 1, -1, -1):
        lo, hi = BIBO_53_L[d + 1], BIBO_53_H[d]
        out += [(prec + X(hi, lo), 0)] * 2 + [(prec + X(hi, hi), 0)]
    return out


def parse_qcd(cs):
    """(Sqcd, [SPqcd words]) from the first QCD marker segment of a codestream (bytes-like)"""
    b = bytes(cs)
    i = b.find(b"\xff\x5c")
    ln = (b[i + 2] << 8) | b[i + 3]
    sq = b[i + 4]
    body = b[i + 5:i + 2 + ln]
    if (sq & 0x1F) == 0:
        return sq, [(v >> 3, 0) for v in body]
    return sq, [(((body[k] << 8) | body[k + 1]) >> 11, ((body[k] << 8) | body[k + 1]) & 0x7FF)
                for k in range(0, len(body), 2)]
