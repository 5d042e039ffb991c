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
