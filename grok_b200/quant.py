"""ctypes mirror of the host-side quantisation parameters (include/grk_b200.h):
param_qcd::generate and Quantizer::setBandStepSizeAndBps of the reference."""
import ctypes as C

from ._lib import lib, check

lib.grkb_qcd_generate.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
lib.grkb_band_quant.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int,
                                C.c_int, C.c_uint32, C.c_uint32, C.POINTER(C.c_float),
                                C.POINTER(C.c_uint32)]


def qcd_generate(numres, prec, reversible, ht=False):
    """[(expn, mant)] per band in codestream order"""
    n = 3 * numres - 2
    e, m = (C.c_uint32 * n)(), (C.c_uint32 * n)()
    check(lib.grkb_qcd_generate(numres, prec, int(reversible), int(ht), e, m), "grkb_qcd_generate")
    return [(e[i], m[i]) for i in range(n)]


def band_quant(qcd, resno, bandindex, orient, prec, reversible, compress, ht=False, numgbits=2,
               roishift=0):
    """(stepsize, numbps) of one band"""
    expn, mant = qcd[(0 if resno == 0 else 3 * resno - 2) + bandindex]
    s, nb = C.c_float(), C.c_uint32()
    check(lib.grkb_band_quant(expn, mant, orient, prec, int(reversible), int(compress), int(ht),
                              numgbits, roishift, C.byref(s), C.byref(nb)), "grkb_band_quant")
    return s.value, nb.value
