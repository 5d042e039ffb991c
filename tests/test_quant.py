"""host-side quantisation parameters of the library (grkb_qcd_generate / grkb_band_quant, no
device work) against the oracle's restatement, which is pinned against the QCD markers the stock
encoder writes (tests/test_e2e_oracle.py)."""
import numpy as np
import pytest

from grok_b200 import quant as Q
from oracle import quant as OQ


@pytest.mark.parametrize("reversible", [0, 1])
def test_qcd_part1(reversible):
    for numres in range(1, 12):
        for prec in (1, 8, 12, 16):
            assert Q.qcd_generate(numres, prec, reversible) == OQ.qcd_part1(numres, bool(reversible), prec)


def test_qcd_ht_reversible():
    for numres in range(1, 12):
        for prec in (8, 12, 16):
            assert Q.qcd_generate(numres, prec, 1, ht=True) == OQ.qcd_ht_rev(numres, prec)
    with pytest.raises(Exception):
        Q.qcd_generate(6, 12, 0, ht=True)


def test_band_quant():
    for reversible in (0, 1):
        for prec in (8, 12, 16):
            qcd = OQ.qcd_part1(6, bool(reversible), prec)
            for compress in (0, 1):
                for resno in range(6):
                    for bi, orient in ([(0, 0)] if resno == 0 else [(0, 1), (1, 2), (2, 3)]):
                        want = OQ.band_params(qcd, resno, bi, orient, prec, bool(reversible),
                                              bool(compress), numgbits=2, roishift=0)
                        got = Q.band_quant(qcd, resno, bi, orient, prec, reversible, compress)
                        assert np.float32(got[0]) == want[0] and got[1] == want[1]
