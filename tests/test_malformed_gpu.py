"""Malformed codestreams through the GPU decoder (tools/fuzz_decode.py, a short run): no crash, no hang, and
the device still decodes valid streams identically afterwards.  Runs in a process of its own with a time limit."""
import os
import subprocess
import sys

import pytest

from oracle import oracle as O

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_malformed_streams_do_not_take_the_decoder_down(gpu):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_decode.py"), "100", "112"],
                       capture_output=True, text=True, timeout=240, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-500:] + r.stderr[-1500:]
    assert "valid streams still decode identically" in r.stdout
