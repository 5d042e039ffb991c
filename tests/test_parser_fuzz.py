"""Memory safety of the host-side marker / packet parser (grok_b200/csrc/codestream.cpp, t2_parse.cpp): the
sources are built with AddressSanitizer + UBSan into a small driver (tools/fuzz_parser/driver.cpp) that opens
every file given, parses every tile (plain, reduced, layer-limited + windowed) and checks that the tables it
would hand to the device stay inside the bytes it was given.  Fed with the stock encoder's codestreams cut
off, spliced and with bytes flipped."""
import os
import subprocess

import numpy as np
import pytest

import codec_helpers as H
import test_random_gpu as R
from oracle import oracle as O

pytestmark = pytest.mark.skipif(not O.have_ref(), reason="compiled reference (oracle/_ref) not present")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("fuzz") / "driver")
    s = os.path.join(ROOT, "grok_b200", "csrc")
    cmd = ["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-I" + s,
           "-I/usr/local/cuda/include", os.path.join(ROOT, "tools", "fuzz_parser", "driver.cpp")] + \
          [os.path.join(s, f) for f in ("codestream.cpp", "t2_parse.cpp", "geometry.cpp", "quant.cpp")] + ["-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        pytest.skip("no sanitizer build here: " + r.stderr[-300:])
    return out


def test_parser_survives_malformed_codestreams(driver, tmp_path):
    rng = np.random.default_rng(77)
    files = []
    for seed in (100, 103, 105, 109, 119, 124, 135, 1001):
        c = R.draw(seed)
        img = H.synth_image(c["w"], c["h"], c["ncomp"], c["prec"], seed=seed)
        try:
            cs = R.stock_encode(img, c, R.stock_opts(c, "rates=40:10" if not c["ht"] else ""))
        except BaseException:  # (the stock encoder refuses the combination)
            continue
        if cs.size < 200:
            continue
        for trial in range(30):
            b = cs.copy()
            kind = trial % 5
            if kind == 0:
                b = b[:int(rng.integers(10, b.size))]
            elif kind == 1:
                for _ in range(int(rng.integers(1, 6))):
                    b[int(rng.integers(2, min(b.size, 400)))] = int(rng.integers(0, 256))
            elif kind == 2:
                for _ in range(int(rng.integers(1, 12))):
                    b[int(rng.integers(2, b.size))] = int(rng.integers(0, 256))
            elif kind == 3:
                a, d = sorted(rng.integers(2, b.size, size=2).tolist())
                b = np.concatenate([b[:a], b[d:]])
            else:
                b[int(rng.integers(2, min(b.size, 300)))] = int(rng.choice([0, 0xFF, 0x7F, 0x80]))
            f = str(tmp_path / ("%d_%d.j2k" % (seed, trial)))
            b.tofile(f)
            files.append(f)
        f = str(tmp_path / ("%d.j2k" % seed))
        cs.tofile(f)
        files.append(f)
    assert len(files) > 150
    r = subprocess.run([driver] + files, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "Sanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-2000:]
    assert r.stdout.startswith("parsed ")
