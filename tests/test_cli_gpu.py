"""The stock reference COMMAND-LINE tools driving the plugin (north_star: "drops in as the accelerator
behind grk_decompress / grk_compress").  oracle/_ref/bin/grk_compress is the reference's own
src/bin/jp2/grk_compress.cpp, compiled where it lies (oracle/Makefile, target cli) and linked against
the unmodified library; run with `-g <plugin dir> -G 0 -y in -a out` it loads
grok_b200/plugin/_build/libgrokj2k_plugin.so through grk_plugin_load and pushes a directory of
frames through grk_plugin_batch_compress.  Every output file must be byte-identical (cmp) to what
the same binary writes WITHOUT the plugin.

Decode cannot be driven from the stock grk_decompress binary in this snapshot of the reference:
grk_plugin_internal_decode_callback drops user_data (src/lib/jp2/grok.cpp:740-753), which the
binary's callback dereferences; the decode doors are driven by oracle/ref_harness.cpp instead
(tests/test_plugin_gpu.py)."""
import os
import subprocess

import numpy as np
import pytest

from synth import synth_image
from oracle import oracle as O

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(O.HERE), "oracle", "_ref", "bin")
GRK_COMPRESS = os.path.join(BIN, "grk_compress")


def _run(args, cwd):
    r = subprocess.run(args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    return r.stdout


@pytest.mark.skipif(not (os.path.exists(GRK_COMPRESS) and O.have_plugin()),
                    reason="stock grk_compress or the plugin is not built (needs /root/reference at build time)")
@pytest.mark.parametrize("name,w,h,prec,flags", [
    ("cinema4k", 640, 360, 12, ["-x", "24"]),                      # BASELINE configs[4] switches, small frames
    ("rate97", 320, 200, 8, ["-I", "-r", "12"]),                   # 9/7 + ICT, rate controlled
    ("rate53", 257, 131, 8, ["-r", "6", "-b", "32,32"]),           # 5/3 + RCT, rate controlled, odd sizes
    # HTJ2K, every block kept.  Not `-r 1`: the CLI turns that into "no rate target", and a plugin-driven
    # encode without one writes empty packets (reference defect: makeSingleLosslessLayer() runs before
    # compress_synch_with_plugin, tile/TileProcessor.cpp:1062 vs :1086)
    ("htj2k", 320, 200, 12, ["-M", "64", "-r", "1.0001"]),
])
def test_stock_grk_compress_binary_with_plugin_is_byte_identical(gpu, tmp_path, name, w, h, prec, flags):
    ind = tmp_path / "in"
    ind.mkdir()
    for f in range(3):
        O.write_pnm(str(ind / ("f%03d.ppm" % f)), synth_image(w, h, 3, prec, seed=900 + f), prec)
    cpu, acc = tmp_path / "cpu", tmp_path / "acc"
    cpu.mkdir()
    acc.mkdir()
    base = [GRK_COMPRESS, "-y", str(ind), "-O", "j2k"] + flags
    _run(base + ["-a", str(cpu)], str(tmp_path))
    log = _run(base + ["-a", str(acc), "-g", O.PLUGIN_DIR, "-G", "0"], str(tmp_path))
    for f in range(3):
        a = np.fromfile(str(cpu / ("f%03d.j2k" % f)), dtype=np.uint8)
        b = np.fromfile(str(acc / ("f%03d.j2k" % f)), dtype=np.uint8)
        assert a.size > 1000
        assert a.size == b.size and np.array_equal(a, b), "%s frame %d differs\n%s" % (name, f, log[-1500:])


@pytest.mark.skipif(not (os.path.exists(GRK_COMPRESS) and O.have_plugin()),
                    reason="stock grk_compress or the plugin is not built")
def test_stock_grk_compress_binary_dci4k_frame(gpu, tmp_path):
    """one full DCI-4K 12-bit frame, `grk_compress -x 24 -g ...` against `grk_compress -x 24`"""
    ind = tmp_path / "in"
    ind.mkdir()
    O.write_pnm(str(ind / "f000.ppm"), synth_image(4096, 2160, 3, 12, seed=24), 12)
    cpu, acc = tmp_path / "cpu", tmp_path / "acc"
    cpu.mkdir()
    acc.mkdir()
    base = [GRK_COMPRESS, "-y", str(ind), "-O", "j2k", "-x", "24"]
    _run(base + ["-a", str(cpu)], str(tmp_path))
    _run(base + ["-a", str(acc), "-g", O.PLUGIN_DIR, "-G", "0"], str(tmp_path))
    a = np.fromfile(str(cpu / "f000.j2k"), dtype=np.uint8)
    b = np.fromfile(str(acc / "f000.j2k"), dtype=np.uint8)
    assert 1000000 < a.size <= 1302083
    assert a.size == b.size and np.array_equal(a, b)
