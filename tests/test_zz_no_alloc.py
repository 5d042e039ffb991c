"""The allocation-free door (include/grk_b200.h section D, csrc/noalloc.cu): the inverse wavelet of
one tile-component on device buffers with its ping-pong planes in the library's built-in
__device__ array.  Checked bit for bit against the existing door (grkb_dwt_host, which cudaMallocs
its own planes) on two calls in a row that share the array, and the size limit is refused.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from common import plane, same_bits

pytestmark = pytest.mark.gpu


def _noalloc_idwt(gpu, torch, src, x0, y0, numres, reversible):
    """src: host plane (h, w); device buffers come from torch (the caller's side of the contract),
    the library call itself allocates nothing"""
    h, w = src.shape
    stride = (w + 31) // 32 * 32
    host = np.zeros((h, stride), dtype=np.int32)
    host[:, :w] = src.view(np.int32)
    d_in = torch.from_numpy(host).cuda()
    d_out = torch.full((h, stride), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = gpu.lib.grkb_idwt_device_noalloc(d_in.data_ptr(), stride, d_out.data_ptr(), stride, x0, y0,
                                          x0 + w, y0 + h, numres, int(reversible), None)
    gpu._lib.check(rc, "grkb_idwt_device_noalloc")
    torch.cuda.synchronize()
    return d_out.cpu().numpy()[:, :w]


@pytest.mark.parametrize("reversible", [1, 0])
def test_noalloc_idwt_matches_allocating_door_twice(gpu, reversible):
    import torch
    from grok_b200 import stages
    gpu._lib.check(gpu.lib.grkb_set_device(0), "grkb_set_device")
    rng = np.random.default_rng(77 + reversible)
    # two calls in a row on the shared array: a large geometry first, then a smaller, ragged one
    # at an odd origin whose intermediate planes land on what the first call left behind
    for (w, h, x0, y0, nr) in [(1024, 333, 0, 0, 5), (481, 66, 3, 1, 4), (2048, 2048, 0, 0, 6)]:
        src = plane(rng, w, h, reversible)
        want = stages.dwt(src, x0, y0, nr, reversible, 0)
        got = _noalloc_idwt(gpu, torch, src, x0, y0, nr, reversible)
        assert same_bits(want, got), (w, h, x0, y0, nr)


def test_noalloc_rejects_sizes_beyond_the_array(gpu):
    import torch
    gpu._lib.check(gpu.lib.grkb_set_device(0), "grkb_set_device")
    w, h = 4096, 2160  # scratch 2 x 2048 x 1080 x 4 = 17.7 MB > 8 MiB
    d = torch.zeros((8, 4096), dtype=torch.int32, device="cuda")
    rc = gpu.lib.grkb_idwt_device_noalloc(d.data_ptr(), 4096, d.data_ptr() + 4096 * 16, 4096, 0, 0, w, h,
                                          6, 1, None)
    assert rc == -2
    assert b"limit" in gpu.lib.grkb_last_error()
