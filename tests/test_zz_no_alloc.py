"""uqk_bootstrap_noalloc through the C-ABI: the allocation-free bootstrap for callers that cannot allocate device
memory while running (GPU memory owned by an arena / caching allocator; cudaMalloc banned in a hot loop).  Its
fixed device scratch is shared by all calls, so the test makes two calls in a row with different index vectors
and different n_boot and checks each against the allocating entry point uqk_bootstrap (same kernels, scratch from the
stream-ordered allocator); atomics make the summation order vary, hence rtol 1e-12 rather than equality.
Collected last on purpose (file name)."""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

from conftest import PKG_NAME

pytestmark = pytest.mark.gpu


def _call(fn, cols, y, idx, out):
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    b, n = idx.shape
    return fn(cols[0].data_ptr(), cols[1].data_ptr(), cols[2].data_ptr(), cols[3].data_ptr(), y.data_ptr(),
              idx.data_ptr(), int(b), int(n), out.data_ptr(), stream)


def test_noalloc_bootstrap_matches_allocating_entry_point_twice():
    lib_mod = importlib.import_module(PKG_NAME + "._lib")
    lib = lib_mod.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    g = torch.Generator(device="cpu").manual_seed(7)
    n = 50_000
    cols = [torch.rand(n, generator=g).to(dev) for _ in range(4)]
    y = (torch.rand(n, generator=g) < 0.3).to(torch.int32).to(dev)
    for b, seed in ((37, 1), (5, 2)):            # second call: fewer iterations, other indices -> stale sums must not leak
        idx = torch.randint(0, n, (b, n), generator=torch.Generator().manual_seed(seed), dtype=torch.int64).to(dev)
        want = torch.empty((b, 6), dtype=torch.float64, device=dev)
        got = torch.full((b, 6), -1.0, dtype=torch.float64, device=dev)
        lib_mod.check(_call(lib.uqk_bootstrap, cols, y, idx, want), "uqk_bootstrap")
        lib_mod.check(_call(lib.uqk_bootstrap_noalloc, cols, y, idx, got), "uqk_bootstrap_noalloc")
        torch.cuda.synchronize()
        np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-12, atol=0)
        # and against direct NumPy gather-means of every output column, the class-conditional ones included
        # (row layout of uqk_bootstrap: variance overall / class 0 / class 1, then the means of the other three columns)
        ih, yh = idx.cpu().numpy(), y.cpu().numpy()
        c = [t.cpu().numpy().astype(np.float64) for t in cols]
        v, yy = c[0][ih], yh[ih]
        ref = np.stack([v.mean(axis=1), (v * (yy == 0)).sum(axis=1) / (yy == 0).sum(axis=1),
                        (v * (yy == 1)).sum(axis=1) / (yy == 1).sum(axis=1),
                        c[1][ih].mean(axis=1), c[2][ih].mean(axis=1), c[3][ih].mean(axis=1)], axis=1)
        np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=1e-6)          # inputs are float32
    # sizes beyond the fixed scratch are rejected with a clear error, nothing is launched
    idx = torch.zeros((1, 8), dtype=torch.int64, device=dev)
    out = torch.empty((2000, 6), dtype=torch.float64, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.uqk_bootstrap_noalloc(cols[0].data_ptr(), cols[1].data_ptr(), cols[2].data_ptr(), cols[3].data_ptr(),
                                   y.data_ptr(), idx.data_ptr(), 2000, 8, out.data_ptr(), stream)
    assert rc == -2 and b"exceeds the fixed scratch" in lib.uqk_last_error()
