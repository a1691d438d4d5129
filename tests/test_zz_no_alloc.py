"""cdx_grid_surface_static: the allocation-free variant of cdx_grid_surface (block counts in a file-scope static __device__
array, include/cdx.h).  It has no counterpart in the reference; the check is that it returns exactly what the existing entry
point returns (integer / index work: bit-exact), on two calls in a row that share the static scratch, and that it rejects
grids beyond its stated maximum.  Collected last on purpose (file name)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_STEPS = 256


def _surface(lib, fn, ctx, mean_dev, steps):
    from compliancedex_b200 import _lib
    idx = torch.full((steps ** 3,), -1, dtype=torch.int64, device=mean_dev.device)
    cnt = torch.zeros(1, dtype=torch.int64, device=mean_dev.device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(fn(ctx, mean_dev.data_ptr(), int(steps), idx.data_ptr(), cnt.data_ptr(), st))
    torch.cuda.synchronize()
    n = int(cnt.item())
    return n, idx[:n].cpu().numpy()


def _grid(steps, seed):
    r = np.random.default_rng(seed)
    ax = np.linspace(-1.0, 1.0, steps, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    m = np.sqrt(x * x + y * y + z * z) - 0.6 + 0.08 * r.standard_normal((steps, steps, steps)).astype(np.float32)
    return np.ascontiguousarray(m, dtype=np.float32)


def test_static_scratch_matches_allocating_entry_point_twice():
    from compliancedex_b200 import _lib
    lib = _lib.load()
    ctx = _lib.new_context(0)
    dev = torch.device("cuda", 0)
    # two calls in a row on different grids and sizes: the second runs on whatever the first left in the shared scratch
    # (a larger grid first, so that stale counts beyond the second call's blocks are really there)
    for steps, seed in ((96, 1), (40, 2), (MAX_STEPS, 3)):
        m = torch.from_numpy(_grid(steps, seed)).to(dev)
        n0, i0 = _surface(lib, lib.cdx_grid_surface, ctx, m, steps)
        n1, i1 = _surface(lib, lib.cdx_grid_surface_static, ctx, m, steps)
        n2, i2 = _surface(lib, lib.cdx_grid_surface_static, ctx, m, steps)
        assert n0 > 0
        assert n1 == n0 and n2 == n0
        assert np.array_equal(i1, i0) and np.array_equal(i2, i0)
        assert np.all(np.diff(i1) > 0)
    ctx.close()


def test_static_scratch_rejects_oversize_grid():
    from compliancedex_b200 import _lib
    lib = _lib.load()
    ctx = _lib.new_context(0)
    dev = torch.device("cuda", 0)
    m = torch.zeros(8, dtype=torch.float32, device=dev)      # never read: the size check comes first
    idx = torch.zeros(8, dtype=torch.int64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    rc = lib.cdx_grid_surface_static(ctx, m.data_ptr(), MAX_STEPS + 1, idx.data_ptr(), cnt.data_ptr(), None)
    assert rc == 1                                            # CDX_ERR_ARG
    assert b"exceeds the static scratch" in lib.cdx_last_error()
    ctx.close()
