"""GPU parity for the grid consumers of the GPIS kernels: GPIS.topcd surface extraction (CUDA stencil +
ordered compaction, csrc/cdx_grid.cu) and the *_gpis.npz writer, against the reference-generated golden vector and the
oracle.  Index / byte work: bit-exact."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _gpis(name="banana"):
    from compliancedex_b200 import GPIS, state_path
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    return g


def test_topcd_against_reference_golden(golden_dir):
    ref = np.load(os.path.join(golden_dir, "gpis_topcd_banana18.npz"))
    g = _gpis()
    steps = ref["mean"].shape[0]
    pts, nrm = g.topcd(ref["mean"], ref["normal"], ref["lb"], ref["ub"], steps)
    assert np.array_equal(pts, ref["points"]) and np.array_equal(nrm, ref["normals"])
    pts, nrm = g.topcd(torch.from_numpy(ref["mean"]), torch.from_numpy(ref["normal"]), ref["lb"], ref["ub"], steps)
    assert np.array_equal(pts, ref["points_tensor_in"]) and np.array_equal(nrm, ref["normals_tensor_in"])


@pytest.mark.parametrize("steps", [1, 2, 3, 7, 64, 100])
def test_topcd_random_grids_against_oracle(steps):
    """edge sizes (no interior voxel at all for steps <= 2), block-boundary sizes and the reference's default 100"""
    from oracle.gpis_oracle import topcd
    g = _gpis()
    rng = np.random.default_rng(steps)
    mean = rng.standard_normal((steps, steps, steps)).astype(np.float32)
    if steps >= 7:                                      # a smooth blob so that the surface is a thin shell as well
        ax = np.linspace(-1, 1, steps, dtype=np.float32)
        r = np.sqrt(ax[:, None, None] ** 2 + ax[None, :, None] ** 2 + ax[None, None, :] ** 2)
        mean = np.where(rng.random((steps, steps, steps)) < 0.5, r - 0.6, mean).astype(np.float32)
    normal = rng.standard_normal((steps, steps, steps, 3)).astype(np.float32)
    lb, ub = [-0.1, -0.2, -0.3], [0.1, 0.25, 0.3]
    p0, n0 = topcd(mean, normal, lb, ub, steps)
    p1, n1 = g.topcd(mean, normal, lb, ub, steps)
    assert p1.shape == p0.shape and np.array_equal(p1, p0) and np.array_equal(n1, n0)


def test_topcd_all_inside_and_nan():
    """all-negative grid has no surface; NaN compares as 'not internal' like numpy's `<`"""
    from oracle.gpis_oracle import topcd
    g = _gpis()
    steps = 9
    mean = -np.ones((steps, steps, steps), dtype=np.float32)
    normal = np.zeros((steps, steps, steps, 3), dtype=np.float32)
    p, n = g.topcd(mean, normal, [-1] * 3, [1] * 3, steps)
    assert p.shape == (0, 3) and n.shape == (0, 3)
    mean[4, 4, 4] = np.nan
    p0, _ = topcd(mean, normal, [-1] * 3, [1] * 3, steps)
    p1, _ = g.topcd(mean, normal, [-1] * 3, [1] * 3, steps)
    assert len(p0) == 6 and np.array_equal(p0, p1)


def test_visualization_grid_and_writer(golden_dir, tmp_path):
    """get_visualization_data vs the reference's slab loop (steps = 18) and the *_gpis.npz layout
    (train_gpis_completion.py:60-61: keys mean, var, normal, ub, lb; float32 grids)"""
    ref = np.load(os.path.join(golden_dir, "gpis_topcd_banana18.npz"))
    g = _gpis()
    steps = ref["mean"].shape[0]
    out = str(tmp_path / "banana_gpis.npz")
    mean, std, nrm, lb, ub = g.save_visualization_data(out, ref["lb"].tolist(), ref["ub"].tolist(), steps)
    assert mean.dtype == np.float32 and nrm.shape == (steps, steps, steps, 3)
    assert np.abs(mean - ref["mean"]).max() <= 1e-5 * np.abs(ref["mean"]).max()
    assert np.abs(std.astype(np.float64) ** 2 - ref["std"].astype(np.float64) ** 2).max() <= 2e-6 * float(g.R) ** 3
    assert np.abs(nrm - ref["normal"]).max() <= 2e-5
    z = np.load(out)
    assert sorted(z.files) == ["lb", "mean", "normal", "ub", "var"]
    assert np.array_equal(z["mean"], mean) and np.array_equal(z["var"], std) and z["mean"].dtype == np.float32
    # the grid feeds topcd exactly as in the reference's __main__ (gpis.py:265-267)
    p, n = g.topcd(mean, nrm, lb, ub, steps)
    assert abs(len(p) - len(ref["points"])) <= 2
