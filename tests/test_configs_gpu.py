"""GPU parity / property tests at the sizes of BASELINE.json's other configs:
C1 banana 64^3 grid (get_visualization_data), C4 fit at N=16,384, C5-shaped Allegro batches (sharding across the
tile-variant threshold, arg-min)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_c1_visualization_grid_small_golden(golden_dir):
    """gpis.py:113-125 on the banana state, steps=6, against the reference's own output."""
    from compliancedex_b200 import GPIS, state_path
    ref = np.load(os.path.join(golden_dir, "gpis_viz_banana6.npz"))
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path("banana_state"))
    mean, std, nrm, lb, ub = g.get_visualization_data([-0.15] * 3, [0.15] * 3, steps=6)
    assert mean.dtype == np.float32 and mean.shape == (6, 6, 6) and std.shape == (6, 6, 6) and nrm.shape == (6, 6, 6, 3)
    assert np.abs(mean - ref["mean"]).max() < 1e-6 and np.abs(std - ref["std"]).max() < 1e-6
    assert np.abs(nrm - ref["normal"]).max() < 1e-5
    assert np.array_equal(lb, ref["lb"]) and np.array_equal(ub, ref["ub"])


def test_c1_full_grid_against_oracle_sample():
    """configs[0]: banana state, 64^3 grid.  The full grid on the GPU; 2,048 grid points re-evaluated by the oracle."""
    from compliancedex_b200 import GPIS, state_path
    from oracle.gpis_oracle import GPISOracle, viz_grid
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path("banana_state"))
    mean, std, nrm, _, _ = g.get_visualization_data([-0.15] * 3, [0.15] * 3, steps=64)
    assert mean.shape == (64, 64, 64) and np.isfinite(mean).all() and np.isfinite(std).all()
    o = GPISOracle(0.08, 1.0).load_npz(state_path("banana_state"))
    grid = viz_grid([-0.15] * 3, [0.15] * 3, 64).reshape(-1, 3)
    idx = torch.randperm(grid.shape[0], generator=torch.Generator().manual_seed(0))[:2048]
    m_ref, s_ref = o.pred_diag(grid[idx])
    n_ref = o.compute_normal(grid[idx])
    prior = float(o.R) ** 3
    m = torch.from_numpy(mean.reshape(-1))[idx].double()
    s = torch.from_numpy(std.reshape(-1))[idx].double()
    assert (m - m_ref).abs().max() <= 2e-7 * m_ref.abs().max() + 1e-7          # outputs are float32 (gpis.py:117-119)
    assert (s ** 2 - s_ref ** 2).abs().max() <= 1e-6 * prior
    assert (torch.from_numpy(nrm.reshape(-1, 3))[idx].double() - n_ref).abs().max() < 1e-5


def test_c4_fit_16384_properties():
    """configs[3]: blocked Cholesky fit at N=16,384 (train_gpis_completion.py shape).  Size-independent checks of the
    factorisation against the matrix it was built from (torch float64 matmul is only the checker)."""
    from compliancedex_b200 import GPIS, _lib
    from compliancedex_b200.synthetic import synthetic_training_set
    import ctypes
    N = 16384
    X1, y1, noise = synthetic_training_set(N, seed=0)
    g = GPIS(0.08, 1.0)
    g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    E = g.E11
    assert E.shape == (N, N) and torch.equal(E, E.T)
    # R = max pairwise distance, diagonal = R^3 + noise^2 (gpis.py:36,40)
    R = float(g.R)
    assert abs(R - float(torch.cdist(X1[:14], X1[:14]).max())) < 1e-12         # the exterior corners span the set
    assert (torch.diagonal(E).cpu() - (R ** 3 + noise ** 2)).abs().max() < 1e-15
    lib, st = _lib.load(), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    alpha = torch.empty(N, dtype=torch.float64, device="cuda")
    _lib.check(lib.cdx_gpis_get(g._ctx, _lib.VEC_ALPHA, ctypes.c_void_p(alpha.data_ptr()), st))
    y = (y1.cuda().reshape(-1) - 1.0)
    res = (E @ alpha - y).abs().max() / y.abs().max()
    assert res < 1e-8, res                                                       # alpha = E11^-1 (y - bias)
    L = torch.empty(N, N, dtype=torch.float64, device="cuda")
    _lib.check(lib.cdx_gpis_get(g._ctx, _lib.MAT_CHOL, ctypes.c_void_p(L.data_ptr()), st))
    assert torch.equal(L, torch.tril(L)) and (torch.diagonal(L) > 0).all()
    rows = torch.randperm(N, generator=torch.Generator().manual_seed(1))[:256].cuda()
    rec = L[rows] @ L.T                                                          # 256 random rows of L L^T
    assert (rec - E[rows]).abs().max() <= 1e-12 * E.abs().max()
    del L, rec
    # posterior at training points: mean within a few noise sigmas of the label, std below the prior
    with torch.no_grad():
        m, s = g.pred(X1[14:4110].cuda())                                       # skip the 14 exterior points (sigma 0.3)
    assert (m.cpu() - y1[14:4110].reshape(-1)).abs().max() < 0.05 and (s ** 2 <= R ** 3).all()


def _allegro_opt(B, lo, hi, iters, gp, seed=7):
    from compliancedex_b200 import ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    doc = load_hand("allegro")
    d = np.load(state_path("hammer_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, B, surface_center(d["X1"], d["y1"]), seed=seed)
    opt = ProbabilisticGraspOptimizer("allegro", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm[lo:hi], num_iters=iters, optimize_target=True, ref_q=doc["ref_q"],
                                      mass=0.3, com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True,
                                      noise_seed=99, env_offset=lo)
    res = opt.optimize(q0[lo:hi].cuda(), T0[lo:hi].cuda(), k0[lo:hi].cuda(), 0.5, gp, verbose=False)
    return opt, res


def test_c5_shape_sharding_across_tile_variants():
    """configs[4] shape (Allegro, hammer state): 6,144 candidates in one batch (24,576 extended queries -> 128-query
    tiles) against the same candidates as two shards of 3,072 (-> 32-query tiles): results are bit-identical, and the
    reduced arg-min of the shards equals the arg-min of the full batch."""
    from compliancedex_b200 import GPIS, state_path
    from compliancedex_b200.distributed import reduce_pairs
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("hammer_state"))
    B, iters = 6144, 30
    full, res_full = _allegro_opt(B, 0, B, iters, gp)
    v_full = full.opt_value.clone()
    best_full = full.argmin()
    res_full = [t.clone() for t in res_full]
    pairs = []
    for lo, hi in ((0, 3072), (3072, 6144)):
        part, res = _allegro_opt(B, lo, hi, iters, gp)
        for a, b in zip(res, res_full):
            assert torch.equal(a, b[lo:hi])
        assert torch.equal(part.opt_value, v_full[lo:hi])
        pairs.append(part.argmin())
    assert reduce_pairs(pairs) == best_full
    finite = torch.isfinite(v_full)
    assert finite.float().mean() > 0.5                       # tracking started (s > 20) for most candidates
    i = int(torch.argmin(torch.where(finite, v_full, torch.full_like(v_full, float("inf")))))
    assert best_full[1] == i and best_full[0] == float(v_full[i])


def test_fit_completion_script_shape(golden_dir):
    """train_gpis_completion.py:20-62 through the drop-in: assemble (reference-golden inputs) -> fit -> E11 / y1 / R as the
    reference's fit defines them (gpis.py:33-40), state round trip, grid + surface extraction run."""
    import tempfile
    from compliancedex_b200 import GPIS
    from compliancedex_b200.completion import fit_completion
    from oracle.gpis_oracle import GPISOracle
    g = np.load(os.path.join(golden_dir, "completion_coffeebottle.npz"))
    gp = fit_completion(g["default_observed"], g["default_completed"], bias=1.0, weights=g["default_draws"])
    o = GPISOracle(0.08, 1.0)
    o.fit(torch.from_numpy(g["default_X1"]), torch.from_numpy(g["default_y1"]), torch.from_numpy(g["default_noise"]))
    assert np.abs(gp.E11.cpu().numpy() - o.E11.numpy()).max() <= 1e-15 and float(gp.R) == float(o.R)
    assert np.array_equal(gp.y1.cpu().numpy(), o.y1.numpy())
    X = (torch.rand(200, 3, dtype=torch.float64, generator=torch.Generator().manual_seed(2)) - 0.5) * 0.2
    m, s = gp.pred(X.cuda())
    mo, so = o.pred(X)
    assert (m.cpu() - mo).abs().max() <= 1e-5 * mo.abs().max() and (s.cpu() ** 2 - so ** 2).abs().max() <= 1e-6 * float(o.R) ** 3
    with tempfile.TemporaryDirectory() as d:          # the reference hard-codes the relative directory gpis_states/ (gpis.py:160,163)
        cwd = os.getcwd()
        os.makedirs(os.path.join(d, "gpis_states"))
        os.chdir(d)
        try:
            gp.save_state_data("cb_state")
            g2 = GPIS(0.08, 1.0)
            g2.load_state_data("cb_state")
        finally:
            os.chdir(cwd)
        m2, _ = g2.pred(X.cuda())
        assert (m2 - m).abs().max() <= 1e-9
