"""GPU parity: CUDA GPIS path (through the C ABI / drop-in class) vs the oracle and the reference-generated
golden vectors.  Gates (BASELINE.json): mean, grad_mean <= 1e-5 rel; std^2 within 1e-6 * prior (R^3) abs."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ["mug", "mug2", "coffeebottle", "lego", "hammer", "banana", "dummy"]


def _gpis(name):
    from compliancedex_b200 import GPIS, state_path
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    return g


def _check(got_mean, got_std, got_gm, got_gs, ref, prior):
    gm_scale = np.abs(ref["grad_mean"]).max()
    assert np.abs(got_mean - ref["mean"]).max() <= 1e-5 * np.abs(ref["mean"]).max()
    rel_pt = np.abs(got_mean - ref["mean"]) / np.maximum(np.abs(ref["mean"]), 1e-3)
    assert rel_pt.max() <= 1e-5
    assert np.abs(got_gm - ref["grad_mean"]).max() <= 1e-5 * gm_scale
    assert np.abs(got_std ** 2 - ref["std"] ** 2).max() <= 1e-6 * prior
    if got_gs is not None:
        assert np.abs(got_gs - ref["grad_std"]).max() <= 1e-5 * np.abs(ref["grad_std"]).max()


@pytest.mark.parametrize("name", NAMES)
def test_shipped_states_against_reference_golden(name, golden_dir):
    ref = np.load(os.path.join(golden_dir, f"gpis_{name}.npz"))
    g = _gpis(name)
    X = torch.from_numpy(ref["X"]).cuda()
    prior = float(g.R) ** 3
    mean, std, gm, gs = g.pred_with_grads(X)                                  # extended kernel (E11^-1)
    _check(mean.cpu().numpy(), std.cpu().numpy(), gm.cpu().numpy(), gs.cpu().numpy(), ref, prior)
    with torch.no_grad():
        m2, s2 = g.pred(X)                                                    # unit kernel (L^-1)
    _check(m2.cpu().numpy(), s2.cpu().numpy(), gm.cpu().numpy(), None, ref, prior)
    n = g.compute_normal(X)                                                   # mean-only kernel (TMA-staged)
    assert np.abs(n.cpu().numpy() - ref["normal"]).max() <= 1e-5
    # autograd through pred, as the optimizers use it
    Xr = X.clone().requires_grad_(True)
    m3, s3 = g.pred(Xr)
    (m3.sum() * 2.0 + s3.sum() * 3.0).backward()
    want = 2.0 * ref["grad_mean"] + 3.0 * ref["grad_std"]
    assert np.abs(Xr.grad.cpu().numpy() - want).max() <= 1e-5 * np.abs(want).max()


def test_batched_shape_rule(golden_dir):
    ref = np.load(os.path.join(golden_dir, "gpis_batched_lego.npz"))
    g = _gpis("lego")
    X = torch.from_numpy(ref["X"]).cuda()
    m, s = g.pred(X)
    assert tuple(m.shape) == (5, 4) and tuple(s.shape) == (5, 4) and m.dtype == torch.float64
    assert np.abs(m.cpu().numpy() - ref["mean"]).max() < 1e-6 and np.abs(s.cpu().numpy() - ref["std"]).max() < 1e-6
    n = g.compute_normal(X)
    assert tuple(n.shape) == (5, 4, 3) and np.abs(n.cpu().numpy() - ref["normal"]).max() < 1e-5


@pytest.mark.parametrize("kern", ["tps", "rbf", "joint"])
def test_fit_against_reference_golden(kern, golden_dir):
    from compliancedex_b200 import GPIS
    ref = np.load(os.path.join(golden_dir, "gpis_fit_synth211.npz"))
    g = GPIS(0.08, 1.0, kernel=kern)
    g.fit(torch.from_numpy(ref["X1"]).cuda(), torch.from_numpy(ref["y1"]).cuda(), torch.from_numpy(ref["noise"]).cuda())
    E = g.E11.cpu().numpy()
    assert np.abs(E - ref[f"{kern}_E11"]).max() <= 1e-12 * np.abs(ref[f"{kern}_E11"]).max()
    if kern != "rbf":
        assert abs(float(g.R) - float(ref[f"{kern}_R"])) <= 1e-14
    assert np.abs(g.y1.cpu().numpy() - ref[f"{kern}_y1"]).max() == 0.0
    prior = {"tps": float(ref["tps_R"]) ** 3, "rbf": 1.0, "joint": 0.3 + 0.7 * float(ref["tps_R"]) ** 3}[kern]
    with torch.no_grad():
        m, s = g.pred(torch.from_numpy(ref["Xq"]).cuda())
    assert np.abs(m.cpu().numpy() - ref[f"{kern}_mean"]).max() <= 1e-5 * np.abs(ref[f"{kern}_mean"]).max()
    assert np.abs(s.cpu().numpy() ** 2 - ref[f"{kern}_std"] ** 2).max() <= 1e-6 * prior
    n = g.compute_normal(torch.from_numpy(ref["Xq"]).cuda())
    assert np.abs(n.cpu().numpy() - ref[f"{kern}_normal"]).max() <= 1e-5


def test_factorisation_identities():
    """L L^T = E11, U L = I, P E11 = I on a shipped state (padding must not leak)."""
    from compliancedex_b200 import _lib
    g = _gpis("hammer")
    N = g.X1.shape[0]
    lib = _lib.load()
    mats = {}
    for nm, which in [("L", _lib.MAT_CHOL), ("U", _lib.MAT_CHOL_INV), ("P", _lib.MAT_E11_INV), ("E", _lib.MAT_E11)]:
        t = torch.empty(N, N, dtype=torch.float64, device="cuda")
        _lib.check(lib.cdx_gpis_get(g._ctx, which, ctypes.c_void_p(t.data_ptr()), None))
        mats[nm] = t
    torch.cuda.synchronize()
    E, L, U, P = mats["E"], mats["L"], mats["U"], mats["P"]
    assert torch.equal(E, g.E11)
    assert (L @ L.T - E).abs().max() <= 1e-13 * E.abs().max()
    assert torch.triu(L, 1).abs().max() == 0 and torch.triu(U, 1).abs().max() == 0
    I = torch.eye(N, dtype=torch.float64, device="cuda")
    assert (U @ L - I).abs().max() <= 1e-9
    assert (P @ E - I).abs().max() <= 1e-6
    assert (P - P.T).abs().max() == 0


def test_synthetic_large_state_vs_oracle():
    """N=1024 synthetic fit (SURVEY 8d generator): CUDA fit+query vs the oracle restatement on 512 points; exercises
    the 128-wide tile path and the multi-level triangular inverse."""
    from compliancedex_b200 import GPIS
    from oracle.gpis_oracle import GPISOracle, synthetic_training_set
    X1, y1, noise = synthetic_training_set(1024, seed=0)
    o = GPISOracle(0.08, 1.0).fit(X1, y1, noise)
    g = GPIS(0.08, 1.0)
    g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    X = (torch.rand(512, 3, generator=torch.Generator().manual_seed(4), dtype=torch.float64) * 2 - 1) * 0.1
    m_ref, s_ref = o.pred_diag(X)
    gm_ref, gs_ref = o.grads(X)
    mean, std, gm, gs = g.pred_with_grads(X.cuda())
    prior = float(o.R) ** 3
    assert (mean.cpu() - m_ref).abs().max() <= 1e-5 * m_ref.abs().max()
    assert (gm.cpu() - gm_ref).abs().max() <= 1e-5 * gm_ref.abs().max()
    assert (std.cpu() ** 2 - s_ref ** 2).abs().max() <= 1e-6 * prior
    assert (gs.cpu() - gs_ref).abs().max() <= 1e-4 * gs_ref.abs().max()
    with torch.no_grad():
        m2, s2 = g.pred(X.cuda())
    assert (m2.cpu() - m_ref).abs().max() <= 1e-5 * m_ref.abs().max()
    assert (s2.cpu() ** 2 - s_ref ** 2).abs().max() <= 1e-6 * prior


def test_edge_cases():
    g = _gpis("mug")
    # empty query
    m, s = g.pred(torch.zeros(0, 3, dtype=torch.float64, device="cuda"))
    assert m.numel() == 0 and s.numel() == 0
    # ragged sizes around the 128-query tile, query exactly on training points, host-buffer entry point
    from compliancedex_b200 import _lib
    lib = _lib.load()
    for M in [1, 127, 128, 129, 300]:
        X = torch.vstack([g.X1[:min(M, 50)], (torch.rand(max(M - 50, 0), 3, dtype=torch.float64, device="cuda") - 0.5) * 0.2])[:M]
        mean, std, gm, gs = g.pred_with_grads(X)
        assert torch.isfinite(mean).all() and torch.isfinite(std).all() and torch.isfinite(gm).all()
        Xh = X.cpu().contiguous()
        mh = torch.empty(M, dtype=torch.float64)
        sh = torch.empty(M, dtype=torch.float64)
        gh = torch.empty(M, 3, dtype=torch.float64)
        _lib.check(lib.cdx_gpis_query_host(g._ctx, Xh.data_ptr(), M, mh.data_ptr(), gh.data_ptr(), sh.data_ptr(), None, None))
        with torch.no_grad():
            m2, s2 = g.pred(X)
        assert torch.equal(mh, m2.cpu()) and torch.equal(sh, s2.cpu())
        assert (mean.cpu() - mh).abs().max() < 1e-9
    # non-SPD matrix raises like torch.linalg would
    from compliancedex_b200 import GPIS
    bad = GPIS(0.08, 1.0)
    with pytest.raises(torch.linalg.LinAlgError):
        bad._adopt(g.R, g.X1, g.y1, -g.E11, g.bias)


def test_full_size_properties():
    """Size-independent properties at the bench size (N=4096 synthetic, 2^18 queries): chunking invariance and
    std^2 <= prior, std -> noise level on training points."""
    from compliancedex_b200 import GPIS
    from oracle.gpis_oracle import synthetic_training_set
    X1, y1, noise = synthetic_training_set(4096, seed=0)
    g = GPIS(0.08, 1.0)
    g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    X = (torch.rand(1 << 18, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * 0.1
    X = X.cuda()
    with torch.no_grad():
        m, s = g.pred(X)
        m1, s1 = g.pred(X[1000:5000])
    assert torch.equal(m[1000:5000], m1) and torch.equal(s[1000:5000], s1)
    assert torch.isfinite(m).all() and (s ** 2 <= float(g.R) ** 3 * (1 + 1e-9)).all()
    n = g.compute_normal(X[:4096])
    assert (n.norm(dim=1) - 1).abs().max() < 1e-6


@pytest.mark.parametrize("name", ["lego", "hammer"])
def test_normal_index_and_multinormals_against_reference_golden(name, golden_dir):
    """GPIS.compute_normal(X2, index) (gpis.py:63-87, index branch) and GPIS.compute_multinormals (gpis.py:89-111) through
    cdx_gpis_query_weights, vs goldens produced by the reference's own gpis.py (oracle/make_golden_normals.py)."""
    ref = np.load(os.path.join(golden_dir, "gpis_normals_index.npz"))
    g = _gpis(name)
    X = torch.from_numpy(ref[f"{name}_X"]).cuda()
    for tag in ["prefix", "suffix", "scattered", "single"]:
        n, w = g.compute_normal(X, ref[f"{name}_{tag}_index"].tolist())
        assert tuple(n.shape) == (48, 3) and n.dtype == torch.float64
        assert np.abs(n.cpu().numpy() - ref[f"{name}_{tag}_normal"]).max() <= 1e-5, tag
        wr = float(ref[f"{name}_{tag}_weight"])
        assert abs(float(w) - wr) <= 1e-5 * abs(wr), tag
    nb, _ = g.compute_normal(X[:40].reshape(10, 4, 3), ref[f"{name}_prefix_index"].tolist())
    assert tuple(nb.shape) == (10, 4, 3)
    assert np.abs(nb.cpu().numpy() - ref[f"{name}_prefix_normal_batched"]).max() <= 1e-5
    for S in (3, 5):
        g.fraction = None                                   # the row subsets are cached by the first call (gpis.py:96)
        nm, wm = g.compute_multinormals(X, S)
        assert [len(i) for i in g.indices] == ref[f"{name}_multi{S}_sizes"].tolist()
        assert [i[0] for i in g.indices] == ref[f"{name}_multi{S}_first"].tolist()
        assert tuple(nm.shape) == (48, S, 3) and tuple(wm.shape) == (S,)
        assert np.abs(nm.cpu().numpy() - ref[f"{name}_multi{S}_normals"]).max() <= 1e-5
        assert np.abs(wm.cpu().numpy() - ref[f"{name}_multi{S}_weights"]).max() <= 1e-5 * np.abs(ref[f"{name}_multi{S}_weights"]).max()


@pytest.mark.parametrize("name", NAMES)
def test_extended_query_kernels_agree_and_are_batch_invariant(name, golden_dir):
    """the extended query has three kernels for small states: the warp-specialised all-columns kernel (default, mode 0: k*
    generated once per query tile by generator warps, E11^-1 streamed by bulk TMA), its single-role predecessor (mode 2) and
    the column-tiled one (mode 1).  All against the reference golden; the all-columns kernels' results must not depend on
    the batch (16- vs 32-query tiles, position in the batch)."""
    ref = np.load(os.path.join(golden_dir, f"gpis_{name}.npz"))
    g = _gpis(name)
    prior = float(g.R) ** 3
    X = torch.from_numpy(ref["X"]).cuda()
    outs = {}
    for mode in (0, 1, 2):
        g.set_ext_mode(mode)
        mean, std, gm, gs = g.pred_with_grads(X)
        _check(mean.cpu().numpy(), std.cpu().numpy(), gm.cpu().numpy(), gs.cpu().numpy(), ref, prior)
        outs[mode] = (mean, std, gm, gs)
    for m in (0, 2):
        for a, b in zip(outs[m], outs[1]):
            assert (a - b).abs().max() <= 1e-9 * max(1.0, float(b.abs().max()))
    b = 0.18 if name in ("hammer", "coffeebottle") else (0.15 if name == "banana" else 0.1)
    big = ((torch.rand(148 * 32 * 2 + 37, 3, generator=torch.Generator().manual_seed(3), dtype=torch.float64) * 2 - 1) * b).cuda()
    for mode in (0, 2):
        g.set_ext_mode(mode)
        full = g.pred_with_grads(big)                       # 32-query tiles, persistent CTAs walking several tiles
        part = g.pred_with_grads(big[1001:1001 + 333])      # 16-query tiles, other positions
        for a, b_ in zip(full, part):
            assert torch.equal(a[1001:1001 + 333], b_)
    g.set_ext_mode(0)


def test_refit_across_state_sizes_keeps_the_extended_query_consistent():
    """one context refitted 150 -> 700 -> 330 -> 430 training points: the slab-ordered copy of E11^-1 that the warp-specialised
    extended kernel streams (states of <= 448 padded points) is rebuilt with every fit, dropped / re-created as the state
    crosses 448, and the kernel (mode 0) keeps agreeing with the column-tiled one (mode 1) and with the oracle."""
    from compliancedex_b200 import GPIS
    from oracle.gpis_oracle import GPISOracle, synthetic_training_set
    g = GPIS(0.08, 1.0)
    gen = torch.Generator().manual_seed(11)
    for N in (150, 700, 330, 430):
        X1, y1, noise = synthetic_training_set(N, seed=N)
        g.fit(X1.cuda(), y1.cuda(), noise.cuda())
        X = ((torch.rand(777, 3, generator=gen, dtype=torch.float64) * 2 - 1) * 0.1).cuda()
        outs = {}
        for mode in (0, 1):
            g.set_ext_mode(mode)
            outs[mode] = g.pred_with_grads(X)
        g.set_ext_mode(0)
        for a, b in zip(outs[0], outs[1]):
            assert (a - b).abs().max() <= 1e-9 * max(1.0, float(b.abs().max())), N
        o = GPISOracle(0.08, 1.0).fit(X1, y1, noise)
        m, sd = o.pred_diag(X.cpu()[:64])
        prior = float(o.R) ** 3
        assert (outs[0][0][:64].cpu() - m.reshape(-1)).abs().max() <= 1e-5 * max(1.0, float(m.abs().max())), N
        assert (outs[0][1][:64].cpu() ** 2 - sd.reshape(-1) ** 2).abs().max() <= 1e-6 * prior, N


@pytest.mark.timeout(300)
@pytest.mark.parametrize("N", [3, 16, 17, 64, 65, 447, 448])
def test_extended_query_tiny_and_boundary_states(N):
    """the warp-specialised extended kernel at the corners of its shape space: states of a few points (one k chunk, one column
    block), exactly 64 / 448 padded points, 449 -> the column-tiled kernel; batches of 1, one tile +- 1, more tiles than SMs.
    Mode 0 against mode 1 (same w bit for bit, outputs to 1e-9) and finite everywhere."""
    from compliancedex_b200 import GPIS
    from oracle.gpis_oracle import synthetic_training_set
    X1, y1, noise = synthetic_training_set(max(N, 64), seed=100 + N)
    X1, y1, noise = X1[:N], y1[:N], noise[:N]
    g = GPIS(0.08, 1.0)
    g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    gen = torch.Generator().manual_seed(N)
    for M in (1, 15, 16, 17, 33, 148 * 32 + 1):
        X = ((torch.rand(M, 3, generator=gen, dtype=torch.float64) * 2 - 1) * 0.1).cuda()
        outs = {}
        for mode in (0, 1):
            g.set_ext_mode(mode)
            outs[mode] = g.pred_with_grads(X)
        g.set_ext_mode(0)
        for a, b in zip(outs[0], outs[1]):
            assert torch.isfinite(a).all()
            assert (a - b).abs().max() <= 1e-9 * max(1.0, float(b.abs().max())), (N, M)


@pytest.mark.gpu
def test_fit_is_bit_reproducible_across_runs_and_streams():
    """The 128 x 128 GEMM of the fit synchronises its shared-memory ring with full / empty mbarriers (no CTA barrier in the k
    loop) and the Cholesky runs its serial chain on a second stream: a race would show as run-to-run differences.  Five fits of
    the same N = 2,304 set (one on a side stream) must give the same factor, inverse factor and inverse bit for bit, and the
    factor must satisfy L L^T = E11."""
    from compliancedex_b200 import GPIS, _lib
    from compliancedex_b200.synthetic import synthetic_training_set
    lib = _lib.load()
    X1, y1, noise = synthetic_training_set(2304, seed=3)
    X1, y1, noise = X1.cuda(), y1.cuda(), noise.cuda()
    N = X1.shape[0]

    def mats(g):
        out = []
        for which in (_lib.MAT_CHOL, _lib.MAT_CHOL_INV, _lib.MAT_E11_INV):
            t = torch.empty(N, N, dtype=torch.float64, device="cuda")
            _lib.check(lib.cdx_gpis_get(g._ctx, which, ctypes.c_void_p(t.data_ptr()), None))
            out.append(t)
        torch.cuda.synchronize()
        return out

    g = GPIS(0.08, 1.0)
    g.fit(X1, y1, noise)
    ref = mats(g)
    E = g.E11
    assert (ref[0] @ ref[0].T - E).abs().max() <= 1e-13 * E.abs().max()
    side = torch.cuda.Stream()
    for rep in range(4):
        if rep == 2:
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                g.fit(X1, y1, noise)
            torch.cuda.current_stream().wait_stream(side)
        else:
            g.fit(X1, y1, noise)
        for a, b in zip(ref, mats(g)):
            assert torch.equal(a, b)
