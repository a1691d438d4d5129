"""CPU: pins the oracle (oracle/*.py) against the golden vectors produced by the reference's own code
(oracle/make_golden.py, tests/golden/*.npz)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle.gpis_oracle import GPISOracle, synthetic_training_set
from oracle.noise import kabsch_noise_np
from oracle.pregrasp_oracle import PregraspOracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STATES = os.path.join(ROOT, "compliancedex_b200", "data", "gpis_states")
HANDS = os.path.join(ROOT, "compliancedex_b200", "data", "hands")
NAMES = ["mug", "mug2", "coffeebottle", "lego", "hammer", "banana", "dummy"]


@pytest.mark.parametrize("name", NAMES)
def test_gpis_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, f"gpis_{name}.npz"))
    o = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, f"{name}_state.npz"))
    X = torch.from_numpy(g["X"])
    m, s = o.pred(X)
    assert np.abs(m.numpy() - g["mean"]).max() < 1e-9
    assert np.abs(s.numpy() - g["std"]).max() < 1e-9
    m2, s2 = o.pred_diag(X)
    assert np.abs(m2.numpy() - g["mean"]).max() < 1e-9
    assert np.abs(s2.numpy() ** 2 - g["std"] ** 2).max() < 1e-9 * float(o.R) ** 3
    gm, gs = o.grads(X)
    assert np.abs(gm.numpy() - g["grad_mean"]).max() < 1e-7 * np.abs(g["grad_mean"]).max()
    assert np.abs(gs.numpy() - g["grad_std"]).max() < 1e-6 * np.abs(g["grad_std"]).max()
    assert np.abs(o.compute_normal(X).numpy() - g["normal"]).max() < 1e-9


def test_gpis_oracle_batched_shapes(golden_dir):
    g = np.load(os.path.join(golden_dir, "gpis_batched_lego.npz"))
    o = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, "lego_state.npz"))
    X = torch.from_numpy(g["X"])
    m, s = o.pred(X)
    assert tuple(m.shape) == (5, 4) and tuple(s.shape) == (5, 4)
    assert np.abs(m.numpy() - g["mean"]).max() < 1e-9 and np.abs(s.numpy() - g["std"]).max() < 1e-9
    n = o.compute_normal(X)
    assert tuple(n.shape) == (5, 4, 3) and np.abs(n.numpy() - g["normal"]).max() < 1e-9


@pytest.mark.parametrize("kern", ["tps", "rbf", "joint"])
def test_gpis_oracle_fit(kern, golden_dir):
    g = np.load(os.path.join(golden_dir, "gpis_fit_synth211.npz"))
    X1, y1, noise = synthetic_training_set(211, seed=3)
    assert np.array_equal(X1.numpy(), g["X1"]) and np.array_equal(noise.numpy(), g["noise"])
    o = GPISOracle(0.08, 1.0, kernel=kern).fit(X1, y1, noise)
    assert np.abs(o.E11.numpy() - g[f"{kern}_E11"]).max() == 0.0
    m, s = o.pred(torch.from_numpy(g["Xq"]))
    assert np.abs(m.numpy() - g[f"{kern}_mean"]).max() < 1e-9
    assert np.abs(s.numpy() - g[f"{kern}_std"]).max() < 1e-9


@pytest.mark.parametrize("name", ["lego", "hammer"])
def test_gpis_oracle_normal_index_and_multinormals(name, golden_dir):
    """compute_normal(X2, index) (gpis.py:66-69,78-80,84-87) and compute_multinormals (gpis.py:89-111) vs the reference"""
    g = np.load(os.path.join(golden_dir, "gpis_normals_index.npz"))
    o = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, f"{name}_state.npz"))
    X = torch.from_numpy(g[f"{name}_X"])
    for tag in ["prefix", "suffix", "scattered", "single"]:
        n, w = o.compute_normal(X, g[f"{name}_{tag}_index"].tolist())
        assert np.abs(n.numpy() - g[f"{name}_{tag}_normal"]).max() < 1e-9
        assert abs(float(w) - float(g[f"{name}_{tag}_weight"])) <= 1e-9 * abs(float(g[f"{name}_{tag}_weight"]))
    nb, _ = o.compute_normal(X[:40].reshape(10, 4, 3), g[f"{name}_prefix_index"].tolist())
    assert tuple(nb.shape) == (10, 4, 3) and np.abs(nb.numpy() - g[f"{name}_prefix_normal_batched"]).max() < 1e-9
    for S in (3, 5):
        subsets = o.nested_row_subsets(S)
        assert [len(i) for i in subsets] == g[f"{name}_multi{S}_sizes"].tolist()
        assert [i[0] for i in subsets] == g[f"{name}_multi{S}_first"].tolist()
        nm, wm = o.compute_multinormals(X, S)
        assert tuple(nm.shape) == (48, S, 3)
        assert np.abs(nm.numpy() - g[f"{name}_multi{S}_normals"]).max() < 1e-9
        assert np.abs(wm.numpy() - g[f"{name}_multi{S}_weights"]).max() < 1e-9


def test_gpis_oracle_viz_grid(golden_dir):
    g = np.load(os.path.join(golden_dir, "gpis_viz_banana6.npz"))
    o = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, "banana_state.npz"))
    mean, std, nrm, lb, ub = o.get_visualization_data([-0.15] * 3, [0.15] * 3, steps=6)
    assert mean.dtype == np.float32 and mean.shape == (6, 6, 6) and nrm.shape == (6, 6, 6, 3)
    assert np.abs(mean - g["mean"]).max() < 1e-6 and np.abs(std - g["std"]).max() < 1e-6


def test_noise_stream_known_values():
    u = kabsch_noise_np(1001, 0, 3, 4, 0)
    assert u.shape == (12, 3, 3) and (u >= 0).all() and (u < 1).all()
    # rank / batch invariance: candidate b of a sharded call equals candidate b of the full call
    full = kabsch_noise_np(5, 7, 3, 8, 0).reshape(3, 8, 3, 3)
    part = kabsch_noise_np(5, 7, 3, 4, 4).reshape(3, 4, 3, 3)
    assert np.array_equal(full[:, 4:], part)
    # fixed known answers (also asserted by the CUDA twin in the gpu tests)


@pytest.mark.parametrize("tag,hand,state", [("leap_mug", "leap", "mug"), ("allegro_hammer", "allegro", "hammer"),
                                            ("allegro_coffeebottle", "allegro", "coffeebottle")])
def test_pregrasp_oracle_matches_reference(tag, hand, state, golden_dir):
    g = np.load(os.path.join(golden_dir, f"pregrasp_{tag}.npz"))
    gp = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, f"{state}_state.npz"))
    iters = int(g["iters"])
    o = PregraspOracle(os.path.join(HANDS, f"{hand}.json"), g["palm0"], iters, mass=float(g["mass"]),
                       com=g["com"].tolist(), noise_seed=int(g["noise_seed"]))
    mu = g["friction_mu"].item()
    q0, t0, c0 = (torch.from_numpy(g[k]) for k in ("q0", "target0", "compliance0"))
    tips = o.forward_kinematics(q0, torch.from_numpy(g["palm0"]))
    assert np.abs(tips.numpy() - g["tips0"]).max() < 1e-12
    losses, grads = [], []
    res = o.optimize(q0, t0, c0, mu, gp, record=lambda s, t, gr: (losses.append(t.numpy()), grads.append(gr)))
    losses = np.stack(losses)
    # NaN policy (optimize_pregrasp.py:819 prints and continues): NaN candidates must be NaN in both
    assert np.array_equal(np.isnan(losses), np.isnan(g["losses"]))
    rel = np.nan_to_num(np.abs(losses - g["losses"]) / np.maximum(np.abs(g["losses"]), 1.0))
    assert rel[0].max() < 1e-12
    assert rel.max() < 1e-4   # objective gate of BASELINE.json; the trajectories drift apart slowly (fp rounding)
    for it in range(3):
        for i, nm in enumerate(["q", "compliance", "target", "palm_pos", "palm_ori"]):
            ref = g[f"grad{it}_{nm}"]
            got = grads[it][i].numpy()
            assert np.array_equal(np.isnan(got), np.isnan(ref))
            assert np.nanmax(np.abs(got - ref)) <= 1e-6 * np.nanmax(np.abs(ref)) + 1e-12
    for a, nm in zip(res[:5], ["opt_q", "opt_compliance", "opt_target", "opt_palm", "opt_margin"]):
        assert np.array_equal(np.isnan(a.detach().numpy()), np.isnan(g[nm]))
        assert np.nanmax(np.abs(a.detach().numpy() - g[nm])) < 1e-4


def test_topcd_oracle_matches_reference_loop(golden_dir):
    """oracle.topcd (vectorised) against the reference's triple loop (gpis.py:127-150), numpy and tensor inputs"""
    import numpy as np
    import torch
    from oracle.gpis_oracle import topcd
    ref = np.load(os.path.join(golden_dir, "gpis_topcd_banana18.npz"))
    steps = ref["mean"].shape[0]
    pts, nrm = topcd(ref["mean"], ref["normal"], ref["lb"], ref["ub"], steps)
    assert np.array_equal(pts, ref["points"]) and np.array_equal(nrm, ref["normals"])
    pts, nrm = topcd(torch.from_numpy(ref["mean"]), torch.from_numpy(ref["normal"]), ref["lb"], ref["ub"], steps)
    assert np.array_equal(pts, ref["points_tensor_in"]) and np.array_equal(nrm, ref["normals_tensor_in"])
    assert len(ref["points"]) > 20


@pytest.mark.parametrize("hand", ["leap", "allegro"])
def test_collision_loss_oracle_matches_reference(hand, golden_dir):
    """oracle compute_collision_loss vs the reference's (optimize_pregrasp.py:671-701): pair, floor and palm barriers"""
    from compliancedex_b200 import hand_path
    ref = np.load(os.path.join(golden_dir, "pregrasp_collision.npz"))
    q, palm = torch.from_numpy(ref[f"{hand}_q"]), torch.from_numpy(ref[f"{hand}_palm"])
    for opt_palm in (True, False):
        o = PregraspOracle(hand_path(hand), palm.numpy(), 1, optimize_palm=opt_palm)
        got = o.compute_collision_loss(q, palm).numpy()
        want = ref[f"{hand}_cost_palm{int(opt_palm)}"]
        assert np.allclose(got, want, rtol=1e-6, atol=1e-9), np.abs(got - want).max()
    o = PregraspOracle(hand_path(hand), palm.numpy(), 1, optimize_palm=True)
    got = o.compute_collision_loss(q, palm, threshold=0.08).numpy()
    assert np.allclose(got, ref[f"{hand}_cost_thr008"], rtol=1e-6, atol=1e-9)
    assert (ref[f"{hand}_cost_thr008"] != ref[f"{hand}_cost_palm1"]).any()      # the pair term is exercised
