"""CPU: the per-candidate math the pregrasp kernels run (cdx_pg_math.cuh, compiled for the host by tests/hostcheck)
against the oracle's autograd - FK + reference-compatible Jacobian, 3x3 SVD, closure value and all leaf gradients."""
import json
import os

import numpy as np
import pytest
import torch

from compliancedex_b200.hand import load_hand, tree_for
from oracle.gpis_oracle import GPISOracle
from oracle.pregrasp_oracle import PregraspOracle
from tests.hostcheck import hc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STATES = os.path.join(ROOT, "compliancedex_b200", "data", "gpis_states")
HANDS = os.path.join(ROOT, "compliancedex_b200", "data", "hands")


def _host(hand):
    doc = load_hand(hand)
    return hc.Host(tree_for(doc, doc["ee_link_name"], doc["ee_link_offset"])), doc


@pytest.mark.parametrize("hand", ["leap", "allegro"])
def test_fk_and_reference_jacobian(hand):
    host, doc = _host(hand)
    o = PregraspOracle(os.path.join(HANDS, f"{hand}.json"), np.zeros((1, 6)))
    g = torch.Generator().manual_seed(3)
    ref_q = torch.tensor(doc["ref_q"], dtype=torch.float64)
    for trial in range(12):
        scale = 0.1 if trial < 8 else 1.5       # large angles reach the other quaternion branches
        q = ref_q + scale * torch.randn(16, generator=g, dtype=torch.float64)
        q32 = q.float().unsqueeze(0).requires_grad_(True)
        tips = o.hand.tips_local(q32, o.ee_names, o.ee_offsets).view(4, 3)
        Jref = torch.stack([torch.autograd.grad(tips.reshape(-1)[i], q32, retain_graph=True)[0][0] for i in range(12)])
        tl, J = host.fk(q.numpy())
        assert np.abs(tl - tips.detach().double().numpy()).max() < 2e-7
        Jfull = np.zeros((12, 16), dtype=np.float32)
        tree = tree_for(doc, doc["ee_link_name"], doc["ee_link_offset"])
        for f in range(4):
            chain, b = [], tree.tip_body[f]
            while b > 0:
                if tree.dof[b] >= 0:
                    chain.append(tree.dof[b])
                b = tree.parent[b]
            for c, d in enumerate(reversed(chain)):
                Jfull[f * 3:(f + 1) * 3, d] = J[f, :, c]
        assert np.abs(Jfull - Jref.numpy()).max() < 2e-6 * max(1.0, np.abs(Jref.numpy()).max())


def test_svd3_small_singular_values():
    host, _ = _host("leap")
    rng = np.random.default_rng(0)
    for trial in range(50):
        a, b = rng.standard_normal(3), rng.standard_normal(3)
        A = 50.0 * np.outer(a, b) + 1e-6 * rng.random((3, 3))       # rank one + noise, as at the standard init
        if trial % 3 == 0:
            A = rng.standard_normal((3, 3))
        U, S, V = host.svd3(A)
        assert np.abs(U @ np.diag(S) @ V.T - A).max() < 1e-13 * max(1.0, np.abs(A).max())
        assert np.abs(U.T @ U - np.eye(3)).max() < 1e-12 and np.abs(V.T @ V - np.eye(3)).max() < 1e-12
        Sref = np.linalg.svd(A, compute_uv=False)
        assert np.all(np.abs(S - Sref) <= 1e-14 * Sref[0]) and S[0] >= S[1] >= S[2]     # LAPACK is absolute-accurate only


def _gpis_inputs(gp, o, q, k, T, palm, setup):
    """what the points kernel emits and the GPIS kernels return, computed with the oracle"""
    tips = o.forward_kinematics(torch.from_numpy(q), torch.from_numpy(palm)).detach()      # [B,4,3]
    Tt = torch.from_numpy(T)
    out = []
    for b in range(q.shape[0]):
        ext = torch.stack([Tt[b] + torch.tensor([setup.coeff[p][f] for f in range(4)]).view(4, 1) * (tips[b] - Tt[b])
                           for p in range(setup.P) if setup.uniq[p] == len(set(setup.uniq[:p]))])   # [U,4,3]
        X = ext.reshape(-1, 3)
        dist, sd = gp.pred_diag(X)
        gmu, gsd = gp.grads(X)
        X9 = torch.vstack([Tt[b], tips[b], torch.from_numpy(palm[b, :3]).view(1, 3)])
        m9, _ = gp.pred_diag(X9)
        g9, _ = gp.grads(X9)
        out.append((dist.numpy(), sd.numpy(), gmu.numpy(), gsd.numpy(), m9.numpy(), g9.numpy()))
    return out


@pytest.mark.parametrize("tag,hand,state", [("leap_mug", "leap", "mug"), ("allegro_hammer", "allegro", "hammer"),
                                            ("allegro_coffeebottle", "allegro", "coffeebottle")])
@pytest.mark.parametrize("perturb", [False, True])
def test_closure_value_and_gradients(tag, hand, state, perturb, golden_dir):
    g = np.load(os.path.join(golden_dir, f"pregrasp_{tag}.npz"))
    gp = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, f"{state}_state.npz"))
    host, doc = _host(hand)
    mu = g["friction_mu"].item()
    mass, com, seed = float(g["mass"]), g["com"].tolist(), int(g["noise_seed"])
    palm = g["palm0"].copy()
    q, T, k = g["q0"].copy(), g["target0"].copy(), g["compliance0"].copy()
    it = 0
    if perturb:      # a generic iterate: distinct targets (full-rank H), compliance at the clamp, different noise key
        rng = np.random.default_rng(7)
        T += 0.01 * rng.standard_normal(T.shape)
        k = 80.0 + 20.0 * rng.random(k.shape)
        q += 0.02 * rng.standard_normal(q.shape)
        palm[:, 3:] += 0.05 * rng.standard_normal((palm.shape[0], 3))
        it = 5
    o = PregraspOracle(os.path.join(HANDS, f"{hand}.json"), palm, 1, mass=mass, com=com, noise_seed=seed)
    cos_mu = float(torch.sqrt(1 / (1 + torch.tensor(mu) ** 2)))
    setup = hc.make_setup(((0.8,) * 4,) * 3, (0.1, 0.8, 0.1), doc["ref_q"], 16, True, cos_mu, com, 10.0 * mass / 2.0,
                          True, 20.0)
    assert setup.U == 1
    total, margin, grads = o.value_and_grads(torch.from_numpy(q), torch.from_numpy(k), torch.from_numpy(T),
                                             torch.from_numpy(palm[:, :3]), torch.from_numpy(palm[:, 3:]), mu, gp, it=it)
    if not perturb:
        assert np.allclose(np.nan_to_num(total.numpy()), np.nan_to_num(g["losses"][0]), rtol=1e-12)
    gin = _gpis_inputs(gp, o, q, k, T, palm, setup)
    B = q.shape[0]
    tl_ref = o.hand.tips_local(torch.from_numpy(q).float(), o.ee_names, o.ee_offsets).double().view(-1, 4, 3).numpy()
    scale = [max(1e-30, float(np.nanmax(np.abs(x.numpy())))) for x in grads]
    for b in range(B):
        tot, tm, gq, gk, gT, gpalm = host.closure(setup, q[b], k[b], T[b], palm[b], *gin[b], seed, it, b, tl=tl_ref[b])
        ref = total[b].item()
        assert np.isnan(tot) == np.isnan(ref)
        if not np.isnan(ref):
            assert abs(tot - ref) <= 1e-9 * max(100.0, abs(ref)), (b, tot, ref)     # terms are O(1e2..1e3) and cancel
        assert np.abs(tm - margin[b].numpy()).max() < 1e-6
        for got, want, sc, nm in [(gq, grads[0][b], scale[0], "q"), (gk, grads[1][b], scale[1], "k"),
                                  (gT, grads[2][b], scale[2], "T"), (gpalm[:3], grads[3][b], scale[3], "palm_pos"),
                                  (gpalm[3:], grads[4][b], scale[4], "palm_ori")]:
            want = want.numpy()
            assert np.array_equal(np.isnan(got), np.isnan(want)), (nm, b)
            err = np.nanmax(np.abs(got - want)) if not np.all(np.isnan(want)) else 0.0
            # the q gradient passes through the float32 FK graph: float32 rounding of O(1e4) values
            assert err <= (2e-5 if nm == "q" else 2e-6) * sc, (nm, b, err, sc)


@pytest.mark.parametrize("tag,hand,state", [("kin_leap_mug", "leap", "mug"), ("kin_allegro_hammer", "allegro", "hammer")])
def test_kin_variant_closure_against_reference_golden(tag, hand, state, golden_dir):
    """the same per-candidate math with KinGPISGraspOptimizer's weights (optimize_pregrasp.py:468-489) against
    the loss and leaf gradients the REFERENCE's autograd produced at the initial iterate (tests/golden/variant_*.npz)."""
    g = np.load(os.path.join(golden_dir, f"variant_{tag}.npz"))
    gp = GPISOracle(0.08, 1.0).load_npz(os.path.join(STATES, f"{state}_state.npz"))
    host, doc = _host(hand)
    mu, mass, com, seed = g["friction_mu"].item(), float(g["mass"]), g["com"].tolist(), int(g["noise_seed"])
    q, T, k = g["q0"], g["target0"], g["compliance0"]
    B = q.shape[0]
    palm = np.zeros((B, 6))
    palm[:, :3] = g["palm3"]
    cos_mu = float(torch.sqrt(1 / (1 + torch.tensor(mu) ** 2)))
    setup = hc.make_setup(((1.0,) * 4,), (1.0,), doc["ref_q"], 16, False, cos_mu, com, 10.0 * mass / 2.0, True, 10.0)
    setup.variant, setup.w_ref, setup.w_pre = 1, 20.0, 0.0
    setup.C.w_reward, setup.C.w_cm, setup.C.w_tar, setup.C.force_clamp, setup.C.w_center = 5.0, 0.0, 10.0, 1.0, 10.0
    tips = g["tips0"].reshape(B, 4, 3)
    tl = tips - g["palm3"]                                   # hand-frame tips of the reference's recursive FK
    want_T = g["grad0_target"] if int(g["optimize_target"]) else None
    for b in range(B):
        X = torch.from_numpy(tips[b])
        dist, sd = gp.pred_diag(X)
        gmu, gsd = gp.grads(X)
        X9 = torch.vstack([torch.from_numpy(T[b]), X, torch.zeros(1, 3, dtype=torch.float64)])
        m9, _ = gp.pred_diag(X9)
        g9, _ = gp.grads(X9)
        tot, tm, gq, gk, gT, _ = host.closure(setup, q[b], k[b], T[b], palm[b], dist.numpy(), sd.numpy(), gmu.numpy(),
                                              gsd.numpy(), m9.numpy(), g9.numpy(), seed, 0, b, tl=tl[b])
        ref = g["losses"][0][b]
        assert abs(tot - ref) <= 1e-8 * max(100.0, abs(ref)), (b, tot, ref)
        assert np.abs(tm - g["margins"][0][b]).max() < 1e-6
        for got, want, nm in [(gq, g["grad0_q"][b], "q"), (gk, g["grad0_compliance"][b], "k")] + \
                ([(gT, want_T[b], "T")] if want_T is not None else []):
            sc = max(1e-30, np.abs(g["grad0_" + {"q": "q", "k": "compliance", "T": "target"}[nm]]).max())
            assert np.abs(got - want).max() <= (5e-5 if nm == "q" else 2e-6) * sc, (nm, b, np.abs(got - want).max(), sc)
