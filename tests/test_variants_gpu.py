"""GPU parity for SURVEY.md 8f-2: the drop-in KinGPISGraspOptimizer / GPISGraspOptimizer (variants 1 / 2 of the fused step
kernel, RMSprop) vs golden vectors produced by the reference's own classes (oracle/make_golden_variants.py ->
tests/golden/variant_*.npz).  Gates: objective <= 1e-4 rel per candidate, identical arg-min index and success flag."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

KIN = [("kin_leap_mug", "leap", "mug"), ("kin_allegro_hammer", "allegro", "hammer")]
GP = [("gpis_lego", "lego"), ("gpis_coffeebottle", "coffeebottle")]


def _gpis(state):
    from compliancedex_b200 import GPIS, state_path
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path(f"{state}_state"))
    return gp


def _kin(tag, hand, state, golden_dir, num_iters=None, **kw):
    from compliancedex_b200 import KinGPISGraspOptimizer, load_hand, FINGERTIP_LB, FINGERTIP_UB
    g = np.load(os.path.join(golden_dir, f"variant_{tag}.npz"))
    doc = load_hand(hand)
    opt = KinGPISGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                palm_offset=g["palm3"].tolist(), num_iters=int(g["iters"]) if num_iters is None else num_iters,
                                optimize_target=bool(int(g["optimize_target"])), ref_q=doc["ref_q"],
                                tip_bounding_box=[FINGERTIP_LB, FINGERTIP_UB], mass=float(g["mass"]), com=g["com"].tolist(),
                                gravity=True, uncertainty=10.0, noise_seed=int(g["noise_seed"]), **kw)
    x0, T0, k0 = (torch.from_numpy(g[k]).cuda() for k in ("q0", "target0", "compliance0"))
    return g, _gpis(state), opt, x0, T0, k0, g["friction_mu"].item()


def _gp(tag, state, golden_dir, num_iters=None, **kw):
    from compliancedex_b200 import GPISGraspOptimizer, FINGERTIP_LB, FINGERTIP_UB
    g = np.load(os.path.join(golden_dir, f"variant_{tag}.npz"))
    opt = GPISGraspOptimizer([FINGERTIP_LB, FINGERTIP_UB], num_iters=int(g["iters"]) if num_iters is None else num_iters,
                             optimize_target=bool(int(g["optimize_target"])), mass=float(g["mass"]), com=g["com"].tolist(),
                             gravity=True, uncertainty=20.0, noise_seed=int(g["noise_seed"]), **kw)
    B = g["q0"].shape[0]
    x0 = torch.from_numpy(g["tips0"]).view(B, 4, 3).cuda()
    T0, k0 = (torch.from_numpy(g[k]).cuda() for k in ("target0", "compliance0"))
    return g, _gpis(state), opt, x0, T0, k0, g["friction_mu"].item()


def _check_first_closures(g, opt, x0, T0, k0, mu, gp, leaf):
    opt._init(x0.reshape(x0.shape[0], -1), T0, k0, mu, gp)
    for it in range(3):
        loss, margin, gx, gk, gT, _ = [t.cpu().numpy() for t in opt.evaluate()]
        ref = g["losses"][it]
        assert (np.abs(loss - ref) <= 1e-4 * np.maximum(np.abs(ref), 1.0)).all(), (it, loss, ref)
        assert np.abs(margin - g["margins"][it]).max() <= 1e-5
        checks = [(gx, leaf, 2e-4 if leaf == "q" else 2e-5), (gk, "compliance", 2e-5)]
        if int(g["optimize_target"]):
            checks.append((gT, "target", 2e-5))
        for got, nm, tol in checks:
            want = g[f"grad{it}_{nm}"].reshape(got.shape)
            assert np.abs(got - want).max() <= tol * np.abs(want).max(), (it, nm, np.abs(got - want).max(), np.abs(want).max())
        opt.step(1)


def _check_trajectory(g, opt, x0, T0, k0, mu, gp):
    iters = int(g["iters"])
    opt._init(x0.reshape(x0.shape[0], -1), T0, k0, mu, gp)
    losses = []
    for it in range(iters):
        losses.append(opt.evaluate()[0].cpu().numpy())
        opt.step(1)
    rel = np.abs(np.stack(losses) - g["losses"]) / np.maximum(np.abs(g["losses"]), 1.0)
    # RMSprop's first steps are +-lr per coordinate whatever the gradient's size, so float32-FK rounding is not amplified
    # the way Adam amplifies it on the prob path; same bounds as there
    assert rel[:20].max() <= 1e-4, rel.max(axis=1)
    assert rel.max() <= 1e-3, rel.max(axis=1)
    res = opt.optimize(x0, T0, k0, mu, gp, verbose=False)        # the drop-in call: same 4-tuple as :511 / :406
    assert len(res) == 4
    for a, nm, tol in zip(res[:3], ["opt_x", "opt_compliance", "opt_target"], [1e-4, 1e-3, 1e-4]):
        want = g[nm]
        assert tuple(a.shape) == want.shape and a.dtype == torch.float64
        assert np.abs(a.cpu().numpy() - want).max() <= tol * max(1.0, np.abs(want).max()), nm
    assert res[3].dtype == torch.bool and res[3].dim() == 0 and bool(res[3]) == bool(g["flag"])
    m, it = opt.latched_margin()
    assert int(it.item()) == int(g["opt_margin_iter"])
    assert np.abs(m.cpu().numpy() - g["opt_margin"]).max() <= 1e-4
    assert np.abs(opt.opt_value.cpu().numpy() - g["opt_value"]).max() <= 1e-3 * np.abs(g["opt_value"]).max()
    v, idx = opt.argmin()
    assert idx == int(np.argmin(g["opt_value"]))


@pytest.mark.parametrize("tag,hand,state", KIN)
def test_kin_recursive_fk(tag, hand, state, golden_dir):
    g, gp, opt, q0, T0, k0, mu = _kin(tag, hand, state, golden_dir)
    tips = opt.forward_kinematics(q0)
    assert tips.dtype == torch.float64 and tuple(tips.shape) == (q0.shape[0] * 4, 3)
    assert np.abs(tips.cpu().numpy() - g["tips0"]).max() < 5e-7


@pytest.mark.parametrize("tag,hand,state", KIN)
def test_kin_first_three_closures(tag, hand, state, golden_dir):
    g, gp, opt, q0, T0, k0, mu = _kin(tag, hand, state, golden_dir)
    _check_first_closures(g, opt, q0, T0, k0, mu, gp, "q")


@pytest.mark.parametrize("tag,hand,state", KIN)
def test_kin_trajectory_and_return(tag, hand, state, golden_dir):
    g, gp, opt, q0, T0, k0, mu = _kin(tag, hand, state, golden_dir)
    _check_trajectory(g, opt, q0, T0, k0, mu, gp)


@pytest.mark.parametrize("tag,state", GP)
def test_gpis_optimizer_first_three_closures(tag, state, golden_dir):
    g, gp, opt, x0, T0, k0, mu = _gp(tag, state, golden_dir)
    _check_first_closures(g, opt, x0, T0, k0, mu, gp, "tip")


@pytest.mark.parametrize("tag,state", GP)
def test_gpis_optimizer_trajectory_and_return(tag, state, golden_dir):
    g, gp, opt, x0, T0, k0, mu = _gp(tag, state, golden_dir)
    _check_trajectory(g, opt, x0, T0, k0, mu, gp)


def test_latch_is_the_last_improving_iteration(golden_dir):
    """opt_margin is the WHOLE batch's margin at the last iteration in which ANY candidate improved (:491-494): replay the
    bookkeeping on the host from the per-iteration losses and compare iteration index, margins and opt_value bit for bit."""
    g, gp, opt, q0, T0, k0, mu = _kin(*KIN[0], golden_dir, num_iters=1)
    opt._init(q0, T0, k0, mu, gp)
    best = np.full(q0.shape[0], np.inf)
    last, margins = -1, []
    for it in range(40):
        loss, margin = [t.cpu().numpy() for t in opt.evaluate()[:2]]
        margins.append(margin)
        upd = loss < best
        if upd.any():
            last = it
            best[upd] = loss[upd]
        opt.step(1)
    m, it = opt.latched_margin()
    assert int(it.item()) == last
    assert np.array_equal(m.cpu().numpy(), margins[last])
    oq, ok, oT, _, _, ov = opt.results()
    assert np.array_equal(ov.cpu().numpy(), best)


def test_variant_sharding_invariance(golden_dir):
    g, gp, opt, q0, T0, k0, mu = _kin(*KIN[0], golden_dir, num_iters=15)
    full = opt.optimize(q0, T0, k0, mu, gp, verbose=False)
    full = [t.clone() for t in full[:3]]
    _, _, part, _, _, _, _ = _kin(*KIN[0], golden_dir, num_iters=15, env_offset=3)
    res = part.optimize(q0[3:], T0[3:], k0[3:], mu, gp, verbose=False)
    for a, b in zip(res[:3], full):
        assert torch.equal(a, b[3:])


def test_variant_errors():
    from compliancedex_b200 import KinGPISGraspOptimizer, load_hand, optimizers
    doc = load_hand("leap")
    with pytest.raises(ValueError):
        KinGPISGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], palm_offset=np.zeros((6, 6)), ref_q=doc["ref_q"])
    assert set(optimizers) == {"prob", "kingpis", "gpis"}
