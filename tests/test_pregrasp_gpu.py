"""GPU parity: fused pregrasp path (C ABI / drop-in ProbabilisticGraspOptimizer) vs the reference-generated golden
vectors (tests/golden/pregrasp_*.npz, made by oracle/make_golden.py from the reference's own classes) and vs the
oracle at other sizes.  Gates (BASELINE.json): objective <= 1e-4 rel per candidate, identical arg-min index."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [("leap_mug", "leap", "mug"), ("allegro_hammer", "allegro", "hammer"),
         ("allegro_coffeebottle", "allegro", "coffeebottle")]


def _setup(tag, hand, state, golden_dir, num_iters=None, B=None):
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path, FINGERTIP_LB, FINGERTIP_UB
    g = np.load(os.path.join(golden_dir, f"pregrasp_{tag}.npz"))
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path(f"{state}_state"))
    doc = load_hand(hand)
    sl = slice(0, B)
    opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=g["palm0"][sl], num_iters=int(g["iters"]) if num_iters is None else num_iters,
                                      optimize_target=True, ref_q=doc["ref_q"], tip_bounding_box=[FINGERTIP_LB, FINGERTIP_UB],
                                      mass=float(g["mass"]), com=g["com"].tolist(), gravity=True, uncertainty=20.0,
                                      optimize_palm=True, noise_seed=int(g["noise_seed"]))
    q0, T0, k0 = (torch.from_numpy(g[k][sl]).cuda() for k in ("q0", "target0", "compliance0"))
    return g, gp, opt, q0, T0, k0, g["friction_mu"].item()


@pytest.mark.parametrize("tag,hand,state", CASES)
def test_forward_kinematics(tag, hand, state, golden_dir):
    g, gp, opt, q0, T0, k0, mu = _setup(tag, hand, state, golden_dir)
    tips = opt.forward_kinematics(q0, torch.from_numpy(g["palm0"]).cuda())
    assert tips.dtype == torch.float64 and tuple(tips.shape) == (q0.shape[0], 4, 3)
    assert np.abs(tips.cpu().numpy() - g["tips0"]).max() < 5e-7          # float32 FK, different op order


@pytest.mark.parametrize("tag,hand,state", CASES)
def test_first_three_closures_match_reference(tag, hand, state, golden_dir):
    """total_loss and every leaf gradient of iterations 0..2 against the reference's autograd."""
    g, gp, opt, q0, T0, k0, mu = _setup(tag, hand, state, golden_dir)
    opt._init(q0, T0, k0, mu, gp)
    for it in range(3):
        loss, margin, gq, gk, gT, gp6 = [t.cpu().numpy() for t in opt.evaluate()]
        ref = g["losses"][it]
        assert np.array_equal(np.isnan(loss), np.isnan(ref))
        ok = ~np.isnan(ref)
        assert (np.abs(loss[ok] - ref[ok]) <= 1e-4 * np.abs(ref[ok])).all(), (it, loss, ref)
        for got, nm, tol in [(gq, "q", 2e-4), (gk, "compliance", 2e-5), (gT, "target", 2e-5), (gp6[:, :3], "palm_pos", 2e-5),
                             (gp6[:, 3:], "palm_ori", 2e-5)]:
            want = g[f"grad{it}_{nm}"]
            assert np.array_equal(np.isnan(got), np.isnan(want)), (it, nm)
            assert np.nanmax(np.abs(got - want)) <= tol * np.nanmax(np.abs(want)), (it, nm)
        opt.step(1)


@pytest.mark.parametrize("tag,hand,state", CASES)
def test_full_trajectory_matches_reference(tag, hand, state, golden_dir):
    g, gp, opt, q0, T0, k0, mu = _setup(tag, hand, state, golden_dir)
    iters = int(g["iters"])
    opt._init(q0, T0, k0, mu, gp)
    losses = []
    for it in range(iters):
        losses.append(opt.evaluate()[0].cpu().numpy())
        opt.step(1)
    losses = np.stack(losses)
    ref = g["losses"]
    assert np.array_equal(np.isnan(losses), np.isnan(ref))
    rel = np.nan_to_num(np.abs(losses - ref) / np.maximum(np.abs(ref), 1.0))
    # every single evaluation is far inside the 1e-4 gate (test_first_three_closures...); along an Adam trajectory the
    # float32 FK rounding differences (<= 3e-8 m in the tips) are amplified by the optimisation dynamics, like the
    # oracle's own drift against the reference: 1e-4 over the first 20 iterations, 1e-3 to the end
    assert rel[:20].max() <= 1e-4, rel.max(axis=1)
    assert rel.max() <= 1e-3, rel.max(axis=1)
    # the drop-in call: same 5-tuple as optimize_pregrasp.py:839
    res = opt.optimize(q0, T0, k0, mu, gp, verbose=False)
    for a, nm, tol in zip(res, ["opt_q", "opt_compliance", "opt_target", "opt_palm", "opt_margin"], [1e-4, 1e-3, 1e-4, 1e-4, 1e-4]):
        want = g[nm]
        assert tuple(a.shape) == want.shape and a.dtype == torch.float64
        assert np.array_equal(np.isnan(a.cpu().numpy()), np.isnan(want))
        assert np.nanmax(np.abs(a.cpu().numpy() - want)) <= tol * max(1.0, np.nanmax(np.abs(want))), nm
    # arg-min over opt_value (BASELINE.json: identical arg-min candidate index)
    best_ref = np.where(np.isnan(ref[21:]), np.inf, ref[21:]).min(axis=0)
    v, idx = opt.argmin()
    assert idx == int(np.argmin(best_ref)) and abs(v - best_ref.min()) <= 1e-4 * abs(best_ref.min())


def test_sharding_invariance(golden_dir):
    """candidates [4:8] run alone with env_offset=4 reproduce rows 4:8 of the full batch bit for bit."""
    g, gp, opt, q0, T0, k0, mu = _setup(*CASES[0], golden_dir, num_iters=25)
    full = [t.clone() for t in opt.optimize(q0, T0, k0, mu, gp, verbose=False)]
    full_v = opt.opt_value.clone()
    from compliancedex_b200 import ProbabilisticGraspOptimizer, load_hand
    doc = load_hand("leap")
    part = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                       palm_offset=g["palm0"][4:8], num_iters=25, optimize_target=True, ref_q=doc["ref_q"],
                                       mass=float(g["mass"]), com=g["com"].tolist(), optimize_palm=True,
                                       noise_seed=int(g["noise_seed"]), env_offset=4)
    res = part.optimize(q0[4:8], T0[4:8], k0[4:8], mu, gp, verbose=False)
    for a, b in zip(res, full):
        assert torch.equal(a, b[4:8])
    assert torch.equal(part.opt_value, full_v[4:8])
    v, idx = part.argmin()
    assert 4 <= idx < 8


@pytest.mark.parametrize("groups", [2, 3, 4])
def test_candidate_groups_are_bit_identical(groups):
    """cdx_pregrasp_run pipelines the candidates in groups on separate streams (default 2 from 768 candidates): whatever the
    group count, every candidate's trajectory is the one it has in a single group."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("mug_state"))
    doc = load_hand("leap")
    d = np.load(state_path("mug_state"))
    B = 100   # not a multiple of the group size or of the 8 candidates a CTA holds
    q0, T0, k0, palm = synthetic_candidates(doc, B, surface_center(d["X1"], d["y1"]), seed=7)

    def run(G):
        opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=palm, num_iters=30, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                          com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True,
                                          noise_seed=11)
        opt.set_tuning(groups=G)
        res = [t.clone() for t in opt.optimize(q0.cuda(), T0.cuda(), k0.cuda(), 1.0, gp, verbose=False)]
        return res + [opt.opt_value.clone()]

    one = run(1)
    many = run(groups)
    for a, b in zip(one, many):
        assert torch.equal(a, b)


@pytest.mark.parametrize("B", [101, 2500])
def test_step_kernel_occupancy_variants_are_bit_identical(B):
    """pg_step_kernel is compiled twice (2 resident CTAs per SM / 255 registers for launches of one wave, 3 / 168 registers
    for larger ones; cdx_pregrasp_set_tuning forces one): same arithmetic, so the choice - which depends on the launch size, i.e. on the
    shard - never shows in the results.  B = 101 also has a warp whose second candidate lies past the end of the batch."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("mug_state"))
    doc = load_hand("leap")
    d = np.load(state_path("mug_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, B, surface_center(d["X1"], d["y1"]), seed=9)

    def run(occ):
        opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=palm, num_iters=25, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                          com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True,
                                          noise_seed=5)
        opt.set_tuning(occupancy=occ)
        res = [t.clone() for t in opt.optimize(q0.cuda(), T0.cuda(), k0.cuda(), 1.0, gp, verbose=False)]
        return res + [opt.opt_value.clone()]

    two = run(2)
    three = run(3)
    for a, b in zip(two, three):
        assert torch.equal(a, b)


@pytest.mark.parametrize("B,iters", [(100, 107), (1030, 53)])
def test_graph_loop_is_bit_identical_to_direct_launches(B, iters):
    """cdx_pregrasp_run replays the loop as a CUDA graph of 50-iteration chunks (+ directly launched remainder); the captured
    nodes read the chunk's base iteration from device memory.  Same results as launching every iteration from the host,
    also when the cached graph is re-used for a second run with other inputs, and after a run of a different length."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("mug_state"))
    doc = load_hand("leap")
    d = np.load(state_path("mug_state"))
    center = surface_center(d["X1"], d["y1"])

    def make(seed):
        q0, T0, k0, palm = synthetic_candidates(doc, B, center, seed=seed)
        opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=palm, num_iters=iters, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                          com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True, noise_seed=3)
        return opt, q0.cuda(), T0.cuda(), k0.cuda()

    def run(opt, q0, T0, k0, graph):
        opt.set_tuning(use_graph=graph)
        res = [t.clone() for t in opt.optimize(q0, T0, k0, 1.0, gp, verbose=False)]
        return res + [opt.opt_value.clone()] + [t.clone() for t in opt.current()]

    for seed in (21, 22):                       # the second seed re-uses the executable graph captured for the first
        opt, q0, T0, k0 = make(seed)
        direct = run(opt, q0, T0, k0, False)
        graphed = run(opt, q0, T0, k0, True)
        for a, b in zip(direct, graphed):
            assert torch.equal(a, b)
    # split runs: 60 iterations (graph chunk + remainder), then the rest, continue the same trajectory
    opt, q0, T0, k0 = make(22)
    opt.set_tuning(use_graph=True)
    opt._init(q0, T0, k0, 1.0, gp)
    opt.step(min(60, iters))
    opt.step(iters - min(60, iters))
    for a, b in zip(opt.results() + opt.current(), graphed[:5] + [graphed[5]] + graphed[6:]):
        assert torch.equal(a, b)


def test_two_hands_share_one_gpis_context():
    """Optimizers with different hands (Leap, Allegro: both 16 dof) on ONE GPIS object alternate: each run uses its own
    kinematics (the context keeps the loaded hand's identity), and an optimizer whose state was taken over refuses to read
    it instead of returning the other batch."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    d = np.load(state_path("mug_state"))
    center = surface_center(d["X1"], d["y1"])

    def make(hand, B, gp_seed):
        doc = load_hand(hand)
        q0, T0, k0, palm = synthetic_candidates(doc, B, center, seed=gp_seed)
        opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=palm, num_iters=12, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                          optimize_palm=True, noise_seed=1)
        return opt, q0.cuda(), T0.cuda(), k0.cuda()

    shared = GPIS(0.08, 1.0)
    shared.load_state_data(state_path("mug_state"))
    a, qa, Ta, ka = make("leap", 24, 1)
    b, qb, Tb, kb = make("allegro", 40, 2)
    ra1 = [t.clone() for t in a.optimize(qa, Ta, ka, 1.0, shared, verbose=False)]
    rb = [t.clone() for t in b.optimize(qb, Tb, kb, 1.0, shared, verbose=False)]
    with pytest.raises(RuntimeError):
        a.results()                                      # b owns the context's pregrasp state now (and it holds 40 candidates)
    ra2 = a.optimize(qa, Ta, ka, 1.0, shared, verbose=False)
    for x, y in zip(ra1, ra2):
        assert torch.equal(x, y)                         # not Allegro's kinematics
    own = GPIS(0.08, 1.0)
    own.load_state_data(state_path("mug_state"))
    b2, _, _, _ = make("allegro", 40, 2)
    for x, y in zip(rb, b2.optimize(qb, Tb, kb, 1.0, own, verbose=False)):
        assert torch.equal(x, y)
    # the C ABI itself rejects buffers sized for another batch
    import ctypes
    from compliancedex_b200 import _lib
    buf = torch.empty(24, dtype=torch.float64, device="cuda")
    rc = _lib.load().cdx_pregrasp_results(own._ctx, 24, None, None, None, None, None, ctypes.c_void_p(buf.data_ptr()), None)
    assert rc == _lib.CDX_ERR_ARG


def test_nan_record():
    """the reference prints "NaN detected" and continues (optimize_pregrasp.py:819); here the step kernel counts NaN
    candidate-iterations on the device.  One candidate starts from a NaN joint angle: it stays NaN for every iteration, the
    others are untouched (bit-identical to a run without it)."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("mug_state"))
    doc = load_hand("leap")
    d = np.load(state_path("mug_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, 20, surface_center(d["X1"], d["y1"]), seed=4)
    opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm, num_iters=30, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                      optimize_palm=True, noise_seed=8)
    clean = [t.clone() for t in opt.optimize(q0.cuda(), T0.cuda(), k0.cuda(), 1.0, gp, verbose=False)]
    assert opt.nan_count() == (0, -1)
    qn = q0.clone()
    qn[7, 3] = float("nan")
    res = opt.optimize(qn.cuda(), T0.cuda(), k0.cuda(), 1.0, gp, verbose=False)
    assert opt.nan_count() == (30, 0)
    keep = [i for i in range(20) if i != 7]
    for x, y in zip(clean, res):
        assert torch.equal(x[keep], y[keep])
    assert torch.isinf(opt.opt_value[7])                 # a NaN loss never becomes the best iterate (:822-829)
    assert opt.argmin()[1] != 7


def test_distinct_pregrasp_coefficients_against_oracle(golden_dir):
    """coefficient rows that differ (U = 3 GPIS query blocks) + optimize_palm/target off, vs the oracle."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path, hand_path
    from oracle.gpis_oracle import GPISOracle
    from oracle.pregrasp_oracle import PregraspOracle
    g = np.load(os.path.join(golden_dir, "pregrasp_leap_mug.npz"))
    coeffs = [[0.7, 0.75, 0.8, 0.85], [0.8, 0.8, 0.8, 0.8], [0.9, 0.85, 0.8, 0.75]]
    doc = load_hand("leap")
    for opt_palm, opt_target in [(True, True), (False, False)]:
        gp = GPIS(0.08, 1.0)
        gp.load_state_data(state_path("mug_state"))
        opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=g["palm0"], num_iters=6, optimize_target=opt_target, ref_q=doc["ref_q"],
                                          pregrasp_coefficients=coeffs, mass=0.1, optimize_palm=opt_palm, noise_seed=77)
        q0, T0, k0 = (torch.from_numpy(g[k]) for k in ("q0", "target0", "compliance0"))
        T0 = T0 + 0.01 * torch.randn(T0.shape, generator=torch.Generator().manual_seed(1), dtype=torch.float64)
        o = PregraspOracle(hand_path("leap"), g["palm0"], 6, optimize_target=opt_target, optimize_palm=opt_palm,
                           pregrasp_coefficients=coeffs, mass=0.1, noise_seed=77)
        go = GPISOracle(0.08, 1.0).load_npz(state_path("mug_state"))
        ref_losses = []
        o.optimize(q0.clone(), T0.clone(), k0.clone(), 1, go, record=lambda s, t, gr: ref_losses.append(t.numpy()))
        opt._init(q0.cuda(), T0.cuda(), k0.cuda(), 1, gp)
        for it in range(6):
            loss = opt.evaluate()[0].cpu().numpy()
            ref = ref_losses[it]
            assert np.array_equal(np.isnan(loss), np.isnan(ref))
            ok = ~np.isnan(ref)
            # the loss is a sum of cancelling terms of magnitude 1e2..1e3 (1000 |dist|, 200 log barriers): floor the scale at 10
            assert (np.abs(loss[ok] - ref[ok]) <= 1e-4 * np.maximum(np.abs(ref[ok]), 10.0)).all(), (it, loss, ref)
            opt.step(1)


def test_errors():
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand
    doc = load_hand("leap")
    opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=np.zeros((2, 6)), num_iters=1, ref_q=doc["ref_q"])
    with pytest.raises(RuntimeError):
        opt.optimize(torch.zeros(2, 16), torch.zeros(2, 4, 3), torch.ones(2, 4), 1, GPIS(0.08, 1.0))   # unfitted GPIS


@pytest.mark.parametrize("hand", ["leap", "allegro"])
def test_collision_loss_against_reference_golden(hand, golden_dir):
    """compute_collision_loss (optimize_pregrasp.py:671-701; dormant on the prob path, provided for API parity): CUDA FK of
    the anchor links + barrier kernel vs the reference's own outputs.  float32 FK: libdevice sin/cos differ from torch
    CPU's in the last bit, amplified by 1/z near the floor -> 1e-4 relative."""
    from compliancedex_b200 import ProbabilisticGraspOptimizer, load_hand
    ref = np.load(os.path.join(golden_dir, "pregrasp_collision.npz"))
    doc = load_hand(hand)
    q, palm = torch.from_numpy(ref[f"{hand}_q"]), torch.from_numpy(ref[f"{hand}_palm"])
    for opt_palm, thr, key in ((True, 0.02, "cost_palm1"), (False, 0.02, "cost_palm0"), (True, 0.08, "cost_thr008")):
        opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                          palm_offset=palm.numpy(), num_iters=1, ref_q=doc["ref_q"], optimize_palm=opt_palm,
                                          anchor_link_names=doc["collision_links"], anchor_link_offsets=doc["collision_offsets"],
                                          collision_pairs=doc["collision_pairs"], collision_pair_threshold=thr)
        got = opt.compute_collision_loss(q.cuda(), palm.cuda()).cpu().numpy()
        want = ref[f"{hand}_{key}"]
        assert got.shape == want.shape
        assert np.allclose(got, want, rtol=1e-4, atol=1e-7), np.abs(got - want).max()
        got2 = opt.compute_collision_loss(q.cuda()).cpu().numpy()            # default palm_poses = palm_offset (:672-673)
        assert np.array_equal(got, got2)
