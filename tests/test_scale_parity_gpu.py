"""GPU parity at the BASELINE config sizes (C2: Leap / mug, 1,024 candidates x 200 iterations; C5 per-GPU shard: Allegro /
hammer, 8,192 of the 65,536 candidates x 500 iterations) against
  * tests/golden/pregrasp_scale_{c2,c5}.npz - the REAL reference's own trajectories for the leading rows of those batches
    (oracle/make_golden_scale.py runs the unmodified optimize_pregrasp.py), and
  * tests/golden/pregrasp_scale_c2_port.npz - the oracle port's result for ALL 1,024 candidates of C2.
What is asserted: identical arg-min candidate index (BASELINE.json's last clause), identical NaN / never-improved pattern,
every candidate's total_loss within 1e-4 for the first 25 iterations, and lower bounds on how many candidates still agree to
1e-4 / 1e-3 / 1e-2 in opt_value after the whole run.  Why not 1e-4 for all of them at the end: the reference evaluates FK in
float32 with torch CPU's SLEEF sin/cos and MKL 3x3 products, whose last bits no other implementation (and no other CPU)
reproduces; a 3e-8 m difference in a fingertip grows by about 2e4 over 200 Adam iterations (measured: median relative
difference 4e-8 at iteration 0, 8e-4 at iteration 199 - the oracle port, itself torch on CPU, drifts from the reference by
the same amount).  The measured fractions are recorded in profiles/r02_scale_parity_*.json."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(hand, state, total, rows, iters, record_iters):
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path(f"{state}_state"))
    doc = load_hand(hand)
    d = np.load(state_path(f"{state}_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, total, surface_center(d["X1"], d["y1"]), seed=100)
    q0, T0, k0, palm = q0[:rows].cuda(), T0[:rows].cuda(), k0[:rows].cuda(), palm[:rows]
    opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm, num_iters=iters, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                      com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True, noise_seed=2024)
    opt._init(q0, T0, k0, 1, gp)
    early = []
    for _ in range(record_iters):
        early.append(opt.evaluate()[0].cpu().numpy())
        opt.step(1)
    res = opt.optimize(q0, T0, k0, 1, gp, verbose=False)
    return np.stack(early), opt.opt_value.cpu().numpy(), [t.cpu().numpy() for t in res], opt.argmin()


def _agreement(ov, ref):
    fin = np.isfinite(ref)
    rel = np.abs(ov[fin] - ref[fin]) / np.maximum(np.abs(ref[fin]), 1.0)
    return {t: float((rel <= t).mean()) for t in (1e-4, 1e-3, 1e-2)}, float(np.median(rel))


def _argmin(v):
    return int(np.argmin(np.where(np.isfinite(v), v, np.inf)))


def test_c2_full_size_against_reference_and_port(golden_dir):
    ref = np.load(os.path.join(golden_dir, "pregrasp_scale_c2.npz"))
    port = np.load(os.path.join(golden_dir, "pregrasp_scale_c2_port.npz"))
    early, ov, res, best = _run("leap", "mug", 1024, 1024, 200, 25)
    n = int(ref["take"])
    # --- the reference itself, rows [0, 256)
    want = ref["losses"][:25]
    assert np.array_equal(np.isnan(early[:, :n]), np.isnan(want))
    rel = np.nan_to_num(np.abs(early[:, :n] - want) / np.maximum(np.abs(want), 1.0))
    assert rel.max() <= 1e-4, rel.max(axis=1)
    assert np.array_equal(np.isfinite(ov[:n]), np.isfinite(ref["opt_value"]))
    assert _argmin(ov[:n]) == int(ref["argmin"]) == _argmin(ref["opt_value"])           # identical arg-min index
    frac, med = _agreement(ov[:n], ref["opt_value"])
    assert frac[1e-4] >= 0.40 and frac[1e-3] >= 0.55 and frac[1e-2] >= 0.75, (frac, med)  # measured 0.52 / 0.69 / 0.88
    top_ref = np.argsort(np.where(np.isfinite(ref["opt_value"]), ref["opt_value"], np.inf))[:3]
    top = np.argsort(np.where(np.isfinite(ov[:n]), ov[:n], np.inf))[:3]
    assert top.tolist() == top_ref.tolist()
    assert np.allclose(ov[top], ref["opt_value"][top_ref], rtol=1e-4)                      # the winners agree to 1e-4
    # --- the oracle port, all 1,024 candidates
    assert np.array_equal(np.isfinite(ov), np.isfinite(port["opt_value"]))
    assert best[1] == int(port["argmin"]) == _argmin(ov)
    assert abs(best[0] - float(port["opt_value"][best[1]])) <= 1e-3 * abs(best[0])
    frac_p, med_p = _agreement(ov, port["opt_value"])
    assert frac_p[1e-4] >= 0.40 and frac_p[1e-2] >= 0.75, (frac_p, med_p)
    # the 5-tuple of the winner (optimize_pregrasp.py:839) against the reference's
    w = int(ref["argmin"])
    for a, nm in zip(res, ["opt_q", "opt_compliance", "opt_target", "opt_palm", "opt_margin"]):
        assert np.allclose(a[w], ref[nm][w], rtol=2e-3, atol=2e-4), nm
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump({"vs_reference_256": {"within": frac, "median_rel": med}, "vs_port_1024": {"within": frac_p, "median_rel": med_p},
               "argmin": best}, open(os.path.join(ROOT, "gpurun_out", "test_scale_parity_c2.json"), "w"))


def test_c5_shard_against_reference(golden_dir):
    ref = np.load(os.path.join(golden_dir, "pregrasp_scale_c5.npz"))
    total = int(ref["full"])
    early, ov, res, best = _run("allegro", "hammer", total, 8192, 500, 12)
    n = int(ref["take"])
    want = ref["losses"][:12]
    assert np.array_equal(np.isnan(early[:, :n]), np.isnan(want))
    rel = np.nan_to_num(np.abs(early[:, :n] - want) / np.maximum(np.abs(want), 1.0))
    assert rel.max() <= 1e-4, rel.max(axis=1)
    assert np.array_equal(np.isfinite(ov[:n]), np.isfinite(ref["opt_value"]))
    assert _argmin(ov[:n]) == int(ref["argmin"])                                           # identical arg-min index
    frac, med = _agreement(ov[:n], ref["opt_value"])
    assert frac[1e-2] >= 0.5, (frac, med)                                                  # 500 iterations: measured 0.32 / 0.45 / 0.65
    assert np.isfinite(best[0]) and 0 <= best[1] < 8192 and best[0] <= np.nanmin(np.where(np.isfinite(ov), ov, np.nan))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump({"vs_reference": {"rows": n, "within": frac, "median_rel": med}, "argmin_shard": best},
              open(os.path.join(ROOT, "gpurun_out", "test_scale_parity_c5.json"), "w"))
