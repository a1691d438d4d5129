"""TEST TOOLING (build container only): config-scale pregrasp fixtures from the REFERENCE's own code.

Runs the unmodified `ProbabilisticGraspOptimizer.optimize` (optimize_pregrasp.py:771-839, imported from /root/reference
through oracle/ref_harness.py) on the FIRST `take` candidates of the batches bench.py times:
  * C2: Leap on the mug state, the 1,024-candidate batch of rank 0 (seed 100, noise seed 2024), 200 iterations;
  * C5: Allegro on the hammer state, the 65,536-candidate batch that bench.py splits over the ranks (seed 100, noise seed
    2024), 500 iterations.
Candidates are independent and the Kabsch noise is keyed by the global candidate index, so rows [0, take) of the full
batch are reproduced exactly by running those rows alone.  Stored: per-iteration total_loss [iters, take], opt_value
(min over s > 20 of total_loss, the reference's internal `opt_value`, :802,:822-829), its arg-min, the 5-tuple.

    python -m oracle.make_golden_scale [c2|c5] [take]
"""
import contextlib
import io
import json
import os
import sys
import time

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.make_golden import GOLD, HANDS, STATES, surface_center
from oracle.pregrasp_oracle import synthetic_candidates

CASES = {"c2": dict(hand="leap", state="mug", full=1024, iters=200, take=256),
         "c5": dict(hand="allegro", state="hammer", full=65536, iters=500, take=128)}


def opt_value_from_losses(losses, track_after=20):
    """the reference's best-iterate rule (:822-829): strict improvement, only for s > 20, NaN never updates"""
    ov = np.full(losses.shape[1], np.inf)
    for s in range(losses.shape[0]):
        if s > track_after:
            upd = losses[s] < ov
            ov = np.where(upd, losses[s], ov)
    return ov


def run(case, take=None):
    c = CASES[case]
    take = take or c["take"]
    g = rh.load_gpis(f"{c['state']}_state")
    doc = json.load(open(os.path.join(HANDS, f"{c['hand']}.json")))
    d = np.load(os.path.join(STATES, f"{c['state']}_state.npz"))
    center = surface_center(d["X1"], d["y1"])
    q, target, comp, palm = synthetic_candidates(doc, c["full"], center, seed=100)
    q, target, comp, palm = q[:take].clone(), target[:take].clone(), comp[:take].clone(), palm[:take].copy()
    ropt, _ = rh.make_prob_optimizer(c["hand"], palm, c["iters"], mass=0.1, com=[0.0, 0.0, 0.0])
    losses = []
    t0 = time.time()
    with contextlib.redirect_stdout(io.StringIO()), rh.pinned_kabsch_noise(2024):
        orig = ropt.closure

        def closure(*a, **k):
            r = orig(*a, **k)
            losses.append(ropt.total_loss.detach().clone().numpy())
            return r

        ropt.closure = closure
        res = ropt.optimize(q, target, comp, 1, g)
    losses = np.stack(losses)
    ov = opt_value_from_losses(losses)
    out = dict(hand=np.asarray(c["hand"]), state=np.asarray(c["state"]), full=np.asarray(c["full"]), take=np.asarray(take),
               iters=np.asarray(c["iters"]), seed=np.asarray(100), noise_seed=np.asarray(2024), losses=losses, opt_value=ov,
               argmin=np.asarray(int(np.nanargmin(ov))))
    for nm, a in zip(["opt_q", "opt_compliance", "opt_target", "opt_palm", "opt_margin"], res):
        out[nm] = a.detach().numpy()
    path = os.path.join(GOLD, f"pregrasp_scale_{case}.npz")
    np.savez_compressed(path, **out)
    print("scale golden", case, "take", take, "iters", c["iters"], "%.0f s" % (time.time() - t0), "argmin", int(out["argmin"]),
          "opt_value min", float(np.nanmin(ov)), flush=True)


def run_port(case, take=None):
    """the same batch through the ORACLE PORT (oracle/pregrasp_oracle.py) at the full config size, where the reference itself
    is too slow to run (1,024 x 200 takes minutes instead of the better part of an hour): opt_value of every candidate, its
    arg-min and every tenth iteration's total_loss -> tests/golden/pregrasp_scale_<case>_port.npz"""
    from compliancedex_b200 import hand_path, state_path
    from oracle.gpis_oracle import GPISOracle
    from oracle.pregrasp_oracle import PregraspOracle
    c = CASES[case]
    take = take or c["full"]
    doc = json.load(open(os.path.join(HANDS, f"{c['hand']}.json")))
    d = np.load(os.path.join(STATES, f"{c['state']}_state.npz"))
    q, target, comp, palm = synthetic_candidates(doc, c["full"], surface_center(d["X1"], d["y1"]), seed=100)
    go = GPISOracle(0.08, 1.0).load_npz(state_path(f"{c['state']}_state"))
    o = PregraspOracle(hand_path(c["hand"]), palm[:take], c["iters"], mass=0.1, noise_seed=2024)
    rows = []
    t0 = time.time()
    o.optimize(q[:take], target[:take], comp[:take], 1, go, record=lambda s, t, gr: rows.append(t.numpy().copy()))
    losses = np.stack(rows)
    ov = opt_value_from_losses(losses)
    np.savez_compressed(os.path.join(GOLD, f"pregrasp_scale_{case}_port.npz"), take=np.asarray(take), iters=np.asarray(c["iters"]),
                        opt_value=ov, argmin=np.asarray(int(np.nanargmin(ov))), losses_every10=losses[::10])
    print("scale golden (oracle port)", case, "take", take, "%.0f s" % (time.time() - t0), "argmin", int(np.nanargmin(ov)), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "port":
        run_port(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else None)
        sys.exit(0)
    rh.install()
    torch.manual_seed(0)
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    take = int(sys.argv[2]) if len(sys.argv) > 2 else None
    for case in (["c2", "c5"] if which == "all" else [which]):
        run(case, take)
