"""TEST TOOLING (build container only): golden vector for ProbabilisticGraspOptimizer.compute_collision_loss
(optimize_pregrasp.py:671-701) from the REFERENCE's own code, Leap and Allegro, with candidates pushed towards the floor
and with fingertips brought together so that every barrier term is exercised.  Run: python -m oracle.make_golden_collision"""
import json
import os

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.pregrasp_oracle import synthetic_candidates

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
HANDS = os.path.join(ROOT, "compliancedex_b200", "data", "hands")


def main():
    rh.install()
    out = {}
    for hand in ("leap", "allegro"):
        doc = json.load(open(os.path.join(HANDS, f"{hand}.json")))
        q, _, _, palm = synthetic_candidates(doc, 24, [0.0, 0.0, 0.03], seed=7)
        g = torch.Generator().manual_seed(11)
        q = q + 0.6 * torch.randn(q.shape, generator=g, dtype=torch.float64)        # curl fingers: tips come close
        palm = torch.from_numpy(palm).clone()
        palm[:, 2] = torch.linspace(-0.03, 0.12, 24, dtype=torch.float64)           # below / at / above the floor
        for opt_palm in (True, False):
            ropt, _ = rh.make_prob_optimizer(hand, palm.numpy(), 1, optimize_palm=opt_palm)
            with torch.no_grad():
                cost = ropt.compute_collision_loss(q, palm)
            out[f"{hand}_cost_palm{int(opt_palm)}"] = cost.numpy()
        # a larger threshold so that the pair term is active for most candidates
        ropt, _ = rh.make_prob_optimizer(hand, palm.numpy(), 1, optimize_palm=True, collision_pair_threshold=0.08)
        with torch.no_grad():
            out[f"{hand}_cost_thr008"] = ropt.compute_collision_loss(q, palm).numpy()
        out[f"{hand}_q"] = q.numpy()
        out[f"{hand}_palm"] = palm.numpy()
        print(hand, out[f"{hand}_cost_palm1"][:6], out[f"{hand}_cost_thr008"][:6])
    np.savez(os.path.join(GOLD, "pregrasp_collision.npz"), **out)


if __name__ == "__main__":
    main()
