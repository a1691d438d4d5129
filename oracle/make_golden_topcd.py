"""TEST TOOLING (build container only): golden vector for GPIS.topcd (gpis.py:127-150) produced by the REFERENCE's own
triple loop on a 18^3 visualisation grid of the banana state.  Run:  python -m oracle.make_golden_topcd"""
import os

import numpy as np
import torch

from oracle import ref_harness as rh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    rh.install()
    g = rh.load_gpis("banana_state")
    steps = 18
    lb, ub = [-0.15] * 3, [0.15] * 3
    mean, std, nrm, lb_a, ub_a = g.get_visualization_data(lb, ub, steps=steps)
    pts, nrms = g.topcd(mean, nrm, lb, ub, steps=steps)                              # numpy inputs
    pts_t, nrms_t = g.topcd(torch.from_numpy(mean), torch.from_numpy(nrm), lb, ub, steps=steps)   # tensor inputs (flipped)
    np.savez(os.path.join(GOLD, "gpis_topcd_banana18.npz"), mean=mean, std=std, normal=nrm, lb=lb_a, ub=ub_a,
             points=pts, normals=nrms, points_tensor_in=pts_t, normals_tensor_in=nrms_t)
    print("topcd golden:", pts.shape, pts_t.shape)


if __name__ == "__main__":
    main()
