import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path, hand_path
from oracle.gpis_oracle import GPISOracle
from oracle.pregrasp_oracle import PregraspOracle
g = np.load("tests/golden/pregrasp_leap_mug.npz")
coeffs = [[0.7, 0.75, 0.8, 0.85], [0.8, 0.8, 0.8, 0.8], [0.9, 0.85, 0.8, 0.75]]
doc = load_hand("leap")
order = [(False, False), (True, True)] if len(sys.argv) > 1 else [(True, True), (False, False)]
for opt_palm, opt_target in order:
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("mug_state"))
    opt = ProbabilisticGraspOptimizer("leap", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=g["palm0"], num_iters=6, optimize_target=opt_target, ref_q=doc["ref_q"],
                                      pregrasp_coefficients=coeffs, mass=0.1, optimize_palm=opt_palm, noise_seed=77)
    q0, T0, k0 = (torch.from_numpy(g[k]) for k in ("q0", "target0", "compliance0"))
    T0 = T0 + 0.01 * torch.randn(T0.shape, generator=torch.Generator().manual_seed(1), dtype=torch.float64)
    o = PregraspOracle(hand_path("leap"), g["palm0"], 2, optimize_target=opt_target, optimize_palm=opt_palm,
                       pregrasp_coefficients=coeffs, mass=0.1, noise_seed=77)
    go = GPISOracle(0.08, 1.0).load_npz(state_path("mug_state"))
    ref = []
    o.optimize(q0, T0, k0, 1, go, record=lambda s, t, gr: ref.append(t.numpy()))
    opt._init(q0.cuda(), T0.cuda(), k0.cuda(), 1, gp)
    for it in range(2):
        loss = opt.evaluate()[0].cpu().numpy()
        print(opt_palm, opt_target, it, "gpu", loss[:4], "ref", ref[it][:4])
        opt.step(1)
