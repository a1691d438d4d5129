"""pregrasp iteration time against the number of candidate groups (cdx_pregrasp_set_tuning): C2 shape (Leap / mug, 1,024
candidates) and the C5 per-GPU shape (Allegro / hammer, 8,192).  Results are bit-identical under every setting; the arg-min
value is printed as the check.  `python scripts/pg_groups_time.py`"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, load_hand, state_path  # noqa: E402
from compliancedex_b200.synthetic import surface_center, synthetic_candidates  # noqa: E402

for hand, obj, B, iters in (("leap", "mug", 1024, 200), ("leap", "mug", 2048, 200), ("allegro", "hammer", 8192, 100)):
    doc = load_hand(hand)
    d = np.load(state_path(f"{obj}_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, B, surface_center(d["X1"], d["y1"]), seed=100)
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path(f"{obj}_state"))
    opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm, num_iters=iters, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                      com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True, device="cuda",
                                      noise_seed=2024, env_offset=0)
    q0d, T0d, k0d = q0.cuda(), T0.cuda(), k0.cuda()
    for G in (1, 2, 3, 4):
        opt.set_tuning(groups=G)
        best = None
        for rep in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            opt._init(q0d, T0d, k0d, 1, gp)
            e0.record()
            opt.step(iters)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            if rep and (best is None or ms < best):
                best = ms
        val = float(opt.results()[-1].min()) if hasattr(opt, "results") else float("nan")
        print(f"{hand}/{obj} B={B} groups={G}: {best / iters * 1e3:8.1f} us per iteration  (min opt_value {val:.9e})", flush=True)
