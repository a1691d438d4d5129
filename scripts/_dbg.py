import os, sys
import numpy as np, torch
sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, state_path
for name in ("coffeebottle", "hammer", "mug", "banana", "lego", "dummy"):
    try:
        ref = np.load(f"tests/golden/gpis_{name}.npz")
    except Exception as e:
        print(name, "no golden", e); continue
    g = GPIS(0.08, 1.0); g.load_state_data(state_path(f"{name}_state"))
    X = torch.from_numpy(ref["X"]).cuda()
    outs = {}
    for mode in (0, 1, 2):
        g.set_ext_mode(mode); outs[mode] = g.pred_with_grads(X)
    print(name, "N", g.X1.shape[0], "prior", float(g.R) ** 3)
    for m in (0, 2):
        print("  mode", m, "vs 1:", ["%.2e / %.2e" % (float((a - b).abs().max()), float(b.abs().max())) for a, b in zip(outs[m], outs[1])])
    print("  std min", float(outs[1][1].min()))
