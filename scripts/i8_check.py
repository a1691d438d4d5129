"""GPU check of the tcgen05 INT8 std path against the FP64 DMMA path (same state, same queries)."""
import sys, time, json
import torch
sys.path.insert(0, ".")
from compliancedex_b200 import GPIS
from compliancedex_b200.synthetic import synthetic_training_set, synthetic_queries

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
M = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
X1, y1, noise = synthetic_training_set(N)
g = GPIS(0.08, 1.0)
g.fit(X1.cuda(), y1.cuda(), noise.cuda())
X = synthetic_queries(M, 0, M).cuda()
X[: M // 8] = X1[torch.randint(0, N, (M // 8,))].cuda() + 1e-4 * torch.randn(M // 8, 3, dtype=torch.float64, device="cuda")
prior = float(g.R) ** 3
out = {"N": N, "M": M}
g.set_var_mode(-1)
m0, s0 = g.pred(X)
torch.cuda.synchronize()
for S in (4, 5, 6):
    g.set_var_mode(S)
    m1, s1 = g.pred(X)
    torch.cuda.synchronize()
    err = ((s1 ** 2 - s0 ** 2).abs().max() / prior).item()
    merr = ((m1 - m0).abs().max() / m0.abs().max()).item()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(3):
        g.pred(X)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / 3
    out[f"S{S}"] = {"var_err_over_prior": err, "mean_rel": merr, "ms": ms, "qps": M / ms * 1e3,
                    "nan": int(torch.isnan(s1).sum())}
g.set_var_mode(-1)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for _ in range(3):
    g.pred(X)
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / 3
out["dmma"] = {"ms": ms, "qps": M / ms * 1e3}
print(json.dumps(out))
