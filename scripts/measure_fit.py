"""Times GPIS.fit (assembly + blocked Cholesky + L^-1 + E11^-1 + alpha) at the C4 shape and reports FP64 TF/s against
N^3/3 (factorisation only) and N^3 (factor + both inverses).  python scripts/measure_fit.py > gpurun_out/fit.json"""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from compliancedex_b200 import GPIS
from compliancedex_b200.synthetic import synthetic_training_set

out = []
for N in (4096, 8192, 16384):
    X1, y1, noise = synthetic_training_set(N, seed=0)
    X1, y1, noise = X1.cuda(), y1.cuda(), noise.cuda()
    g = GPIS(0.08, 1.0)
    g.fit(X1, y1, noise)                       # warm-up (allocations, attribute calls)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.fit(X1, y1, noise)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    t0 = time.perf_counter()
    Lc = torch.linalg.cholesky(g.E11.cpu()) if N <= 8192 else None      # host LAPACK for scale (all cores)
    cpu_s = time.perf_counter() - t0 if Lc is not None else None
    out.append({"N": N, "fit_ms_best_of_3": ms, "tflops_cholesky_only_N3_over_3": N ** 3 / 3 / (ms * 1e-3) / 1e12,
                "tflops_all_three_N3": N ** 3 / (ms * 1e-3) / 1e12, "cpu_torch_cholesky_s": cpu_s,
                "note": "fit = assembly + Cholesky + L^-1 + E11^-1 + alpha + E11 read-back copy"})
    del g
    torch.cuda.empty_cache()
print(json.dumps(out))
