"""Quick GPU check of the INT8 std variants (cdx_gpis_set_i8_variant 0 / 1 / 2) against the FP64 tensor-MMA path:
max |d var| / prior, chunk invariance, and timing at N = 4096.  Usage: python scripts/i8x_check.py [variants...]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, state_path  # noqa: E402
from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set  # noqa: E402

variants = [int(v) for v in sys.argv[1:]] or [2, 1, 0]


def check(g, X, tag, variants):
    prior = float(g.R) ** 3
    with torch.no_grad():
        g.set_var_mode(-1)
        _, s0 = g.pred(X)
        for v in variants:
            for S in (5, 4, 6):
                g.set_i8_variant(v)
                g.set_var_mode(S)
                _, s1 = g.pred(X)
                torch.cuda.synchronize()
                err = ((s1 ** 2 - s0 ** 2).abs().max() / prior).item()
                _, s2 = g.pred(X[37:37 + 500])
                same = torch.equal(s2, s1[37:37 + 500])
                print(f"{tag} variant {v} S={S}: max|dvar|/prior = {err:.3e} nan={int(torch.isnan(s1).sum())} chunk-invariant={same}", flush=True)


g = GPIS(0.08, 1.0)
g.load_state_data(state_path("hammer_state"))
X = ((torch.rand(1000, 3, generator=torch.Generator().manual_seed(0), dtype=torch.float64) - 0.5) * 0.36).cuda()
check(g, X, "hammer N=317 M=1000", variants)
X1, y1, noise = synthetic_training_set(1100, seed=1)
g2 = GPIS(0.08, 1.0)
g2.fit(X1.cuda(), y1.cuda(), noise.cuda())
check(g2, synthetic_queries(3000, 0, 3000).cuda(), "synthetic N=1100 M=3000", variants)
X1, y1, noise = synthetic_training_set(4096, seed=0)
g3 = GPIS(0.08, 1.0)
g3.fit(X1.cuda(), y1.cuda(), noise.cuda())
M = 1 << 20
X = synthetic_queries(M, 0, M).cuda()
check(g3, X[:20000], "synthetic N=4096 M=20000", variants)
for v in variants:
    g3.set_i8_variant(v)
    g3.set_var_mode(5)
    with torch.no_grad():
        g3.pred(X)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            g3.pred(X)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 3
    print(f"N=4096 variant {v} S=5: {M / dt:.3e} unit queries/s ({dt * 1e3:.1f} ms per {M} queries, mean kernel included)", flush=True)
