"""timing of the experimental builds of the rows-as-M INT8 kernel (csrc/build/exp, -DEXP=n; results are wrong by design)"""
import sys
import time

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import _lib  # noqa: E402

exp, variant = sys.argv[1], int(sys.argv[2])
if exp != "0":
    _lib.LIB_PATH = "compliancedex_b200/csrc/build/exp/libcdx_exp%s.so" % exp
from compliancedex_b200 import GPIS  # noqa: E402
from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set  # noqa: E402

X1, y1, noise = synthetic_training_set(4096, seed=0)
g = GPIS(0.08, 1.0)
g.fit(X1.cuda(), y1.cuda(), noise.cuda())
M = 1 << 19
X = synthetic_queries(M, 0, M).cuda()
g.set_i8_variant(variant)
g.set_var_mode(5)
std = torch.empty(M, dtype=torch.float64, device="cuda")
import ctypes  # noqa: E402
lib = _lib.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def step():
    _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, None, None, std.data_ptr(), None, st))


step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    step()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 3
print(f"EXP={exp} variant {variant}: std kernel alone {M / dt:.3e} q/s, {dt * 1e3:.1f} ms per {M} queries", flush=True)
