"""in-kernel clock64 breakdown of the rows-as-M INT8 kernel (experimental build -DEXP=9, block 0): generator warp 0 and the MMA warp"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import _lib  # noqa: E402

_lib.LIB_PATH = "compliancedex_b200/csrc/build/exp/libcdx_exp9.so"
from compliancedex_b200 import GPIS  # noqa: E402
from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set  # noqa: E402

variant = int(sys.argv[1])
X1, y1, noise = synthetic_training_set(4096, seed=0)
g = GPIS(0.08, 1.0)
g.fit(X1.cuda(), y1.cuda(), noise.cuda())
waves = int(sys.argv[2]) if len(sys.argv) > 2 else 4
M = 96 * (74 if variant == 1 else 148) * waves
X = synthetic_queries(M, 0, M).cuda()
g.set_i8_variant(variant)
g.set_var_mode(5)
std = torch.empty(M, dtype=torch.float64, device="cuda")
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
_lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, None, None, std.data_ptr(), None, st))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
z = (ctypes.c_longlong * 16)()
e0.record()
_lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, None, None, std.data_ptr(), None, st))
e1.record()
torch.cuda.synchronize()
print("kernel (2nd launch) %.3f ms for %d waves -> %.3f ms per wave" % (e0.elapsed_time(e1), waves, e0.elapsed_time(e1) / waves))
t = (ctypes.c_longlong * 16)()
raw.cdx_debug_i8x_ticks(t)
names = ["gen wait a_full", "gen compute", "gen wait empty", "gen store+fence+arrive", "mma wait acc_empty", "mma wait a_full",
         "mma wait b_full", "mma issue+commit", "gen items", "mma stages"]
for n, v in zip(names, t):
    print(f"{n:28s} {v:12d}")
print("per generator item: compute %.0f, waits %.0f + %.0f, store %.0f cycles" % (t[1] / max(t[8], 1), t[0] / max(t[8], 1), t[2] / max(t[8], 1), t[3] / max(t[8], 1)))
print("block 0: %d cycles in %d ns -> %.3f GHz; last block: %d cycles in %d ns -> %.3f GHz (both launches summed for the phase counters)" % (t[10], t[11], t[10] / max(t[11], 1), t[12], t[13], t[12] / max(t[13], 1)))
print("mma fence_after per stage %.0f; mma loop total per stage %.0f" % (t[14] / max(t[9], 1), t[15] / max(t[9], 1)))
print("per mma stage: wait acc %.0f, a_full %.0f, b_full %.0f, issue %.0f cycles" % (t[4] / max(t[9], 1), t[5] / max(t[9], 1), t[6] / max(t[9], 1), t[7] / max(t[9], 1)))
