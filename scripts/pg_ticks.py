"""Section timing of pg_step_kernel: run with CDX_LIB=.../libcdx_ticks.so (scripts/build_variant.sh _ticks cdx_pregrasp.cu
-DCDX_PG_TICKS).  Thread 0 of CTA 0 stores clock64() at 13 points; prints the deltas of the LAST iteration in cycles."""
import ctypes, json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench_pregrasp import _make
from compliancedex_b200 import _lib

dev = torch.device("cuda", 0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
gp, opt, q0, T0, k0, palm, N = _make(0, B, 30, dev)
lib = _lib.load()
opt._init(q0.to(dev), T0.to(dev), k0.to(dev), 1, gp)
opt.step(30)
torch.cuda.synchronize()
raw = ctypes.CDLL(_lib.LIB_PATH)
buf = (ctypes.c_longlong * 16)()
assert raw.cdx_debug_pg_ticks(buf) == 0
t = list(buf)
names = [("start->row (loads, Euler, noise)", 7, 8), ("row (4 lanes per copy)", 8, 6), ("row end -> reduced over copies", 6, 9),
         ("chain rule + grad", 9, 10), ("tracking + Adam", 10, 11), ("FK new iterate", 11, 12), ("whole kernel", 7, 12)]
print(json.dumps({"B": B, "cycles": {n: t[b] - t[a] for n, a, b in names}}))
