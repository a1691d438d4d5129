"""standalone timing of the unit query {mean, grad mean, std} on the shipped states (FP64 DMMA kernel gpis_var_kernel<unit>):
CUDA events, queries resident.  `python scripts/unit_time.py [path/to/libcdx.so]` (default: the product library)"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import _lib  # noqa: E402

if len(sys.argv) > 1:
    _lib.LIB_PATH = sys.argv[1]
from compliancedex_b200 import GPIS, state_path  # noqa: E402

lib = _lib.load()
for name, box in (("mug", 0.1), ("coffeebottle", 0.1), ("lego", 0.1), ("hammer", 0.18), ("banana", 0.15), ("dummy", 0.1)):
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    N = g.X1.shape[0]
    for M in (4608, 262144):
        X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * box).cuda()
        out = [torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda"),
               torch.empty(M, dtype=torch.float64, device="cuda")]
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

        def step():
            _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), None, st))
        for _ in range(3):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            step()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        tf = (N * N + 26 * N) * M / (us * 1e-6) / 1e12
        print(f"{name} N={N} M={M}: {us:9.1f} us  {tf:5.2f} TF/s  checksum {float(out[2].sum()):.12e}", flush=True)
