"""standalone timing of the extended query (mean, grad mean, std, grad std) on the shipped states: all-columns kernel (mode 0)
vs column-tiled kernel (mode 1), CUDA events, queries resident; fraction of the FP64 DMMA peak given on the command line"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, _lib, state_path  # noqa: E402

peak = float(sys.argv[1]) if len(sys.argv) > 1 else 37.0
lib = _lib.load()
for name, box in (("mug", 0.1), ("hammer", 0.18), ("dummy", 0.1)):
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    N = g.X1.shape[0]
    for M in (2048, 4096, 16384, 131072):
        X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * box).cuda()
        out = [torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda"),
               torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda")]
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        res = []
        for mode in (0, 2, 1):
            g.set_ext_mode(mode)

            def step():
                _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(), st))
            for _ in range(3):
                step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                step()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 20 * 1e3
            tf = (2 * N * N + 34 * N) * M / (us * 1e-6) / 1e12
            res.append("mode %d: %8.1f us %5.2f TF/s %.2f" % (mode, us, tf, tf / peak))
        print(f"{name} N={N} M={M}: " + " | ".join(res), flush=True)
