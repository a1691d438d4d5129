import ctypes, sys
import torch
sys.path.insert(0, ".")
from compliancedex_b200 import _lib
if len(sys.argv) > 1 and sys.argv[1] != "base":
    _lib.LIB_PATH = sys.argv[1]
from compliancedex_b200 import GPIS, state_path
lib = _lib.load()
for name, box in (("hammer", 0.18),):
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    N = g.X1.shape[0]
    for M in (16384, 131072):
        X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * box).cuda()
        out = [torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda"),
               torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda")]
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        def step():
            _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(), st))
        for _ in range(3): step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): step()
        e1.record(); torch.cuda.synchronize()
        print(sys.argv[1] if len(sys.argv) > 1 else "base", name, M, "%.1f us" % (e0.elapsed_time(e1) / 20 * 1e3), flush=True)
