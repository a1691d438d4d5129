"""timing of one build of the extended query on hammer (N = 317): `python scripts/ext_exp.py base` = the product library,
`python scripts/ext_exp.py <n>` = csrc/build/exp/libcdx_ext_exp<n>.so (scripts/make_ext_exp.py; results wrong by design)"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import _lib  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "base"
if tag != "base":
    _lib.LIB_PATH = "compliancedex_b200/csrc/build/exp/libcdx_ext_exp%s.so" % tag
from compliancedex_b200 import GPIS, state_path  # noqa: E402

lib = _lib.load()
g = GPIS(0.08, 1.0)
g.load_state_data(state_path("hammer_state"))
for M in (16384, 131072):
    X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * 0.18).cuda()
    out = [torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda"),
           torch.empty(M, dtype=torch.float64, device="cuda"), torch.empty(M, 3, dtype=torch.float64, device="cuda")]
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

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
    print(tag, "hammer", M, "%.1f us" % (e0.elapsed_time(e1) / 20 * 1e3), flush=True)
