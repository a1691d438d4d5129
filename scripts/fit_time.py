"""phase timing of cdx_gpis_fit (assembly, Cholesky, L^-1, E11^-1, alpha + packing) at one or more sizes, CUDA events inside the
library (cdx_gpis_fit_profile).  `python scripts/fit_time.py [libcdx.so] [N ...]`"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import _lib  # noqa: E402

args = sys.argv[1:]
if args and args[0].endswith(".so"):
    _lib.LIB_PATH = args.pop(0)
from compliancedex_b200.synthetic import synthetic_training_set  # noqa: E402

lib = _lib.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for N in [int(a) for a in args] or [4096, 16384]:
    X1, y1, noise = synthetic_training_set(N, seed=0)
    X1d, y1d, nzd = X1.cuda(), y1.reshape(-1).contiguous().cuda(), noise.cuda()
    ctx = _lib.new_context(0)
    for _ in range(2):
        _lib.check(lib.cdx_gpis_fit(ctx, X1d.data_ptr(), y1d.data_ptr(), nzd.data_ptr(), 0.0, N, 1.0, 0, 0.08, st))
    _lib.check(lib.cdx_gpis_fit_profile(ctx, 1))
    best = None
    for _ in range(3):
        _lib.check(lib.cdx_gpis_fit(ctx, X1d.data_ptr(), y1d.data_ptr(), nzd.data_ptr(), 0.0, N, 1.0, 0, 0.08, st))
        ph = (ctypes.c_double * 5)()
        _lib.check(lib.cdx_gpis_fit_profile_read(ctx, ph, st))
        if best is None or sum(ph) < sum(best):
            best = list(ph)
    print("N=%d: assembly %.2f cholesky %.2f inverse_factor %.2f inverse_matrix %.2f alpha %.2f  total %.2f ms" % (N, *best, sum(best)), flush=True)
    ctx.close()
