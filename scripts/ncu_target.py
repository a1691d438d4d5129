"""small single-purpose drivers for ncu captures (profiles/r02_*): each runs the kernel of interest a few times on a B200.
   python scripts/ncu_target.py ext  [state] [M]     extended GPIS query (gpis_ext_kernel) on a shipped state
   python scripts/ncu_target.py i8   [waves]         INT8 std kernel (gpis_std_i8x_kernel) at N = 4096, `waves` waves of 74 CTA pairs
   python scripts/ncu_target.py fit  [N]             GPIS fit (dgemm_kernel, potrf_diag_kernel)
   python scripts/ncu_target.py unit [state] [M]     unit GPIS query {mean, grad mean, std} (gpis_var_kernel<unit>) on a shipped state
   python scripts/ncu_target.py c5   [B] [iters]     pregrasp iterations at the C5 per-GPU shape (Allegro, hammer)"""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, _lib, state_path  # noqa: E402

what = sys.argv[1]
lib = _lib.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
if what == "ext":
    name = sys.argv[2] if len(sys.argv) > 2 else "hammer"
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * 0.18).cuda()
    for _ in range(4):
        out = g.pred_with_grads(X)
    torch.cuda.synchronize()
elif what == "unit":
    name = sys.argv[2] if len(sys.argv) > 2 else "banana"
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 262144
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path(f"{name}_state"))
    X = ((torch.rand(M, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64) * 2 - 1) * 0.15).cuda()
    std = torch.empty(M, dtype=torch.float64, device="cuda")
    mean = torch.empty(M, dtype=torch.float64, device="cuda")
    gm = torch.empty(M, 3, dtype=torch.float64, device="cuda")
    for _ in range(3):
        _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, mean.data_ptr(), gm.data_ptr(), std.data_ptr(), None, st))
    torch.cuda.synchronize()
elif what == "i8":
    from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set
    waves = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    X1, y1, noise = synthetic_training_set(4096, seed=0)
    g = GPIS(0.08, 1.0)
    g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    M = 96 * 74 * waves
    X = synthetic_queries(M, 0, M).cuda()
    std = torch.empty(M, dtype=torch.float64, device="cuda")
    mean = torch.empty(M, dtype=torch.float64, device="cuda")
    gm = torch.empty(M, 3, dtype=torch.float64, device="cuda")
    for _ in range(3):      # the unit query of bench.py's gpis_c3: {mean, grad_mean, std}
        _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, mean.data_ptr(), gm.data_ptr(), std.data_ptr(), None, st))
    torch.cuda.synchronize()
elif what == "fit":
    from compliancedex_b200.synthetic import synthetic_training_set
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    X1, y1, noise = synthetic_training_set(N, seed=0)
    g = GPIS(0.08, 1.0)
    for _ in range(2):
        g.fit(X1.cuda(), y1.cuda(), noise.cuda())
    torch.cuda.synchronize()
elif what == "c5":
    from compliancedex_b200 import ProbabilisticGraspOptimizer, load_hand
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path("hammer_state"))
    doc = load_hand("allegro")
    d = np.load(state_path("hammer_state"))
    q0, T0, k0, palm = synthetic_candidates(doc, B, surface_center(d["X1"], d["y1"]), seed=100)
    opt = ProbabilisticGraspOptimizer("allegro", ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm, num_iters=iters, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                      optimize_palm=True, noise_seed=2024)
    opt.set_tuning(use_graph=False)
    opt.optimize(q0.cuda(), T0.cuda(), k0.cuda(), 1, gp, verbose=False)
    torch.cuda.synchronize()
print("done", what)
