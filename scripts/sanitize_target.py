"""small launches of the kernels with hand-rolled synchronisation, for compute-sanitizer (racecheck / memcheck / synccheck):
the tcgen05 INT8 std kernel in its three launch shapes (mbarrier rings, tcgen05.commit, st.async into a peer CTA, CTA pair),
the all-columns extended kernel (cp.async ring) and the column-tiled extended kernel with its cluster / DSMEM reduction."""
import sys

import torch

sys.path.insert(0, ".")
from compliancedex_b200 import GPIS, state_path  # noqa: E402
from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set  # noqa: E402

X1, y1, noise = synthetic_training_set(300, seed=1)
g = GPIS(0.08, 1.0)
g.fit(X1.cuda(), y1.cuda(), noise.cuda())
X = synthetic_queries(200, 0, 200).cuda()
g.set_var_mode(-1)
with torch.no_grad():
    _, s0 = g.pred(X)
for v in (3, 1, 2):
    g.set_i8_variant(v)
    g.set_var_mode(5)
    with torch.no_grad():
        m, s = g.pred(X)
    torch.cuda.synchronize()
    print("i8 variant", v, "max |d var| / prior", float(((s ** 2 - s0 ** 2).abs().max() / float(g.R) ** 3)))
h = GPIS(0.08, 1.0)
h.load_state_data(state_path("mug_state"))
Xh = ((torch.rand(100, 3, generator=torch.Generator().manual_seed(0), dtype=torch.float64) - 0.5) * 0.2).cuda()
for mode in (0, 1):
    h.set_ext_mode(mode)
    out = h.pred_with_grads(Xh)
    torch.cuda.synchronize()
    print("ext mode", mode, float(out[1].sum()))
print("done")
