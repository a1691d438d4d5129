import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from compliancedex_b200 import GPIS, state_path
g = GPIS(0.08, 1.0); g.load_state_data(state_path("mug_state"))
X = ((torch.rand(20000, 3, generator=torch.Generator().manual_seed(0), dtype=torch.float64) - 0.5) * 0.2).cuda()
big = g.pred_with_grads(X)          # TQ=128? (20000/128=157 >= 148)
for M in (32, 72, 100, 268, 4096):
    small = g.pred_with_grads(X[:M].contiguous())
    print(M, [float((a[:M] - b).abs().max()) for a, b in zip(big, small)])
    with torch.no_grad():
        m, s = g.pred(X[:M].contiguous())
    print('  unit', float((big[0][:M]-m).abs().max()), float((big[1][:M]-s).abs().max()))
    n = torch.empty(M, 3, dtype=torch.float64, device='cuda'); mm = torch.empty(M, dtype=torch.float64, device='cuda')
    g._query(X[:M].contiguous(), mm, n, None, None)
    print('  mean', float((big[0][:M]-mm).abs().max()), float((big[2][:M]-n).abs().max()))
