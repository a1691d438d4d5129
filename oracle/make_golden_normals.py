"""TEST TOOLING (build container only): goldens for GPIS.compute_normal(X2, index=...) (gpis.py:63-87, the `index`
branch :66-69,78-80,84-87) and GPIS.compute_multinormals (gpis.py:89-111), produced by the reference's own gpis.py
(imported unmodified from /root/reference through oracle/ref_harness.py).   python -m oracle.make_golden_normals
"""
import os

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.make_golden import BOX, GOLD


def main():
    rh.install()
    out = {}
    for name in ["lego", "hammer"]:
        g = rh.load_gpis(f"{name}_state")
        N = len(g.X1)
        b = BOX.get(name, 0.1)
        gen = torch.Generator().manual_seed(77 + N)
        X = (torch.rand(48, 3, generator=gen, dtype=torch.float64) * 2 - 1) * b
        Xb = X[:40].reshape(10, 4, 3)                                    # the [B,F,3] shape rule of :64-65,:85-87
        subsets = {"prefix": list(range(N // 2)),
                   "suffix": list(range(N // 3, N)),
                   "scattered": sorted(torch.randperm(N, generator=gen)[: (2 * N) // 3].tolist()),
                   "single": [N - 1]}
        out[f"{name}_X"] = X.numpy()
        for tag, idx in subsets.items():
            n, w = g.compute_normal(X, idx)
            out[f"{name}_{tag}_index"] = np.asarray(idx, dtype=np.int64)
            out[f"{name}_{tag}_normal"] = n.detach().numpy()
            out[f"{name}_{tag}_weight"] = np.asarray(w.detach().numpy())
        nb, wb = g.compute_normal(Xb, subsets["prefix"])
        out[f"{name}_prefix_normal_batched"] = nb.detach().numpy()
        for S in (3, 5):
            g.fraction = None                                             # the reference caches the row subsets of the first call
            nm, wm = g.compute_multinormals(X, S)
            out[f"{name}_multi{S}_normals"] = nm.detach().numpy()
            out[f"{name}_multi{S}_weights"] = wm.detach().numpy()
            out[f"{name}_multi{S}_sizes"] = np.asarray([len(i) for i in g.indices], dtype=np.int64)
            out[f"{name}_multi{S}_first"] = np.asarray([i[0] for i in g.indices], dtype=np.int64)
    np.savez_compressed(os.path.join(GOLD, "gpis_normals_index.npz"), **out)
    print("normals/index goldens:", sorted(out)[:6], "...", len(out), "arrays")


if __name__ == "__main__":
    main()
