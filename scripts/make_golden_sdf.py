"""Runs ON THE GPU BOX: the reference's own TorchSDF kernel (oracle/_ref/torchsdf_ref.so, built from the unmodified sources by
oracle/build_ref_torchsdf.py) on seeded inputs -> gpurun_out/sdf_reference_b200.npz, committed as
tests/golden/sdf_reference_b200.npz.  Also prints how cdx_sdf_forward compares (bit-exact fractions)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_torchsdf, sdf_oracle as so  # noqa: E402
from compliancedex_b200.sdf import compute_sdf  # noqa: E402


def soup(F, seed, dtype):
    rng = np.random.default_rng(seed)
    c = rng.random((F, 1, 3)) * 2 - 1
    return (c + 0.15 * rng.standard_normal((F, 3, 3))).astype(dtype)


def main():
    assert ref_torchsdf.load() is not None, "oracle/_ref/torchsdf_ref.so missing"
    out = {}
    cases = [("cube_f32", so.cube_mesh(0.5, np.float32), 1.0), ("sphere_f32", so.icosphere(3, 1.0, np.float32), 1.3),
             ("soup_f32", soup(3000, 4, np.float32), 1.2), ("sphere_f64", so.icosphere(2, 0.7, np.float64), 1.0)]
    for i, (tag, fv, box) in enumerate(cases):
        rng = np.random.default_rng(10 + i)
        p = ((rng.random((2000, 3)) * 2 - 1) * box).astype(fv.dtype)
        pt, ft = torch.from_numpy(p).cuda(), torch.from_numpy(fv).cuda()
        d, s, n, c = ref_torchsdf.compute_sdf(pt, ft)
        g = ref_torchsdf.backward(torch.ones_like(d) * 0.5, pt, c)
        out.update({f"{tag}_points": p, f"{tag}_faces": fv, f"{tag}_dist": d.cpu().numpy(), f"{tag}_sign": s.cpu().numpy(),
                    f"{tag}_normals": n.cpu().numpy(), f"{tag}_closest": c.cpu().numpy(), f"{tag}_grad": g.cpu().numpy()})
        d2, s2, n2, c2 = compute_sdf(pt, ft)
        print(tag, "cdx vs reference kernel: dist bit-exact %.4f, max |d| %.3e; sign equal %.4f; closest bit-exact %.4f, max %.3e;"
              " normals max %.3e" % ((d2 == d).float().mean().item(), (d2 - d).abs().max().item(), (s2 == s).float().mean().item(),
                                     (c2 == c).all(dim=1).float().mean().item(), (c2 - c).abs().max().item(),
                                     (n2 - n).abs().max().item()))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "sdf_reference_b200.npz"), **out)


if __name__ == "__main__":
    main()
