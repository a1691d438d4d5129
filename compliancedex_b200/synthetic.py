"""Seeded synthetic inputs of the benchmark shapes: a GPIS training set shaped like
train_gpis_completion.py:20-58 and pregrasp candidates following run_experiments.py:87-90 /
optimize_pregrasp.py:19-24,889-891,940-941,956.  Data generators only - no GPIS or objective arithmetic.
(oracle/ carries its own copies; tests/test_host_logic.py checks both produce identical arrays.)"""
import math

import numpy as np
import torch


def synthetic_training_set(N, seed=0, bound=0.1, radius=0.05):
    """14 exterior points (+bound, sigma .3) | surface points on a sphere (label 0; sigma .005 for the first 2/3
    'observed', .02 for the 'completed' rest) | ceil(.03N) interior softmax-convex combinations (label -bound,
    sigma .02).  Returns X1[N,3], y1[N,1] (before bias subtraction), noise[N]."""
    g = torch.Generator().manual_seed(seed)
    b = bound
    ext = torch.tensor([[-b, -b, -b], [b, -b, -b], [-b, b, -b], [b, b, -b],
                        [-b, -b, b], [b, -b, b], [-b, b, b], [b, b, b],
                        [-b, 0., 0.], [0., -b, 0.], [b, 0., 0.], [0., b, 0.],
                        [0., 0., b], [0., 0., -b]], dtype=torch.float64)
    n_int = int(np.ceil(0.03 * N))
    n_surf = N - 14 - n_int
    v = torch.randn(n_surf, 3, generator=g, dtype=torch.float64)
    surf = radius * v / v.norm(dim=1, keepdim=True)
    w = torch.softmax(10.0 * torch.rand(n_int, n_surf, generator=g, dtype=torch.float64), dim=1)
    interior = w @ surf
    X1 = torch.vstack([ext, surf, interior])
    y1 = torch.vstack([b * torch.ones(14, 1), torch.zeros(n_surf, 1), -b * torch.ones(n_int, 1)]).double()
    n_obs = (2 * n_surf) // 3
    noise = torch.tensor([0.3] * 14 + [0.005] * n_obs + [0.02] * (n_surf - n_obs) + [0.02] * n_int, dtype=torch.float64)
    return X1, y1, noise


def synthetic_queries(m_total, lo, hi, bound=0.1, seed=1234):
    """queries [lo:hi) of the m_total-point stream, uniform in [-bound,bound]^3; generated in fixed 2^20 blocks so a
    rank's slice does not depend on the world size"""
    g = torch.Generator().manual_seed(seed)
    blk = 1 << 20
    parts = []
    for b0 in range(0, m_total, blk):
        n = min(blk, m_total - b0)
        x = (torch.rand(n, 3, generator=g, dtype=torch.float64) * 2 - 1) * bound
        s, e = max(lo, b0), min(hi, b0 + n)
        if s < e:
            parts.append(x[s - b0:e - b0])
    return torch.cat(parts) if parts else torch.empty(0, 3, dtype=torch.float64)


def surface_center(X1, y1_bias_subtracted):
    """centre of the axis-aligned box of the label-0 (surface) training points (object centre, :889-891)"""
    X1 = np.asarray(X1)
    surf = X1[np.abs(np.asarray(y1_bias_subtracted).reshape(-1) + 1.0) < 1e-9]
    return ((surf.max(0) + surf.min(0)) / 2).tolist()


def synthetic_candidates(hand_doc, B, center, seed=0):
    """q = ref_q + 0.1 N(0,1); palm = tiled WRIST_OFFSET rows + centre shift + N(0,0.01) on xyz; targets = centre x4;
    compliance [10,10,10,20].  -> q [B,16], target [B,4,3], compliance [B,4] (float64 tensors), palm [B,6] (numpy)"""
    g = torch.Generator().manual_seed(seed)
    ref_q = torch.tensor(hand_doc["ref_q"], dtype=torch.float64)
    q = ref_q.unsqueeze(0) + 0.1 * torch.randn(B, 16, generator=g, dtype=torch.float64)
    wrist = torch.tensor([[-0.06, 0.0, 0.05, 0.0, 0.0, 0.0],
                          [-0.04, 0.03, 0.05, 0.0, 0.0, -math.pi / 4],
                          [-0.01, 0.0, 0.05, 0.0, 0.0, math.pi / 4],
                          [0.1, 0.06, 0.03, -math.pi / 2, math.pi / 2, 0.0],
                          [-0.0, -0.06, 0.05, 0.0, 0.0, math.pi / 2],
                          [0.02, -0.04, 0.05, 0.0, 0.0, 3 * math.pi / 4]], dtype=torch.float64)
    c = torch.as_tensor(center, dtype=torch.float64)
    palm = wrist[torch.arange(B) % 6].clone()
    palm[:, 0] += c[0]
    palm[:, 1] += c[1]
    palm[:, 2] += 2 * c[2]
    palm[:, :3] += 0.01 * torch.randn(B, 3, generator=g, dtype=torch.float64)
    target = c.view(1, 1, 3).repeat(B, 4, 1)
    compliance = torch.tensor([[10.0, 10.0, 10.0, 20.0]], dtype=torch.float64).repeat(B, 1)
    return q, target, compliance, palm.numpy()


def icosphere_faces(subdiv=3, radius=1.0, dtype=np.float32):
    """Synthetic closed mesh for the compute_sdf workload: an icosahedron subdivided `subdiv` times, vertices pushed to the
    sphere of `radius` -> face_vertices [20 * 4^subdiv, 3, 3] with outward-facing normals."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    verts = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1),
             (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [np.asarray(v, dtype=np.float64) / np.linalg.norm(v) for v in verts]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
             (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7),
             (9, 8, 1)]
    for _ in range(subdiv):
        midpoint, split = {}, []
        for tri in faces:
            m = []
            for a, b in ((tri[0], tri[1]), (tri[1], tri[2]), (tri[2], tri[0])):
                key = (a, b) if a < b else (b, a)
                if key not in midpoint:
                    v = verts[a] + verts[b]
                    verts.append(v / np.linalg.norm(v))
                    midpoint[key] = len(verts) - 1
                m.append(midpoint[key])
            split += [(tri[0], m[0], m[2]), (tri[1], m[1], m[0]), (tri[2], m[2], m[1]), (m[0], m[1], m[2])]
        faces = split
    return (np.asarray(verts)[np.asarray(faces)] * radius).astype(dtype)
