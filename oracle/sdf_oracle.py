"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

CPU restatement (numpy) of the reference's point-to-mesh distance kernel
(thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu:176-254, backward :256-270; Python wrapper
torchsdf/sdf.py:34-64).  Vectorised over points, Python loop over faces.  Arithmetic in the dtype of the inputs, with the
reference's two float truncations kept: the edge parameter passed to point_at (`float t`, :170) and the squared distance
(`float dist`, :235).  numpy does not contract a*b+c into FMAs and its rsqrt is 1/sqrt, so float32 results agree with the
CUDA kernels to rounding (~1e-6 relative), not bit for bit; near-ties between faces may resolve differently.

Pinned against (tests/test_sdf_oracle.py, tests/test_sdf_gpu.py): closed-form distances to an axis-aligned cube and a
sphere, and tests/golden/sdf_reference_b200.npz = outputs of the REFERENCE's own kernel compiled unmodified for sm_100a
(oracle/build_ref_torchsdf.py) and run on a B200 by scripts/make_golden_sdf.py.
"""
import numpy as np


def _dot(a, b):
    return a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1] + a[..., 2] * b[..., 2]


def _cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1], a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)


def face_closest(p, v1, v2, v3):
    """closest point of ONE face to points p [P,3] (:192-231) -> closest [P,3], face normal [3]"""
    T = p.dtype.type
    e12, e23, e31 = v2 - v1, v3 - v2, v1 - v3
    n = _cross(v1 - v2, e31)
    with np.errstate(divide="ignore", invalid="ignore"):
        uab = _dot(p - v1, e12) / _dot(e12, e12)
        uca = _dot(p - v3, e31) / _dot(e31, e31)
        ubc = _dot(p - v2, e23) / _dot(e23, e23)
        inv = T(1) / np.sqrt(_dot(n, n))
    un = n * inv
    d = _dot(p - v1, un)
    cp = p - un * d[:, None]                                                 # project_plane (:137-146)
    done = np.zeros(p.shape[0], dtype=bool)

    def take(mask, val):
        m = mask & ~done
        cp[m] = val[m] if val.ndim == 2 else val
        done[m] = True

    def on_edge(v, e, u):                                                    # point_at(v, e, float t) (:170)
        return v + e * u.astype(np.float32).astype(p.dtype)[:, None]

    take((uca > 1) & (uab < 0), v1)
    take((uab > 1) & (ubc < 0), v2)
    take((ubc > 1) & (uca < 0), v3)
    take((uab <= 1) & (uab >= 0) & (_dot(_cross(n, e12), p - v1) <= 0), on_edge(v1, e12, uab))
    take((ubc <= 1) & (ubc >= 0) & (_dot(_cross(n, e23), p - v2) <= 0), on_edge(v2, e23, ubc))
    take((uca <= 1) & (uca >= 0) & (_dot(_cross(n, e31), p - v3) <= 0), on_edge(v3, e31, uca))
    return cp, n


def compute_sdf(points, face_vertices):
    """-> (dist [P] squared distance (float-rounded), sign [P] int32, normals [P,3], closest [P,3], face index [P])"""
    p = np.ascontiguousarray(points)
    fv = np.ascontiguousarray(face_vertices).astype(p.dtype)
    P = p.shape[0]
    best = np.zeros(P, dtype=p.dtype)
    sign = np.zeros(P, dtype=np.int32)
    normals = np.zeros((P, 3), dtype=p.dtype)
    closest = np.zeros((P, 3), dtype=p.dtype)
    index = np.full(P, -1, dtype=np.int64)
    for f in range(fv.shape[0]):
        cp, n = face_closest(p, fv[f, 0], fv[f, 1], fv[f, 2])
        dv = p - cp
        d2 = _dot(dv, dv)
        dist = d2.astype(np.float32).astype(p.dtype)                         # `float dist` (:235)
        gn = dv * (p.dtype.type(1) / np.sqrt(np.float32(1e-16).astype(p.dtype) + d2))[:, None]
        sg = np.where(_dot(dv, n) >= 0, 1, -1).astype(np.int32)
        upd = (best > dist) if f else np.ones(P, dtype=bool)                 # first face always taken, then strict < (:236)
        best[upd], sign[upd], normals[upd], closest[upd], index[upd] = dist[upd], sg[upd], gn[upd], cp[upd], f
    return best, sign, normals, closest, index


def backward(grad_dist, points, closest):
    """:256-270"""
    return (points - closest) * (points.dtype.type(2.0) * grad_dist)[:, None]


# ------------------------------------------------------------------ test meshes (generated, no files needed)
def cube_mesh(h=0.5, dtype=np.float32):
    """axis-aligned cube [-h,h]^3, 12 outward-facing triangles -> face_vertices [12,3,3]"""
    v = np.array([[x, y, z] for x in (-h, h) for y in (-h, h) for z in (-h, h)], dtype=np.float64)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    faces = []
    for a, b, c, d in quads:
        faces += [(a, b, c), (a, c, d)]
    return v[np.array(faces)].astype(dtype)


def icosphere(subdiv=3, radius=1.0, dtype=np.float32):
    """unit icosahedron subdivided `subdiv` times (20 * 4^subdiv outward-facing triangles) -> face_vertices [F,3,3]"""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1),
         (-t, 0, -1), (-t, 0, 1)]
    v = [np.array(x, dtype=np.float64) / np.linalg.norm(x) for x in v]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
         (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]

        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return (np.array(v)[np.array(f)] * radius).astype(dtype)


def cube_sdf_exact(p, h=0.5):
    """signed distance to the cube [-h,h]^3 (positive outside)"""
    q = np.abs(p) - h
    return np.linalg.norm(np.maximum(q, 0.0), axis=1) + np.minimum(q.max(axis=1), 0.0)
