"""CPU: the numpy restatement of TorchSDF's kernel (oracle/sdf_oracle.py) against closed-form distances and, when the
golden file exists, against outputs of the reference's own kernel run on a B200 (tests/golden/sdf_reference_b200.npz)."""
import os

import numpy as np
import pytest

from oracle import sdf_oracle as so


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_cube_known_answers(dtype):
    rng = np.random.default_rng(0)
    p = (rng.random((3000, 3)) * 2 - 1).astype(dtype)
    d2, sign, normals, closest, idx = so.compute_sdf(p, so.cube_mesh(0.5, dtype))
    exact = so.cube_sdf_exact(p.astype(np.float64), 0.5)
    assert np.abs(np.sqrt(d2.astype(np.float64)) - np.abs(exact)).max() < 2e-6
    # outward normals: + outside.  The sign is taken from the closest FACE's normal (:234), which is only the true
    # inside/outside test when the closest point is interior to that face (exactly one coordinate on the surface)
    interior = (np.abs(np.abs(closest) - 0.5) < 1e-6).sum(axis=1) == 1
    assert interior.mean() > 0.3 and np.array_equal(sign[interior], np.where(exact > 0, 1, -1)[interior])
    assert np.abs(np.abs(closest).max(axis=1) - 0.5).max() < 1e-6            # closest points lie on the surface
    assert np.abs(np.linalg.norm(normals, axis=1) - 1).max() < 1e-5
    assert np.abs(normals * np.sqrt(d2)[:, None] - (p - closest)).max() < 1e-6
    # a point straight above the middle of a face, an edge and a corner
    q = np.array([[0.1, 0.2, 1.0], [1.0, 1.0, 0.0], [1.0, 1.0, 1.0], [0.0, 0.0, 0.0]], dtype=dtype)
    d2, sign, _, closest, _ = so.compute_sdf(q, so.cube_mesh(0.5, dtype))
    assert np.allclose(d2, [0.25, 0.5, 0.75, 0.25], atol=1e-6)
    assert np.allclose(closest[:3], [[0.1, 0.2, 0.5], [0.5, 0.5, 0.0], [0.5, 0.5, 0.5]], atol=1e-6)
    assert list(sign) == [1, 1, 1, -1]


def test_sphere_and_backward():
    rng = np.random.default_rng(1)
    p = (rng.standard_normal((2000, 3)) * 0.8).astype(np.float32)
    fv = so.icosphere(3, 1.0)                                                # 1280 faces, inscribed polyhedron
    d2, sign, normals, closest, idx = so.compute_sdf(p, fv)
    r = np.linalg.norm(p.astype(np.float64), axis=1)
    assert np.abs(np.sqrt(d2) - np.abs(r - 1.0)).max() < 0.012              # chordal error of the subdivision
    far = np.abs(r - 1.0) > 0.02
    assert np.array_equal(sign[far], np.where(r[far] > 1, 1, -1))
    g = so.backward(np.ones_like(d2), p, closest)
    assert np.array_equal(g, 2 * (p - closest))
    # float truncation of the squared distance is kept for float64 inputs
    d64 = so.compute_sdf(p.astype(np.float64), fv.astype(np.float64))[0]
    assert d64.dtype == np.float64 and np.array_equal(d64, d64.astype(np.float32).astype(np.float64))


def test_first_face_wins_ties():
    tri = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], dtype=np.float32)
    fv = np.concatenate([tri, tri, tri + np.float32(5)])
    _, _, _, _, idx = so.compute_sdf(np.array([[0.2, 0.2, 1.0], [5.2, 5.2, 6.0]], dtype=np.float32), fv)
    assert list(idx) == [0, 2]


def test_oracle_against_reference_kernel_golden(golden_dir):
    path = os.path.join(golden_dir, "sdf_reference_b200.npz")
    if not os.path.isfile(path):
        pytest.skip("tests/golden/sdf_reference_b200.npz not generated yet (scripts/make_golden_sdf.py on a GPU box)")
    g = np.load(path)
    for tag in ("cube_f32", "sphere_f32", "soup_f32", "sphere_f64"):
        p, fv = g[f"{tag}_points"], g[f"{tag}_faces"]
        d2, sign, normals, closest, idx = so.compute_sdf(p, fv)
        scale = max(1.0, np.abs(p).max()) ** 2
        assert np.abs(d2 - g[f"{tag}_dist"]).max() <= 2e-6 * scale, tag
        same = np.abs(closest - g[f"{tag}_closest"]).max(axis=1) <= 1e-5    # near-ties between faces may pick another face
        assert same.mean() > 0.995, (tag, same.mean())
        clear = same & (g[f"{tag}_dist"] > 1e-8)
        assert np.array_equal(sign[clear], g[f"{tag}_sign"][clear]) or (sign[clear] != g[f"{tag}_sign"][clear]).mean() < 2e-3, tag
        assert np.abs(normals[clear] - g[f"{tag}_normals"][clear]).max() <= 2e-3, tag
