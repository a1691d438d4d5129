"""Drop-in for `torchsdf` (reference: thirdparty/TorchSDF/torchsdf/sdf.py:1-64) backed by libcdx.so - SURVEY.md 8f-3.

    from compliancedex_b200.sdf import compute_sdf, index_vertices_by_faces

`compute_sdf(pointclouds [P,3], face_vertices [F,3,3]) -> (squared distance [P], sign [P] int32, normals [P,3],
closest points [P,3])`, float32 or float64 CUDA tensors; differentiable w.r.t. the points through the squared distance
only (the other three outputs are marked non-differentiable, sdf.py:50-51).  There is no CPU path.
"""
import ctypes

import torch

from . import _lib


def index_vertices_by_faces(vertices_features, faces):
    """sdf.py:5-31: per-vertex features [V,3] + face indices [F,3] -> per-face vertex features [F,3,3]."""
    assert vertices_features.ndim == 2, "vertices_features must have 2 dimensions of shape (batch_sizenum_points, knum)"
    assert faces.ndim == 2, "faces must have 2 dimensions of shape (num_faces, num_vertices)"
    return vertices_features[faces.reshape(-1)].reshape(faces.shape[0], faces.shape[1], vertices_features.shape[-1])


def _check(t, name, shape_tail):
    # csrc/unbatched_triangle_distance.cpp:48-66 (CHECK_CUDA / CHECK_CONTIGUOUS / CHECK_SIZES)
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if t.ndim != 1 + len(shape_tail) or tuple(t.shape[1:]) != shape_tail:
        raise RuntimeError(f"{name} must have shape (n, {', '.join(map(str, shape_tail))})")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


class _UnbatchedTriangleDistanceCuda(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, face_vertices):
        _check(points, "points", (3,))
        _check(face_vertices, "face_vertices", (3, 3))
        if points.dtype not in (torch.float32, torch.float64) or face_vertices.dtype != points.dtype:
            raise RuntimeError("points and face_vertices must both be float32 or both float64")   # data_ptr<scalar_t> in the reference
        if face_vertices.device != points.device:
            raise RuntimeError("points and face_vertices must be on the same device")
        P = points.shape[0]
        dev = points.device
        min_dist = torch.empty(P, device=dev, dtype=points.dtype)
        dist_sign = torch.empty(P, device=dev, dtype=torch.int32)
        normals = torch.empty(P, 3, device=dev, dtype=points.dtype)
        clst_points = torch.empty(P, 3, device=dev, dtype=points.dtype)
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(_lib.load().cdx_sdf_forward(_lib.context(dev.index), _ptr(points), P, _ptr(face_vertices),
                                               face_vertices.shape[0], int(points.dtype == torch.float64), _ptr(min_dist),
                                               _ptr(dist_sign), _ptr(normals), _ptr(clst_points), st))
        ctx.save_for_backward(points, clst_points)
        ctx.mark_non_differentiable(dist_sign, normals, clst_points)
        return min_dist, dist_sign, normals, clst_points

    @staticmethod
    def backward(ctx, grad_dist, grad_dist_sign, grad_normals, grad_clst_points):
        points, clst_points = ctx.saved_tensors
        grad_dist = grad_dist.contiguous()
        grad_points = torch.empty_like(points)
        dev = points.device
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(_lib.load().cdx_sdf_backward(_lib.context(dev.index), _ptr(grad_dist), _ptr(points), _ptr(clst_points),
                                                points.shape[0], int(points.dtype == torch.float64), _ptr(grad_points), st))
        return grad_points, None


def set_search_mode(mode, device=None):
    """0 auto (bounding-sphere culled search when the mesh has >= 512 faces and there are >= 128 x SMs points), 1 exhaustive point x face scan (the reference's
    algorithm), 2 culled always.  Every mode returns the same face, bit for bit."""
    index = torch.cuda.current_device() if device is None else torch.device(device).index
    _lib.check(_lib.load().cdx_sdf_set_mode(_lib.context(index), int(mode)))


def compute_sdf(pointclouds, face_vertices):
    return _UnbatchedTriangleDistanceCuda.apply(pointclouds, face_vertices)
