"""TEST INFRASTRUCTURE ONLY.  Loads oracle/_ref/torchsdf_ref.so - the REFERENCE's TorchSDF extension compiled unmodified for
sm_100a by oracle/build_ref_torchsdf.py - and exposes compute_sdf exactly as the reference's Python wrapper drives it
(thirdparty/TorchSDF/torchsdf/sdf.py:38-64).  GPU only; returns None from load() when the library was not built."""
import importlib.util
import os

SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "torchsdf_ref.so")
_mod = None


def load():
    global _mod
    if _mod is None and os.path.isfile(SO):
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        spec = importlib.util.spec_from_file_location("torchsdf_ref", SO)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def compute_sdf(points, face_vertices):
    import torch
    m = load()
    n = points.shape[0]
    dist = torch.zeros(n, device=points.device, dtype=points.dtype)          # sdf.py:42-49
    sign = torch.zeros(n, device=points.device, dtype=torch.int32)
    normals = torch.zeros(n, 3, device=points.device, dtype=points.dtype)
    closest = torch.zeros(n, 3, device=points.device, dtype=points.dtype)
    m.unbatched_triangle_distance_forward_cuda(points, face_vertices, dist, sign, normals, closest)
    return dist, sign, normals, closest


def backward(grad_dist, points, closest):
    import torch
    g = torch.zeros_like(points)
    load().unbatched_triangle_distance_backward_cuda(grad_dist.contiguous(), points, closest, g)
    return g
