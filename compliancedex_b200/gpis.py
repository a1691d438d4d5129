"""Drop-in `GPIS` (reference: gpis.py:4-168) backed by libcdx.so.

Same constructor, method names, argument meaning, tensor shapes / dtypes and state layout as the reference
class; the arithmetic runs in hand-written sm_100a kernels behind the C ABI of include/cdx.h:
  fit / load_state_data  -> cdx_gpis_fit / cdx_gpis_set_state (assembly + blocked Cholesky + inverses, once)
  pred                   -> cdx_gpis_query (mean, std and - when X2 requires grad - both spatial gradients)
  compute_normal         -> cdx_gpis_query (grad of the mean), normalised as gpis.py:83
There is no CPU path: tensors must live on a CUDA device.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib

STATE_DIR = "gpis_states"  # gpis.py:160,163 hard-code this relative directory


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _PredFn(torch.autograd.Function):
    """pred's outputs stay differentiable w.r.t. X2 as in the reference (autograd through gpis.py:51-59)."""

    @staticmethod
    def forward(ctx, X2, gpis):
        M = X2.shape[0]
        mean = torch.empty(M, dtype=torch.float64, device=X2.device)
        std = torch.empty(M, dtype=torch.float64, device=X2.device)
        need_grad = bool(ctx.needs_input_grad[0])
        gm = torch.empty(M, 3, dtype=torch.float64, device=X2.device) if need_grad else None
        gs = torch.empty(M, 3, dtype=torch.float64, device=X2.device) if need_grad else None
        gpis._query(X2, mean, gm, std, gs)
        if need_grad:
            ctx.save_for_backward(gm, gs)
        ctx.has = need_grad
        return mean, std

    @staticmethod
    def backward(ctx, g_mean, g_std):
        if not ctx.has:
            return None, None
        gm, gs = ctx.saved_tensors
        out = torch.zeros_like(gm)
        if g_mean is not None:
            out = out + g_mean.unsqueeze(1) * gm
        if g_std is not None:
            out = out + g_std.unsqueeze(1) * gs
        return out, None


class GPIS:
    def __init__(self, sigma=0.6, bias=2, kernel="tps"):
        # gpis.py:5-14
        if kernel not in _lib.KERNEL_IDS:
            raise ValueError(f"unknown kernel {kernel!r}")
        self.sigma = sigma
        self.bias = bias
        self.fraction = None
        self.kernel = kernel
        self.kernel_function = getattr(self, {"tps": "thin_plate_spline", "rbf": "exponentiated_quadratic",
                                              "joint": "joint_kernel"}[kernel])
        self._ctx = None
        self._device = None

    # ---- covariance functions: API parity only (plain torch, never on the query path) ------------------
    def exponentiated_quadratic(self, xa, xb):
        return torch.exp(-0.5 * torch.cdist(xa, xb) ** 2 / self.sigma ** 2)          # gpis.py:16-19

    def thin_plate_spline(self, xa, xb, R=None):
        R = self.R if R is None else R
        d = torch.cdist(xa, xb)
        return 2 * d ** 3 - 3 * R * d ** 2 + R ** 3                                    # gpis.py:21-26

    def joint_kernel(self, xa, xb):
        return 0.3 * self.exponentiated_quadratic(xa, xb) + 0.7 * self.thin_plate_spline(xa, xb)  # gpis.py:28-29

    # ---- helpers ---------------------------------------------------------------------------------------
    def _ensure_ctx(self, device):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("compliancedex_b200.GPIS runs on CUDA tensors only (no CPU fallback)")
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._ctx is None or self._device != idx:
            if self._ctx is not None:
                self._ctx.close()                     # moving to another device: release the old device's state now
            self._ctx = _lib.new_context(idx)
            self._device = idx
            _lib.check(_lib.load().cdx_gpis_set_var_mode(self._ctx, getattr(self, "_var_mode", 0)))
            if hasattr(self, "_ext_mode"):
                _lib.check(_lib.load().cdx_gpis_set_ext_mode(self._ctx, self._ext_mode))
            if hasattr(self, "_i8_variant"):
                _lib.check(_lib.load().cdx_gpis_set_i8_variant(self._ctx, self._i8_variant))
        return self._ctx

    def _check_factorisation(self):
        lib = _lib.load()
        n, r, b, p, col = ctypes.c_int64(), ctypes.c_double(), ctypes.c_double(), ctypes.c_double(), ctypes.c_int()
        _lib.check(lib.cdx_gpis_info(self._ctx, ctypes.byref(n), ctypes.byref(r), ctypes.byref(b), ctypes.byref(p),
                                     ctypes.byref(col), _stream(self._device)))
        self._prior = p.value
        return r.value

    def _query(self, X, mean, gmean, std, gstd):
        if self._ctx is None:
            raise RuntimeError("GPIS: fit() or load_state_data() first")
        X = X.contiguous()
        _lib.check(_lib.load().cdx_gpis_query(self._ctx, _ptr(X), X.shape[0], _ptr(mean), _ptr(gmean), _ptr(std),
                                              _ptr(gstd), _stream(self._device)))

    def set_var_mode(self, mode):
        """-1: std always by FP64 tensor MMA; 0: auto (tcgen05 INT8 slices for TPS states with padded N in 1024..8192);
        4..6: INT8 path with that many byte slices whenever eligible (include/cdx.h cdx_gpis_set_var_mode)."""
        self._var_mode = int(mode)
        if self._ctx is not None:
            _lib.check(_lib.load().cdx_gpis_set_var_mode(self._ctx, self._var_mode))

    def set_ext_mode(self, mode):
        """kernel of the extended query (mean, std and both gradients) on states of up to 448 padded points: 0 warp-specialised
        all-columns kernel (default), 2 its single-role predecessor, 1 the column-tiled kernel that larger states use
        (include/cdx.h cdx_gpis_set_ext_mode)."""
        self._ext_mode = int(mode)
        if self._ctx is not None:
            _lib.check(_lib.load().cdx_gpis_set_ext_mode(self._ctx, self._ext_mode))

    def set_i8_variant(self, variant):
        """launch shape of the tcgen05 INT8 std path (include/cdx.h cdx_gpis_set_i8_variant): 3 CTA pair with
        tcgen05.mma.cta_group::2 (default), 1 two-CTA cluster exchanging k* over distributed shared memory, 2 no cluster."""
        self._i8_variant = int(variant)
        if self._ctx is not None:
            _lib.check(_lib.load().cdx_gpis_set_i8_variant(self._ctx, self._i8_variant))

    # ---- fit / state -----------------------------------------------------------------------------------
    def fit(self, X1, y1, noise=0.0):
        """gpis.py:33-40.  X1 [N,3], y1 [N,1] float64 CUDA; noise scalar or [N]."""
        ctx = self._ensure_ctx(X1.device)
        X1c = X1.detach().double().contiguous()
        y1c = y1.detach().double().reshape(-1).contiguous()
        N = X1c.shape[0]
        if torch.is_tensor(noise) and noise.numel() > 1:
            nz = noise.detach().double().to(X1c.device).reshape(-1).contiguous()
            assert nz.numel() == N
            nptr, nscalar = _ptr(nz), 0.0
        else:
            nz, nptr, nscalar = None, None, float(noise)
        _lib.check(_lib.load().cdx_gpis_fit(ctx, _ptr(X1c), _ptr(y1c), nptr, nscalar, N, float(self.bias),
                                            _lib.KERNEL_IDS[self.kernel], float(self.sigma), _stream(self._device)))
        R = self._check_factorisation()
        self.X1 = X1
        self.y1 = y1 - self.bias
        self.noise = noise
        if self.kernel in ("tps", "joint"):
            self.R = torch.tensor(R, dtype=torch.float64, device=X1.device)
        E = torch.empty(N, N, dtype=torch.float64, device=X1.device)
        _lib.check(_lib.load().cdx_gpis_get(ctx, _lib.MAT_E11, _ptr(E), _stream(self._device)))
        self.E11 = E

    def _adopt(self, R, X1, y1, E11, bias):
        ctx = self._ensure_ctx(X1.device)
        self.R, self.X1, self.y1, self.E11, self.bias = R, X1, y1, E11, bias
        _lib.check(_lib.load().cdx_gpis_set_state(
            ctx, _ptr(X1.contiguous()), _ptr(y1.reshape(-1).contiguous()), _ptr(E11.contiguous()), float(R),
            float(bias), X1.shape[0], _lib.KERNEL_IDS[self.kernel], float(self.sigma), _stream(self._device)))
        self._check_factorisation()

    def save_state_data(self, name="gpis_state"):
        # gpis.py:155-160 - frozen layout: R, X1, y1 (bias-subtracted), E11, bias
        np.savez(f"{STATE_DIR}/{name}.npz", R=self.R.cpu().numpy(), X1=self.X1.cpu().numpy(), y1=self.y1.cpu().numpy(),
                 E11=self.E11.cpu().numpy(), bias=self.bias if not torch.is_tensor(self.bias) else self.bias.cpu().numpy())

    def load_state_data(self, name="gpis_state", device="cuda"):
        # gpis.py:162-168
        path = name if os.path.isfile(str(name)) else f"{STATE_DIR}/{name}.npz"
        data = np.load(path)
        dev = torch.device(device)
        self._adopt(torch.from_numpy(data["R"]).to(dev), torch.from_numpy(data["X1"]).to(dev),
                    torch.from_numpy(data["y1"]).to(dev), torch.from_numpy(data["E11"]).to(dev),
                    torch.from_numpy(np.asarray(data["bias"])).double().to(dev))

    # ---- posterior -------------------------------------------------------------------------------------
    def pred(self, X2):
        """gpis.py:43-59: X2 [M,3] or [B,F,3] -> (mean, std) shaped like X2 without the last dim."""
        shape = list(X2.shape[:-1])
        mean, std = _PredFn.apply(X2.reshape(-1, 3).double(), self)
        return mean.view(shape), std.view(shape)

    def pred_with_grads(self, X2):
        """mean, std, d mean/dX2, d std/dX2 in one fused call (no autograd graph)."""
        X = X2.detach().reshape(-1, 3).double().contiguous()
        M = X.shape[0]
        mean = torch.empty(M, dtype=torch.float64, device=X.device)
        std = torch.empty_like(mean)
        gm = torch.empty(M, 3, dtype=torch.float64, device=X.device)
        gs = torch.empty_like(gm)
        self._query(X, mean, gm, std, gs)
        return mean, std, gm, gs

    def compute_normal(self, X2, index=None):
        """gpis.py:63-87: unit gradient of the posterior mean (detached); `index` = training-row subset."""
        shape = X2.shape
        X = X2.detach().reshape(-1, 3).double().contiguous()
        g = torch.empty(X.shape[0], 3, dtype=torch.float64, device=X.device)
        if index is None:
            self._query(X, None, g, None, None)
        else:
            N = self.X1.shape[0]
            alpha = torch.empty(N, dtype=torch.float64, device=X.device)
            _lib.check(_lib.load().cdx_gpis_get(self._ctx, _lib.VEC_ALPHA, _ptr(alpha), _stream(self._device)))
            idx = torch.as_tensor(index, device=X.device, dtype=torch.long)
            w = torch.zeros_like(alpha)
            w[idx] = alpha[idx]
            _lib.check(_lib.load().cdx_gpis_query_weights(self._ctx, _ptr(w), _ptr(X), X.shape[0], None, _ptr(g),
                                                          _stream(self._device)))
            weight_sum = alpha[idx].sum()
        n = g / (torch.norm(g, dim=1, keepdim=True) + 1e-8)
        if index is None:
            return n.view(shape)
        return n.view(shape), weight_sum

    def _nested_row_subsets(self, S):
        """The training-row subsets of gpis.py:96-104, cached on first use like the reference (`self.fraction`,
        `self.indices`): S // 2 prefixes [0, int(f N)), f = linspace(0.8, 1, S // 2); all rows; S // 2 ranges
        [int(1 - f), N) - which start at 0 for every f in (0, 1], i.e. all rows again (the reference's arithmetic, kept)."""
        if self.fraction is None:
            N, h = len(self.X1), S // 2
            self.fraction = torch.linspace(0.8, 1, h)
            stops = [int(self.fraction[i] * N) for i in range(h)]
            starts = [int(1 - self.fraction[h - 1 - i]) for i in range(h)]
            self.indices = [list(range(e)) for e in stops] + [list(range(N))] + [list(range(b, N)) for b in starts]
        return self.indices

    def compute_multinormals(self, X2, num_normal_samples):
        """gpis.py:89-111: X2 [M,3] -> (normals [M,S,3], weights [S]); sample i is compute_normal on the i-th nested
        row subset, the weights are the subsets' alpha sums normalised to 1."""
        subsets = self._nested_row_subsets(num_normal_samples)
        per_subset = [self.compute_normal(X2, subsets[i]) for i in range(num_normal_samples)]
        w = torch.hstack([w_i for _, w_i in per_subset])
        return torch.stack([n_i for n_i, _ in per_subset], dim=1), w / w.sum()

    def get_visualization_data(self, lb, ub, steps=100):
        """gpis.py:113-125: the grid driver; one fused query over all steps^3 points instead of `steps` slabs."""
        dev = self.X1.device
        # the three float32 axis vectors come from the host's linspace (the reference's values bit for bit); only they
        # cross the bus (3*steps floats), the steps^3 x 3 grid is expanded on the device
        axes = torch.stack([torch.linspace(lb[a], ub[a], steps) for a in range(3)]).to(dev)
        X = torch.stack(torch.meshgrid(axes[0], axes[1], axes[2], indexing="xy"), dim=3).double().reshape(-1, 3)
        M = X.shape[0]
        mean = torch.empty(M, dtype=torch.float64, device=dev)
        std = torch.empty_like(mean)
        g = torch.empty(M, 3, dtype=torch.float64, device=dev)
        self._query(X, mean, g, std, None)
        # float32 results packed as planes [mean | std | normals] -> ONE device-to-host copy into pinned memory; the
        # returned arrays are contiguous views of that buffer (it lives as long as any of them does)
        packed = torch.empty(5 * M, dtype=torch.float32, device=dev)
        packed[:M].copy_(mean)
        packed[M:2 * M].copy_(std)
        packed[2 * M:].view(M, 3).copy_(g / (torch.norm(g, dim=1, keepdim=True) + 1e-8))
        host = torch.empty(5 * M, dtype=torch.float32, pin_memory=True)
        host.copy_(packed, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        out = host.numpy()
        return (out[:M].reshape(steps, steps, steps), out[M:2 * M].reshape(steps, steps, steps),
                out[2 * M:].reshape(steps, steps, steps, 3), np.asarray(lb), np.asarray(ub))

    def topcd(self, test_mean, test_normal, lb, ub, steps=100):
        """gpis.py:127-150: surface point cloud of a mean grid - voxels with mean < 0 that have a 6-neighbour with
        mean >= 0 (interior voxels only), in C index order, with their normals.  Like the reference, TENSOR inputs are
        flipped along the last axis first (`[:,:,::-1]`), numpy inputs are used as they are.  The stencil and the ordered
        compaction run on the GPU (csrc/cdx_grid.cu); the coordinates come from numpy's linspace so that they are the
        reference's bit for bit."""
        lb, ub = np.asarray(lb), np.asarray(ub)
        if torch.is_tensor(test_mean):
            test_mean = test_mean.cpu().numpy()[:, :, ::-1]
        if torch.is_tensor(test_normal):
            test_normal = test_normal.cpu().numpy()[:, :, ::-1]
        if self._ctx is None:
            self._ensure_ctx(self.X1.device if hasattr(self, "X1") and torch.is_tensor(self.X1) else "cuda")
        dev = torch.device("cuda", self._device)
        m = torch.from_numpy(np.ascontiguousarray(test_mean, dtype=np.float32)).to(dev)
        if tuple(m.shape) != (steps, steps, steps):
            raise ValueError(f"topcd: test_mean must be [{steps},{steps},{steps}], got {tuple(m.shape)}")
        idx = torch.empty(steps ** 3, dtype=torch.int64, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        _lib.check(_lib.load().cdx_grid_surface(self._ctx, _ptr(m), int(steps), _ptr(idx), _ptr(cnt), _stream(self._device)))
        lin = idx[:int(cnt.item())].cpu().numpy()
        i, j, k = lin // (steps * steps), (lin // steps) % steps, lin % steps
        xs, ys, zs = (np.linspace(lb[a], ub[a], steps) for a in range(3))
        points = np.stack([xs[j], ys[i], zs[k]], axis=1)          # meshgrid(indexing="xy"): [i,j,k] -> (x[j], y[i], z[k])
        normals = np.asarray(test_normal)[i, j, k]
        return points, normals

    def save_visualization_data(self, name, lb, ub, steps=100):
        """what train_gpis_completion.py:60-61 / gpis.py:264-266 write next to the state: gpis_states/{name}.npz with
        float32 grids mean, var (the std, as the reference names it), normal and the box ub, lb."""
        mean, std, normal, lb, ub = self.get_visualization_data(lb, ub, steps)
        np.savez(f"gpis_states/{name}.npz" if not str(name).endswith(".npz") else name, mean=mean, var=std, normal=normal,
                 ub=ub, lb=lb)
        return mean, std, normal, lb, ub
