"""Drop-in `ProbabilisticGraspOptimizer` (reference: optimize_pregrasp.py:614-839) backed by libcdx.so.

Same constructor keywords, `.forward_kinematics(...)`, `.optimize(...)` signature and 5-tuple return as the reference
class.  The whole optimisation loop (float32 FK, GPIS queries, Kabsch equilibrium, cost terms, analytic gradient,
Adam, clamps, best-iterate tracking) runs in sm_100a kernels behind the C ABI of include/cdx.h with no host sync per
iteration.  Differences a caller can observe, all deliberate:
  * the per-iteration cost print (:738) and the final `print("Margin:")` (:838) only happen with verbose=True and
    never force a sync inside the loop;
  * the `1e-6 * torch.rand_like(H)` perturbation (:61) comes from a keyed stream (seed, iteration, copy, GLOBAL
    candidate index) so results do not depend on batch size or on how candidates are sharded over GPUs;
  * `opt_value` (best loss per candidate, :802) is kept as `self.opt_value`, and `argmin()` returns its arg-min.
There is no CPU path.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .hand import load_hand, tree_for

# optimize_pregrasp.py:13-30
EE_OFFSETS = [[0.0, -0.04, 0.015], [0.0, -0.04, 0.015], [0.0, -0.04, 0.015], [0.0, -0.05, -0.015]]
WRIST_OFFSET = np.array([[-0.06, 0.0, 0.05, 0.0, 0.0, 0.0],
                         [-0.04, 0.03, 0.05, 0.0, 0.0, -np.pi / 4],
                         [-0.01, 0.0, 0.05, 0.0, 0.0, np.pi / 4],
                         [0.1, 0.06, 0.03, -np.pi / 2, np.pi / 2, 0.0],
                         [-0.0, -0.06, 0.05, 0.0, 0.0, np.pi / 2],
                         [0.02, -0.04, 0.05, 0.0, 0.0, 3 * np.pi / 4]])
z_margin = 0.2
FINGERTIP_LB = [-0.2, -0.2, 0.015] * 4
FINGERTIP_UB = [0.2, 0.2, z_margin] * 4


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _np_ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class ProbabilisticGraspOptimizer:
    def __init__(self, robot_urdf, ee_link_names, ee_link_offsets=EE_OFFSETS, palm_offset=WRIST_OFFSET, num_iters=1000,
                 optimize_target=False, ref_q=None, tip_bounding_box=[FINGERTIP_LB, FINGERTIP_UB],
                 pregrasp_coefficients=[[0.8, 0.8, 0.8, 0.8], [0.8, 0.8, 0.8, 0.8], [0.8, 0.8, 0.8, 0.8]],
                 pregrasp_weights=[0.1, 0.8, 0.1], anchor_link_names=None, anchor_link_offsets=None, collision_pairs=None,
                 collision_pair_threshold=0.02, mass=0.1, com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0,
                 optimize_palm=False, device="cuda", noise_seed=0, env_offset=0):
        """robot_urdf: URDF path (as the reference), or 'leap' / 'allegro' / a hand JSON of data/hands."""
        self.device = torch.device(device)
        self._robot_urdf = robot_urdf
        self.tree = tree_for(robot_urdf, ee_link_names, ee_link_offsets)
        self.ref_q = torch.tensor(ref_q)                                         # float32 (:634)
        self.num_iters = num_iters
        self.ee_link_names = ee_link_names
        self.ee_link_offsets = ee_link_offsets
        self.palm_offset = torch.from_numpy(np.asarray(palm_offset)).double()   # [B,6] (:640)
        self.optimize_target = optimize_target
        self.optimize_palm = optimize_palm
        self.tip_bounding_box = [torch.tensor(tip_bounding_box[0]).view(-1, 3), torch.tensor(tip_bounding_box[1]).view(-1, 3)]
        self.pregrasp_coefficients = torch.tensor(pregrasp_coefficients)        # float32 (:644)
        self.pregrasp_weights = torch.tensor(pregrasp_weights).double()         # float32 values widened (:645)
        self.anchor_link_names = anchor_link_names
        self.anchor_link_offsets = anchor_link_offsets
        self.collision_pairs = collision_pairs
        self.collision_pair_threshold = collision_pair_threshold
        self.mass, self.com, self.gravity, self.uncertainty = mass, com, gravity, uncertainty
        self.noise_seed, self.env_offset = int(noise_seed), int(env_offset)
        self.opt_value = None
        self._ctx = None
        self._generation = None          # cdx_pregrasp_generation of the _init that made this object the owner of the ctx state
        self._tuning = None

    # ---- plumbing --------------------------------------------------------------------------------------
    def _dev_index(self):
        if self.device.type != "cuda":
            raise RuntimeError("compliancedex_b200.ProbabilisticGraspOptimizer runs on CUDA only (no CPU fallback)")
        return self.device.index if self.device.index is not None else torch.cuda.current_device()

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self._dev_index()).cuda_stream)

    def _bind(self, ctx):
        """use `ctx` (the GPIS object's context, which owns the fitted state) and make this optimizer's hand the one loaded
        in it.  Unconditional: the context keeps the loaded hand's identity, so re-loading the same hand is free and two
        optimizers with different hands sharing one GPIS can alternate."""
        self._ctx = ctx
        t = self.tree
        _lib.check(_lib.load().cdx_hand_load(ctx, t.n_bodies, _np_ptr(t.parent), _np_ptr(t.dof), _np_ptr(t.axis),
                                             _np_ptr(t.sign), _np_ptr(t.rot), _np_ptr(t.xyz), t.n_dofs,
                                             _np_ptr(t.tip_body), _np_ptr(t.tip_offset)))

    def _owned(self):
        """the pregrasp state is one per context: refuse to read or advance it after another optimizer re-initialised it"""
        if self._ctx is None or self._generation is None:
            raise RuntimeError("optimize() / _init() first")
        if _lib.load().cdx_pregrasp_generation(self._ctx, None) != self._generation:
            raise RuntimeError("another optimizer has re-initialised this GPIS context's pregrasp state; call optimize() again")

    def set_tuning(self, groups=0, occupancy=0, use_graph=True):
        """launch tuning of the device loop (cdx_pregrasp_set_tuning); results are bit-identical under every setting"""
        self._tuning = (int(groups), int(occupancy), int(bool(use_graph)))
        if self._ctx is not None:
            _lib.check(_lib.load().cdx_pregrasp_set_tuning(self._ctx, *self._tuning))

    def nan_count(self):
        """(candidate-iterations whose total_loss was NaN since optimize() started, first such iteration or -1): the
        reference's `print("NaN detected")` (:819) as a device-side record"""
        self._owned()
        n, first = ctypes.c_int64(0), ctypes.c_int(0)
        _lib.check(_lib.load().cdx_pregrasp_nan_count(self._ctx, ctypes.byref(n), ctypes.byref(first), self._stream()))
        return int(n.value), int(first.value)

    def _own_ctx(self):
        if self._ctx is None:
            self._bind(_lib.new_context(self._dev_index()))
        return self._ctx

    def _params(self, friction_mu):
        p = _lib.PregraspParams()
        coeffs = self.pregrasp_coefficients.numpy()
        P = coeffs.shape[0]
        if P > 4 or coeffs.shape[1] != 4:
            raise ValueError("pregrasp_coefficients must be [P<=4, 4]")
        p.num_copies = P
        w32 = self.pregrasp_weights.float().numpy()
        for a in range(P):
            for f in range(4):
                p.coeffs[a][f] = float(coeffs[a, f])
            p.weights[a] = float(w32[a])
        for i in range(self.tree.n_dofs):
            p.ref_q[i] = float(self.ref_q[i])
        lo, hi = self.tip_bounding_box[0].numpy(), self.tip_bounding_box[1].numpy()
        for f in range(4):
            for r in range(3):
                p.box_lo[f][r] = float(lo[f, r])
                p.box_hi[f][r] = float(hi[f, r])
        com32 = torch.zeros(3)                                                   # dummy_tip is float32 (:90-92)
        com32[:] = torch.tensor(self.com).double()
        for r in range(3):
            p.com[r] = float(com32[r])
        M = 2.0
        p.dummy_k = float((10.0 * self.mass / M * torch.ones(1))[0])             # gravity=10.0 (:726), float32 (:94)
        p.dummy_M = M
        p.gravity = 1 if self.gravity else 0
        p.uncertainty = float(self.uncertainty)
        p.optimize_target = int(bool(self.optimize_target))
        p.optimize_palm = int(bool(self.optimize_palm))
        p.cos_mu = float(torch.sqrt(1 / (1 + torch.tensor(friction_mu) ** 2)))   # exactly :111 / :707 (float32)
        p.noise_seed = self.noise_seed
        p.env_offset = self.env_offset
        for i, lr in enumerate([1e-3, 0.2, 2e-3, 1e-4, 1e-4]):                   # :782-796
            p.lr[i] = lr
        p.beta1, p.beta2, p.eps = 0.9, 0.999, 1e-8
        p.compliance_min = 80.0                                                  # :833
        p.track_after = 20                                                       # :823
        return p

    # ---- reference API ---------------------------------------------------------------------------------
    def forward_kinematics(self, joint_angles, palm_poses=None):
        """optimize_pregrasp.py:657-669: [B,n_dofs] (+ [B,6] palm poses) -> fingertip positions [B,4,3] float64."""
        if palm_poses is None:
            palm_poses = self.palm_offset
        ctx = self._own_ctx()
        dev = self.device
        q = joint_angles.detach().double().to(dev).contiguous()
        palm = palm_poses.detach().double().to(dev).contiguous()
        B = q.shape[0]
        tips = torch.empty(B, 4, 3, dtype=torch.float64, device=dev)
        _lib.check(_lib.load().cdx_hand_fk(ctx, _ptr(q), _ptr(palm), B, _ptr(tips), self._stream()))
        return tips

    def _init(self, init_joint_angles, target_pose, compliance, friction_mu, gpis):
        if getattr(gpis, "_ctx", None) is None:
            raise RuntimeError("gpis must be a fitted compliancedex_b200.GPIS (fit() or load_state_data() first)")
        self.device = torch.device("cuda", gpis._device)
        self._bind(gpis._ctx)
        dev = self.device
        q0 = init_joint_angles.detach().double().to(dev).contiguous()
        k0 = compliance.detach().double().to(dev).contiguous()
        T0 = target_pose.detach().double().to(dev).contiguous()
        palm0 = self.palm_offset.to(dev).contiguous()
        B = q0.shape[0]
        if palm0.shape[0] != B or k0.shape != (B, 4) or T0.shape != (B, 4, 3) or q0.shape[1] != self.tree.n_dofs:
            raise ValueError("shapes: joint_angles [B,n_dofs], target_pose [B,4,3], compliance [B,4], palm_offset [B,6]")
        p = self._params(friction_mu)
        lib = _lib.load()
        _lib.check(lib.cdx_pregrasp_set_tuning(self._ctx, *(self._tuning or (0, 0, 1))))
        _lib.check(lib.cdx_pregrasp_init(self._ctx, ctypes.byref(p), B, _ptr(q0), _ptr(k0), _ptr(T0), _ptr(palm0), self._stream()))
        self._generation = lib.cdx_pregrasp_generation(self._ctx, None)
        self._B = B
        return B

    def evaluate(self):
        """closure() at the current iterate without stepping: total_loss [B], margin [B,4] and the gradients
        autograd leaves in q / compliance / target / palm (.grad after total_loss.sum().backward(), :767-768)."""
        self._owned()
        B, dev, nd = self._B, self.device, self.tree.n_dofs
        out = [torch.empty(s, dtype=torch.float64, device=dev) for s in ((B,), (B, 4), (B, nd), (B, 4), (B, 4, 3), (B, 6))]
        _lib.check(_lib.load().cdx_pregrasp_eval(self._ctx, B, *[_ptr(t) for t in out], self._stream()))
        return out

    def step(self, n_iters):
        self._owned()
        _lib.check(_lib.load().cdx_pregrasp_run(self._ctx, int(n_iters), self._stream()))

    def current(self):
        self._owned()
        B, dev, nd = self._B, self.device, self.tree.n_dofs
        out = [torch.empty(s, dtype=torch.float64, device=dev) for s in ((B, nd), (B, 4), (B, 4, 3), (B, 6))]
        _lib.check(_lib.load().cdx_pregrasp_state(self._ctx, B, *[_ptr(t) for t in out], self._stream()))
        return out

    def results(self):
        self._owned()
        B, dev, nd = self._B, self.device, self.tree.n_dofs
        out = [torch.empty(s, dtype=torch.float64, device=dev) for s in ((B, nd), (B, 4), (B, 4, 3), (B, 6), (B, 4), (B,))]
        _lib.check(_lib.load().cdx_pregrasp_results(self._ctx, B, *[_ptr(t) for t in out], self._stream()))
        return out

    def optimize(self, init_joint_angles, target_pose, compliance, friction_mu, gpis, verbose=True):
        """optimize_pregrasp.py:771-839 -> (opt_joint_angle [B,n_dofs], opt_compliance [B,4], opt_target_pose [B,4,3],
        opt_palm_poses [B,6], opt_margin [B,4])."""
        self._init(init_joint_angles, target_pose, compliance, friction_mu, gpis)
        self.step(self.num_iters)
        oq, ok, oT, opalm, omargin, ovalue = self.results()
        self.opt_value = ovalue
        if verbose:
            print("Margin:", omargin)
        return oq, ok, oT, opalm, omargin

    def argmin(self):
        """(best opt_value, GLOBAL candidate index = env_offset + local index), lowest index on ties; the cross-rank
        reduction of these pairs is compliancedex_b200.distributed.global_argmin."""
        self._owned()
        out = torch.empty(2, dtype=torch.float64, device=self.device)
        self.argmin_into(out)
        v, i = out.tolist()
        return v, int(i)

    def argmin_into(self, out):
        """device-side arg-min into out[2] (float64, on this optimizer's device): no host sync"""
        _lib.check(_lib.load().cdx_pregrasp_argmin(self._ctx, _ptr(out), self._stream()))

    def compute_collision_loss(self, joint_angles, palm_poses=None):
        """optimize_pregrasp.py:671-701 (dormant on the prob path: its call at :765 is commented out, so it is not part
        of the fused step; provided for API parity).  float32 FK of the 4 anchor links -> world frame -> pair / floor /
        palm barriers.  joint_angles [B,16], palm_poses [B,6] (default: the constructor's palm_offset) -> cost [B]."""
        if self.anchor_link_names is None or self.collision_pairs is None:
            raise ValueError("compute_collision_loss needs anchor_link_names, anchor_link_offsets and collision_pairs")
        dev = torch.device("cuda", self._dev_index())
        lib = _lib.load()
        if getattr(self, "_anchor_ctx", None) is None:
            t = tree_for(self._robot_urdf, self.anchor_link_names, self.anchor_link_offsets)   # 4 anchors, like the shipped hands
            ctx = _lib.new_context(self._dev_index())
            _lib.check(lib.cdx_hand_load(ctx, t.n_bodies, _np_ptr(t.parent), _np_ptr(t.dof), _np_ptr(t.axis), _np_ptr(t.sign),
                                         _np_ptr(t.rot), _np_ptr(t.xyz), t.n_dofs, _np_ptr(t.tip_body), _np_ptr(t.tip_offset)))
            pairs = np.asarray(self.collision_pairs, dtype=np.int32).reshape(-1, 2)
            self._anchor_ctx = ctx
            self._pair_left = torch.from_numpy(np.ascontiguousarray(pairs[:, 0])).to(dev)
            self._pair_right = torch.from_numpy(np.ascontiguousarray(pairs[:, 1])).to(dev)
        if palm_poses is None:
            palm_poses = self.palm_offset
        q = joint_angles.detach().to(dev).double().contiguous()
        palm = palm_poses.detach().to(dev).double().contiguous()
        B = q.shape[0]
        anchors = torch.empty(B, 4, 3, dtype=torch.float64, device=dev)
        cost = torch.empty(B, dtype=torch.float64, device=dev)
        _lib.check(lib.cdx_hand_fk(self._anchor_ctx, _ptr(q), _ptr(palm), B, _ptr(anchors), self._stream()))
        _lib.check(lib.cdx_collision_cost(self._anchor_ctx, _ptr(anchors), B, 4, _ptr(self._pair_left), _ptr(self._pair_right),
                                          int(self._pair_left.numel()), float(self.collision_pair_threshold),
                                          _ptr(palm) if self.optimize_palm else None, _ptr(cost), self._stream()))
        return cost


class KinGPISGraspOptimizer(ProbabilisticGraspOptimizer):
    """Drop-in `KinGPISGraspOptimizer` (optimize_pregrasp.py:408-511): the fused step kernel with variant 1 -
    recursive FK on float64 angles + a constant palm offset [3], one copy (the tips themselves are the contact points),
    loss = -5 reward + 1000 sum|dist| + 10 sum tar_dist + 10 |centre diff| + force_cost(max 1) + 20 |q - ref_q|
    + max_f uncertainty log(100 std), torch.optim.RMSprop, compliance >= 40, best-iterate tracking from iteration 0,
    `flag = (opt_margin > 0).all()` with opt_margin the whole batch's margin at the last improving iteration."""
    _variant = 1

    def __init__(self, robot_urdf, ee_link_names, ee_link_offsets=EE_OFFSETS, palm_offset=WRIST_OFFSET, num_iters=1000,
                 optimize_target=False, ref_q=None, tip_bounding_box=[FINGERTIP_LB, FINGERTIP_UB], mass=0.1,
                 com=[0.0, 0.0, 0.0], gravity=True, uncertainty=10.0, device="cuda", noise_seed=0, env_offset=0):
        palm = np.asarray(palm_offset, dtype=np.float64).reshape(-1)
        if palm.size != 3:
            raise ValueError("KinGPISGraspOptimizer adds palm_offset to [B*4,3] tip positions (:437): it must have 3 entries")
        super().__init__(robot_urdf, ee_link_names, ee_link_offsets, palm_offset=np.zeros((1, 6)), num_iters=num_iters,
                         optimize_target=optimize_target, ref_q=ref_q, tip_bounding_box=tip_bounding_box,
                         pregrasp_coefficients=[[1.0, 1.0, 1.0, 1.0]], pregrasp_weights=[1.0], mass=mass, com=com,
                         gravity=gravity, uncertainty=uncertainty, optimize_palm=False, device=device,
                         noise_seed=noise_seed, env_offset=env_offset)
        self.palm_offset = torch.from_numpy(palm)                                # float64 [3] (:423)
        self.opt_margin = None

    def _params(self, friction_mu):
        p = super()._params(friction_mu)
        p.variant = self._variant
        lrs = [2e-3, 0.2, 1e-3] if self.optimize_target else [1e-2, 0.2, 0.0]    # :454-459
        for i, lr in enumerate(lrs + [0.0, 0.0]):
            p.lr[i] = lr
        p.rms_alpha, p.eps = 0.99, 1e-8                                          # torch.optim.RMSprop defaults
        p.compliance_min = 40.0                                                  # :506
        p.track_after = -1                                                       # :496: every iteration
        return p

    def _palm6(self, B):
        palm = torch.zeros(B, 6, dtype=torch.float64)
        palm[:, :3] = self.palm_offset
        return palm

    def forward_kinematics(self, joint_angles):
        """optimize_pregrasp.py:431-438: [B,n_dofs] -> fingertip positions [B*4,3] float64."""
        ctx = self._own_ctx()
        dev = self.device
        q = joint_angles.detach().double().to(dev).contiguous()
        B = q.shape[0]
        palm = self._palm6(B).to(dev)
        tips = torch.empty(B * 4, 3, dtype=torch.float64, device=dev)
        _lib.check(_lib.load().cdx_hand_fk_recursive(ctx, _ptr(q), _ptr(palm), B, _ptr(tips), self._stream()))
        return tips

    def _init(self, joint_angles, target_pose, compliance, friction_mu, gpis):
        self.palm6 = self._palm6(joint_angles.shape[0])
        saved, self.palm_offset = self.palm_offset, self.palm6
        try:
            return super()._init(joint_angles, target_pose, compliance, friction_mu, gpis)
        finally:
            self.palm_offset = saved

    def latched_margin(self):
        """(opt_margin [B,4], iteration at which it was taken; -1 = never)"""
        m = torch.empty(self._B, 4, dtype=torch.float64, device=self.device)
        it = torch.empty(1, dtype=torch.int32, device=self.device)
        self._owned()
        _lib.check(_lib.load().cdx_pregrasp_latched_margin(self._ctx, self._B, _ptr(m), _ptr(it), self._stream()))
        return m, it

    def optimize(self, joint_angles, target_pose, compliance, friction_mu, gpis, verbose=True):
        """optimize_pregrasp.py:440-511 -> (opt_joint_angle [B,n_dofs], opt_compliance [B,4], opt_target_pose [B,4,3], flag)."""
        self._init(joint_angles, target_pose, compliance, friction_mu, gpis)
        self.step(self.num_iters)
        oq, ok, oT, _, _, ovalue = self.results()
        self.opt_value = ovalue
        self.opt_margin, _ = self.latched_margin()
        if verbose:
            print(self.opt_margin)
        flag = (self.opt_margin > 0.0).all()
        return oq, ok, oT, flag


class _NoHand:
    n_dofs = 12


class GPISGraspOptimizer(KinGPISGraspOptimizer):
    """Drop-in `GPISGraspOptimizer` (optimize_pregrasp.py:322-406): the fingertip positions [B,4,3] themselves are optimised
    (no hand, no FK): variant 2 of the fused step - reward x25, variance cost = uncertainty * sum_f std, no ref cost,
    RMSprop 1e-3 / 1e-3 / 0.2, tips clamped to the bounding box, compliance >= 40, target NOT clamped (:402)."""
    _variant = 2

    def __init__(self, tip_bounding_box, num_iters=2000, optimize_target=False, mass=0.1, com=[0.0, 0.0, 0.0], gravity=True,
                 uncertainty=20.0, device="cuda", noise_seed=0, env_offset=0):
        self.device = torch.device(device)
        self.tree = _NoHand()
        self.ref_q = torch.zeros(12)
        self.num_iters = num_iters
        self.palm_offset = torch.zeros(3, dtype=torch.float64)
        self.optimize_target, self.optimize_palm = optimize_target, False
        self.tip_bounding_box = [torch.tensor(tip_bounding_box[0]).view(-1, 3), torch.tensor(tip_bounding_box[1]).view(-1, 3)]
        self.pregrasp_coefficients = torch.tensor([[1.0, 1.0, 1.0, 1.0]])
        self.pregrasp_weights = torch.tensor([1.0]).double()
        self.mass, self.com, self.gravity, self.uncertainty = mass, com, gravity, uncertainty
        self.noise_seed, self.env_offset = int(noise_seed), int(env_offset)
        self.opt_value = self.opt_margin = None
        self._ctx = None
        self._generation = None
        self._tuning = None

    def _bind(self, ctx):
        self._ctx = ctx                                                          # no hand to load

    def _params(self, friction_mu):
        p = super()._params(friction_mu)
        lrs = [1e-3, 0.2, 1e-3 if self.optimize_target else 0.0, 0.0, 0.0]       # :348-353
        for i, lr in enumerate(lrs):
            p.lr[i] = lr
        return p

    def forward_kinematics(self, tip_pose):
        return tip_pose.view(-1, 3)

    def optimize(self, tip_pose, target_pose, compliance, friction_mu, gpis, verbose=True):
        """optimize_pregrasp.py:335-406 -> (opt_tip_pose [B,4,3], opt_compliance [B,4], opt_target_pose [B,4,3], flag)."""
        B = tip_pose.shape[0]
        otip, ok, oT, flag = super().optimize(tip_pose.reshape(B, 12), target_pose, compliance, friction_mu, gpis, verbose)
        return otip.view(B, 4, 3), ok, oT, flag


# optimize_pregrasp.py:842-848; "sdf" / "kin" (mesh SDF) and "wc" (cvxpy conic solve per step) are out of scope
optimizers = {"prob": ProbabilisticGraspOptimizer, "kingpis": KinGPISGraspOptimizer, "gpis": GPISGraspOptimizer}
