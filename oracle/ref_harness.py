"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

Imports the *unmodified* reference (Ericcsr/ComplianceDex, mounted read-only at /root/reference) on
CPU so that golden vectors can be generated from the reference's own code (see make_golden.py).
In the build container the reference is /root/reference; on the GPU box it is the git-ignored unmodified copy
oracle/_ref/reference_py (oracle/build_ref_py.py).  Only goldens generation and bench.py's reference arm /
cpu_baseline leg use this module - always on the HOST cores (CUDA is hidden from the process first).

What is stubbed (none of it is arithmetic on the hot path):
  * open3d, matplotlib, torchsdf, cvxpy, cvxpylayers, pybullet_robot: imported at module top of
    optimize_pregrasp.py:1-10 / math_utils.py:1-4, absent here, unused by the `prob` path.
  * urdf_parser_py: parsing only (urdf_utils.py:9,14); replaced by a stdlib-XML stand-in exposing the
    attributes urdf_utils.py:28-126 reads.
  * torch.jit.script -> identity and Tensor.cuda -> no-op (CPU-only container);
    DifferentiableRobotModel is forced to device="cpu" (optimize_pregrasp.py:635 hard-codes cuda:0).
  * torch.rand_like inside optimal_transformation_batch (optimize_pregrasp.py:61) is replaced by the
    keyed, counter-based stream `oracle.noise.kabsch_noise` so that the oracle, the reference and the
    CUDA kernel all see the same perturbation.
"""
import contextlib
import importlib.util
import io
import os
import sys
import types
import xml.etree.ElementTree as ET

import numpy as np
import torch

def _reference_root():
    """/root/reference in the build container; on the GPU box the git-ignored unmodified copy of the path's files that
    oracle/build_ref_py.py made (oracle/_ref/reference_py travels with the snapshot like oracle/_ref/torchsdf_ref.so)"""
    for cand in (os.environ.get("CDX_REFERENCE_ROOT"), "/root/reference",
                 os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference_py")):
        if cand and os.path.isfile(os.path.join(cand, "gpis.py")):
            return cand
    return "/root/reference"


REF = _reference_root()
LEAP_URDF = f"{REF}/pybullet_robot/src/pybullet_robot/robots/leap_hand/assets/leap_hand/robot.urdf"
ALLEGRO_URDF = f"{REF}/pybullet_robot/src/pybullet_robot/robots/allegro_hand/models/allegro_hand_description_left.urdf"

_state = {"installed": False}


def available():
    return os.path.isdir(REF) and os.path.isfile(f"{REF}/gpis.py")


# ------------------------------------------------------------------ urdf_parser_py stand-in
class _NS(types.SimpleNamespace):
    pass


def _floats(s, default):
    if s is None:
        return list(default)
    return [float(v) for v in s.split()]


class _URDF:
    def __init__(self, links, joints):
        self.links, self.joints = links, joints

    @staticmethod
    def from_xml_file(path):
        root = ET.parse(path).getroot()
        links = []
        for l in root.findall("link"):
            links.append(_NS(name=l.get("name"), inertial=None))
        joints = []
        for j in root.findall("joint"):
            o = j.find("origin")
            origin = _NS(position=_floats(o.get("xyz") if o is not None else None, [0, 0, 0]),
                         rotation=_floats(o.get("rpy") if o is not None else None, [0, 0, 0]))
            ax = j.find("axis")
            lim = j.find("limit")
            limit = None
            if lim is not None:
                limit = _NS(effort=float(lim.get("effort", 0)), lower=float(lim.get("lower", 0)),
                            upper=float(lim.get("upper", 0)), velocity=float(lim.get("velocity", 0)))
            joints.append(_NS(name=j.get("name"), type=j.get("type"),
                              parent=j.find("parent").get("link"), child=j.find("child").get("link"),
                              origin=origin,
                              axis=_floats(ax.get("xyz") if ax is not None else None, [1, 0, 0]),
                              limit=limit, dynamics=_NS(damping=0.0)))
        return _URDF(links, joints)


def install():
    """Make `import gpis`, `import optimize_pregrasp` work against /root/reference on CPU."""
    if _state["installed"]:
        return
    if not available():
        raise RuntimeError(f"reference tree not found at {REF} (only present in the build container)")
    for name in ["open3d", "matplotlib", "matplotlib.pyplot", "torchsdf", "pybullet_robot",
                 "pybullet_robot.robots", "cvxpy", "cvxpylayers", "cvxpylayers.torch"]:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["torchsdf"].compute_sdf = None
    sys.modules["pybullet_robot.robots"].robot_configs = {}
    sys.modules["cvxpylayers.torch"].CvxpyLayer = object
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    up = types.ModuleType("urdf_parser_py")
    upu = types.ModuleType("urdf_parser_py.urdf")
    upu.URDF = _URDF
    up.urdf = upu
    sys.modules["urdf_parser_py"] = up
    sys.modules["urdf_parser_py.urdf"] = upu

    torch.jit.script = lambda f, *a, **k: f
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
    sys.path.insert(0, REF)
    sys.path.insert(0, f"{REF}/thirdparty/differentiable-robot-model")
    with contextlib.redirect_stdout(io.StringIO()):
        import differentiable_robot_model.robot_model as rm
    if not torch.cuda.is_available():
        Base = rm.DifferentiableRobotModel

        class CpuModel(Base):
            def __init__(self, urdf_path, name="", device=None):
                super().__init__(urdf_path, name, device="cpu")

        rm.DifferentiableRobotModel = CpuModel
    _state["installed"] = True


def hand_config(hand):
    """The reference's ROBOT_CONFIG dict (leap_hand_config.py:3-27 / allegro_hand_config.py:3-27)."""
    sub = {"leap": "leap_hand/leap_hand_config.py", "allegro": "allegro_hand/allegro_hand_config.py"}[hand]
    path = f"{REF}/pybullet_robot/src/pybullet_robot/robots/{sub}"
    spec = importlib.util.spec_from_file_location(f"_cdx_{hand}_config", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ROBOT_CONFIG


def hand_urdf(hand):
    return {"leap": LEAP_URDF, "allegro": ALLEGRO_URDF}[hand]


def load_gpis(state_name, bias=1.0, device=None):
    """Reference GPIS object with the state assigned as gpis.py:162-168 does; `.cuda()` only when a device is given (the
    reference-on-GPU parity run of scripts/scale_parity.py --ref-gpu, where CUDA is visible and nothing is patched to the host)."""
    install()
    from gpis import GPIS
    g = GPIS(0.08, bias)
    d = np.load(f"{REF}/gpis_states/{state_name}.npz")
    g.R = torch.from_numpy(d["R"])
    g.X1 = torch.from_numpy(d["X1"])
    g.y1 = torch.from_numpy(d["y1"])
    g.E11 = torch.from_numpy(d["E11"])
    g.bias = torch.from_numpy(d["bias"]).double()
    if device is not None:
        g.R, g.X1, g.y1, g.E11, g.bias = (t.to(device) for t in (g.R, g.X1, g.y1, g.E11, g.bias))
    return g


def gpis_from_arrays(X1, y1, E11, R, bias):
    install()
    from gpis import GPIS
    g = GPIS(0.08, float(bias))
    g.R = torch.as_tensor(R, dtype=torch.float64)
    g.X1 = torch.as_tensor(X1, dtype=torch.float64)
    g.y1 = torch.as_tensor(y1, dtype=torch.float64)
    g.E11 = torch.as_tensor(E11, dtype=torch.float64)
    g.bias = torch.as_tensor(float(bias), dtype=torch.float64)
    return g


def ref_pred_chunked(g, X2, chunk=2048, with_grad=False):
    """Reference pred in chunks (bit-exact w.r.t. one call) - optional autograd grads of
    mean.sum() and std.sum() w.r.t. X2 (what the optimizers back-propagate, optimize_pregrasp.py:733-736)."""
    means, stds, gms, gss = [], [], [], []
    for s in range(0, X2.shape[0], chunk):
        x = X2[s:s + chunk].clone().requires_grad_(with_grad)
        m, sd = g.pred(x)
        m = m.reshape(-1)
        sd = sd.reshape(-1)
        if with_grad:
            gm, = torch.autograd.grad(m.sum(), x, retain_graph=True)
            gs, = torch.autograd.grad(sd.sum(), x)
            gms.append(gm)
            gss.append(gs)
        means.append(m.detach())
        stds.append(sd.detach())
    out = [torch.cat(means), torch.cat(stds)]
    if with_grad:
        out += [torch.cat(gms), torch.cat(gss)]
    return out


@contextlib.contextmanager
def pinned_kabsch_noise(seed, num_envs_offset=0, num_copies=3):
    """Replace torch.rand_like by the keyed stream of oracle/noise.py for the duration of a reference
    run.  The reference calls rand_like exactly once per closure (optimize_pregrasp.py:61) on a
    [3B,3,3] tensor; call k of the context = optimizer iteration k."""
    from oracle.noise import kabsch_noise
    orig = torch.rand_like
    counter = {"it": 0}

    def fake(t, *a, **k):
        E = t.shape[0]
        out = kabsch_noise(seed, counter["it"], E, num_envs_offset, num_copies)
        counter["it"] += 1
        return out.to(device=t.device, dtype=t.dtype)

    torch.rand_like = fake
    try:
        yield counter
    finally:
        torch.rand_like = orig


def make_prob_optimizer(hand, palm_offset, num_iters, mass=0.1, com=(0.0, 0.0, 0.0), gravity=True,
                        uncertainty=20.0, optimize_palm=True, optimize_target=True, **kw):
    """A fresh reference ProbabilisticGraspOptimizer (fresh robot model: SURVEY fact 6)."""
    install()
    with contextlib.redirect_stdout(io.StringIO()):
        import optimize_pregrasp as op
        cfg = hand_config(hand)
        opt = op.ProbabilisticGraspOptimizer(
            hand_urdf(hand),
            ee_link_names=cfg["ee_link_name"],
            ee_link_offsets=cfg["ee_link_offset"].tolist(),
            anchor_link_names=cfg["collision_links"],
            anchor_link_offsets=cfg["collision_offsets"].tolist(),
            collision_pairs=cfg["collision_pairs"],
            tip_bounding_box=[op.FINGERTIP_LB, op.FINGERTIP_UB],
            ref_q=cfg["ref_q"].tolist(),
            optimize_target=optimize_target, optimize_palm=optimize_palm,
            num_iters=num_iters, palm_offset=np.asarray(palm_offset, dtype=np.float64),
            mass=mass, com=list(com), gravity=gravity, uncertainty=uncertainty, **kw)
    return opt, op
