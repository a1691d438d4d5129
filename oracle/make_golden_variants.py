"""TEST TOOLING (build container only): golden vectors for SURVEY.md 8f-2 - the reference's own
KinGPISGraspOptimizer (optimize_pregrasp.py:408-511) and GPISGraspOptimizer (:322-406), imported unmodified from
/root/reference through oracle/ref_harness.py and run on CPU with the Kabsch perturbation (:61) pinned to the keyed
stream of oracle/noise.py.  Run:  python -m oracle.make_golden_variants

Per case: the per-candidate loss `l` of every iteration, the leaves' .grad of the first 3 iterations, the 4-tuple the
optimizer returns, opt_margin (the whole batch's margin at the last improving iteration, :491-494) and the fingertip
positions of forward_kinematics at the initial iterate.
"""
import contextlib
import io
import json
import os
import sys

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.make_golden import GOLD, HANDS, STATES, surface_center
from oracle.pregrasp_oracle import synthetic_candidates


@contextlib.contextmanager
def recording_rmsprop(rec):
    """torch.optim.RMSprop whose step() first copies `l`, `margin` (locals of the calling optimize()) and the leaves' grads"""
    Base = torch.optim.RMSprop

    class Rec(Base):
        def step(self, *a, **k):
            fr = sys._getframe(1)
            while fr is not None and not (fr.f_code.co_name == "optimize" and "l" in fr.f_locals):
                fr = fr.f_back
            rec["loss"].append(fr.f_locals["l"].detach().clone().numpy())
            rec["margin"].append(fr.f_locals["margin"].detach().clone().numpy())
            if len(rec["grads"]) < 3:
                rec["grads"].append([g["params"][0].grad.clone().numpy() for g in self.param_groups])
            return super().step(*a, **k)

    torch.optim.RMSprop = Rec
    try:
        yield
    finally:
        torch.optim.RMSprop = Base


def variant_golden(kind, hand, state, B, iters, seed, friction_mu, mass, com, optimize_target, tag):
    g = rh.load_gpis(f"{state}_state")
    doc = json.load(open(os.path.join(HANDS, f"{hand}.json")))
    d = np.load(os.path.join(STATES, f"{state}_state.npz"))
    center = surface_center(d["X1"], d["y1"])
    q, target, comp, palm = synthetic_candidates(doc, B, center, seed=seed)
    gen = torch.Generator().manual_seed(50 + seed)
    target = target + 0.01 * torch.randn(target.shape, generator=gen, dtype=torch.float64)   # distinct targets: full-rank H
    palm3 = palm[0, :3].copy()
    noise_seed = 2000 + seed
    rh.install()
    with contextlib.redirect_stdout(io.StringIO()):
        import optimize_pregrasp as op
    cfg = rh.hand_config(hand)
    with contextlib.redirect_stdout(io.StringIO()):
        kin = op.KinGPISGraspOptimizer(rh.hand_urdf(hand), ee_link_names=cfg["ee_link_name"],
                                       ee_link_offsets=cfg["ee_link_offset"].tolist(), palm_offset=palm3.tolist(),
                                       num_iters=iters, optimize_target=optimize_target, ref_q=cfg["ref_q"].tolist(),
                                       tip_bounding_box=[op.FINGERTIP_LB, op.FINGERTIP_UB], mass=mass, com=list(com),
                                       gravity=True, uncertainty=10.0)
        tips0 = kin.forward_kinematics(q).detach().numpy()            # [B*4,3]; fresh model, recursive only
    out = dict(q0=q.numpy(), target0=target.numpy(), compliance0=comp.numpy(), palm3=palm3, tips0=tips0,
               friction_mu=np.asarray(friction_mu), mass=np.asarray(mass), com=np.asarray(com),
               noise_seed=np.asarray(noise_seed), iters=np.asarray(iters), optimize_target=np.asarray(int(optimize_target)))
    rec = dict(loss=[], margin=[], grads=[])
    with contextlib.redirect_stdout(io.StringIO()), rh.pinned_kabsch_noise(noise_seed, num_copies=1), recording_rmsprop(rec):
        if kind == "kin":
            res = kin.optimize(q.clone(), target.clone(), comp.clone(), friction_mu, g, verbose=False)
            names = ["q", "target", "compliance"] if optimize_target else ["q", "compliance"]
        else:
            opt = op.GPISGraspOptimizer([op.FINGERTIP_LB, op.FINGERTIP_UB], num_iters=iters, optimize_target=optimize_target,
                                        mass=mass, com=list(com), gravity=True, uncertainty=20.0)
            tip0 = torch.from_numpy(tips0).view(B, 4, 3).clone()
            res = opt.optimize(tip0, target.clone(), comp.clone(), friction_mu, g, verbose=False)
            names = ["tip", "target", "compliance"] if optimize_target else ["tip", "compliance"]
    out["losses"] = np.stack(rec["loss"])                               # [iters,B]
    out["margins"] = np.stack(rec["margin"])                            # [iters,B,4]
    for it, gr in enumerate(rec["grads"]):
        for nm, a in zip(names, gr):
            out[f"grad{it}_{nm}"] = a
    for nm, a in zip(["opt_x", "opt_compliance", "opt_target"], res[:3]):
        out[nm] = a.detach().numpy()
    out["flag"] = np.asarray(bool(res[3]))
    # opt_margin as the reference latches it (:491-494): margin of the last iteration in which any candidate improved
    best = np.full(B, np.inf)
    last = -1
    for it, l in enumerate(out["losses"]):
        upd = l < best
        if upd.any():
            last = it
            best[upd] = l[upd]
    out["opt_margin"] = out["margins"][last]
    out["opt_margin_iter"] = np.asarray(last)
    out["opt_value"] = best
    assert bool((out["opt_margin"] > 0).all()) == bool(out["flag"])
    np.savez(os.path.join(GOLD, f"variant_{tag}.npz"), **out)
    print("variant golden", tag, "loss0", out["losses"][0][:3], "lossT", out["losses"][-1][:3], "flag", out["flag"], "latched at", last)


def main():
    torch.manual_seed(0)
    variant_golden("kin", "leap", "mug", 6, 30, 11, 1, 0.1, [0.0, 0.0, 0.0], True, "kin_leap_mug")
    variant_golden("kin", "allegro", "hammer", 5, 30, 12, 0.5, 0.3, [0.0, 0.0, 0.0], False, "kin_allegro_hammer")
    variant_golden("gpis", "leap", "lego", 6, 40, 13, 1, 0.2, [0.0, 0.0, 0.0], True, "gpis_lego")
    variant_golden("gpis", "allegro", "coffeebottle", 5, 40, 14, 1, 0.5, [0.0, 0.05, 0.0], False, "gpis_coffeebottle")


if __name__ == "__main__":
    main()
