"""TEST TOOLING (build container only): generates tests/golden/*.npz by running the REFERENCE's own code
(gpis.py, optimize_pregrasp.py, vendored differentiable_robot_model - imported unmodified from
/root/reference through oracle/ref_harness.py) on seeded inputs.  Run:  python -m oracle.make_golden

The fitted states of the reference (gpis_states/*_state.npz, frozen layout gpis.py:160) are data the
hot path consumes; they are copied byte-for-byte to compliancedex_b200/data/gpis_states/.
"""
import contextlib
import io
import json
import os
import shutil

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.gpis_oracle import synthetic_training_set
from oracle.pregrasp_oracle import synthetic_candidates

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
STATES = os.path.join(ROOT, "compliancedex_b200", "data", "gpis_states")
HANDS = os.path.join(ROOT, "compliancedex_b200", "data", "hands")
ALL_STATES = ["mug", "mug2", "coffeebottle", "lego", "hammer", "banana", "dummy"]
BOX = {"banana": 0.15, "hammer": 0.18, "coffeebottle": 0.18}


def surface_center(X1, y1):
    surf = X1[np.abs(y1.reshape(-1) + 1.0) < 1e-9]
    return ((surf.max(0) + surf.min(0)) / 2).tolist()


def gpis_goldens():
    for name in ALL_STATES:
        shutil.copyfile(f"{rh.REF}/gpis_states/{name}_state.npz", os.path.join(STATES, f"{name}_state.npz"))
        g = rh.load_gpis(f"{name}_state")
        b = BOX.get(name, 0.1)
        gen = torch.Generator().manual_seed(100 + len(name))
        X = (torch.rand(256, 3, generator=gen, dtype=torch.float64) * 2 - 1) * b
        X = torch.vstack([X, g.X1[:8] + 1e-4, g.X1[8:12]])            # near and exactly on training points
        mean, std, gm, gs = rh.ref_pred_chunked(g, X, chunk=512, with_grad=True)
        normal = g.compute_normal(X)
        np.savez(os.path.join(GOLD, f"gpis_{name}.npz"), X=X.numpy(), mean=mean.numpy(), std=std.numpy(),
                 grad_mean=gm.numpy(), grad_std=gs.numpy(), normal=normal.numpy())
        print("gpis golden", name, X.shape[0], "points; min std", float(std.min()))
    # batched [B,F,3] shape rule (gpis.py:47-50,59) on one state
    g = rh.load_gpis("lego_state")
    Xb = (torch.rand(5, 4, 3, generator=torch.Generator().manual_seed(5), dtype=torch.float64) - 0.5) * 0.2
    m, s = g.pred(Xb)
    n = g.compute_normal(Xb)
    np.savez(os.path.join(GOLD, "gpis_batched_lego.npz"), X=Xb.numpy(), mean=m.numpy(), std=s.numpy(), normal=n.numpy())
    # fit (gpis.py:33-40) on the synthetic generator, tps / rbf / joint
    install_fit = {}
    from gpis import GPIS
    X1, y1, noise = synthetic_training_set(211, seed=3)
    Xq = (torch.rand(64, 3, generator=torch.Generator().manual_seed(9), dtype=torch.float64) - 0.5) * 0.2
    for kern in ["tps", "rbf", "joint"]:
        gg = GPIS(0.08, 1.0, kernel=kern)
        gg.fit(X1, y1, noise)
        m, s = gg.pred(Xq)
        n = gg.compute_normal(Xq)
        install_fit[f"{kern}_E11"] = gg.E11.numpy()
        install_fit[f"{kern}_y1"] = gg.y1.numpy()
        install_fit[f"{kern}_R"] = np.asarray(gg.R.numpy() if hasattr(gg, "R") else 0.0)
        install_fit[f"{kern}_mean"] = m.numpy()
        install_fit[f"{kern}_std"] = s.numpy()
        install_fit[f"{kern}_normal"] = n.numpy()
    np.savez(os.path.join(GOLD, "gpis_fit_synth211.npz"), X1=X1.numpy(), y1=y1.numpy(), noise=noise.numpy(),
             Xq=Xq.numpy(), **install_fit)
    # visualisation grid driver (gpis.py:113-125), small
    g = rh.load_gpis("banana_state")
    mean, std, nrm, lb, ub = g.get_visualization_data([-0.15] * 3, [0.15] * 3, steps=6)
    np.savez(os.path.join(GOLD, "gpis_viz_banana6.npz"), mean=mean, std=std, normal=nrm, lb=lb, ub=ub)
    print("fit + viz goldens done")


def pregrasp_golden(hand, state, B, iters, seed, friction_mu, mass, com, tag):
    g = rh.load_gpis(f"{state}_state")
    doc = json.load(open(os.path.join(HANDS, f"{hand}.json")))
    d = np.load(os.path.join(STATES, f"{state}_state.npz"))
    center = surface_center(d["X1"], d["y1"])
    q, target, comp, palm = synthetic_candidates(doc, B, center, seed=seed)
    noise_seed = 1000 + seed
    out = dict(q0=q.numpy(), target0=target.numpy(), compliance0=comp.numpy(), palm0=palm, center=np.asarray(center),
               friction_mu=np.asarray(friction_mu), mass=np.asarray(mass), com=np.asarray(com),
               noise_seed=np.asarray(noise_seed), iters=np.asarray(iters))
    ropt, op = rh.make_prob_optimizer(hand, palm, iters, mass=mass, com=com)
    losses, grads = [], []
    with contextlib.redirect_stdout(io.StringIO()), rh.pinned_kabsch_noise(noise_seed):
        orig_closure = ropt.closure

        def closure(joint_angles, compliance, target_pose, palm_poses, palm_oris, *a, **k):
            r = orig_closure(joint_angles, compliance, target_pose, palm_poses, palm_oris, *a, **k)
            losses.append(ropt.total_loss.detach().clone().numpy())
            if len(grads) < 3:
                grads.append([p.grad.clone().numpy() for p in (joint_angles, compliance, target_pose, palm_poses, palm_oris)])
            return r

        ropt.closure = closure
        res = ropt.optimize(q, target, comp, friction_mu, g)
        tips0 = ropt.forward_kinematics(q, torch.from_numpy(palm)).detach().numpy()
    out["tips0"] = tips0
    out["losses"] = np.stack(losses)                                   # [iters, B]
    for it, gr in enumerate(grads):
        for nm, a in zip(["q", "compliance", "target", "palm_pos", "palm_ori"], gr):
            out[f"grad{it}_{nm}"] = a
    for nm, a in zip(["opt_q", "opt_compliance", "opt_target", "opt_palm", "opt_margin"], res):
        out[nm] = a.detach().numpy()
    np.savez(os.path.join(GOLD, f"pregrasp_{tag}.npz"), **out)
    print("pregrasp golden", tag, "loss0", losses[0][:3], "lossT", losses[-1][:3])


def main():
    os.makedirs(GOLD, exist_ok=True)
    os.makedirs(STATES, exist_ok=True)
    rh.install()
    torch.manual_seed(0)
    gpis_goldens()
    pregrasp_golden("leap", "mug", 8, 40, 1, 1, 0.1, [0.0, 0.0, 0.0], "leap_mug")
    pregrasp_golden("allegro", "hammer", 6, 30, 2, 0.5, 0.3, [0.0, 0.0, 0.0], "allegro_hammer")
    pregrasp_golden("allegro", "coffeebottle", 6, 30, 3, 1, 0.5, [0.0, 0.05, 0.0], "allegro_coffeebottle")


if __name__ == "__main__":
    main()
