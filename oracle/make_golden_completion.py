"""TEST TOOLING (build container only): golden vectors for the training-set assembly of train_gpis_completion.py (SURVEY 8a,
row a8).  The reference file is a script (argparse + open3d windows), so its assembly statements - lines 24-58, from
"# Prepare auxiliary data" to the end of the gpis.fit(...) call - are executed VERBATIM from /root/reference (read at run
time, not copied) in a namespace that supplies `args`, the observed / completed slices of lines 20-22 and a recording stand-in
for the GPIS object; torch.rand is seeded.  -> tests/golden/completion_coffeebottle.npz (inputs, the uniform draws, X1, y1,
noise as passed to fit).  Run:  python -m oracle.make_golden_completion"""
import os
import types

import numpy as np
import torch

from oracle import ref_harness as rh
from oracle.make_golden import GOLD


class _Recorder:
    def __init__(self, *a):
        self.args = a

    def fit(self, X1, y1, noise):
        self.X1, self.y1, self.noise = X1, y1, noise


def main():
    rh.install()                                   # Tensor.cuda -> no-op on this CPU-only container
    src = open(f"{rh.REF}/train_gpis_completion.py").read().split("\n")
    start = next(i for i, l in enumerate(src) if l.startswith("# Prepare auxiliary data"))
    end = next(i for i, l in enumerate(src) if l.startswith("test_mean, test_var"))
    code = "\n".join(src[start:end])
    assert "gpis.fit(" in code and "torch.rand(" in code
    d = np.load(f"{rh.REF}/assets/coffeebottle/completed_pcd.npz")
    out = {}
    for tag, kw in (("default", dict(num_obs=128, num_gen=64, obs_noise=0.005, gen_noise=0.02, num_internals=5, boundary=0.1, bias=1.0)),
                    ("box018", dict(num_obs=200, num_gen=33, obs_noise=0.01, gen_noise=0.03, num_internals=9, boundary=0.18, bias=1.0))):
        args = types.SimpleNamespace(exp_name="coffeebottle", **kw)
        observed = torch.tensor(d["observed"][:args.num_obs]).double()
        completed = torch.tensor(d["completed"][:args.num_gen]).double()
        torch.manual_seed(7)
        draws = torch.rand(args.num_internals, len(observed) + len(completed))
        torch.manual_seed(7)
        ns = dict(torch=torch, np=np, args=args, observed=observed, completed=completed, GPIS=_Recorder)
        exec(compile(code, "train_gpis_completion.py[24:58]", "exec"), ns)
        g = ns["gpis"]
        out.update({f"{tag}_observed": observed.numpy(), f"{tag}_completed": completed.numpy(), f"{tag}_draws": draws.numpy(),
                    f"{tag}_X1": g.X1.numpy(), f"{tag}_y1": g.y1.numpy(), f"{tag}_noise": g.noise.numpy(),
                    f"{tag}_args": np.array([kw[k] for k in ("num_obs", "num_gen", "obs_noise", "gen_noise", "num_internals", "boundary", "bias")])})
        print(tag, g.X1.shape, g.y1.shape, g.noise.shape, g.args)
    np.savez_compressed(os.path.join(GOLD, "completion_coffeebottle.npz"), **out)


if __name__ == "__main__":
    main()
