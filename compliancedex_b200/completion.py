"""Training-set assembly and fit of train_gpis_completion.py:20-62 as functions.

The reference script builds X1 = [14 exterior points at +-boundary | observed[:num_obs] | completed[:num_gen] | interior points],
labels +boundary / 0 / -boundary, per-class noise 0.3 / obs_noise / gen_noise / 0.02, fits GPIS(0.08, bias) and writes
`{exp}_gpis.npz` (the 100^3 visualisation grid) and `gpis_states/{exp}_state.npz`.  Interior points are softmax(10 U)-weighted
convex combinations of the surface points with U ~ torch.rand (unseeded in the reference): pass `weights` or a `generator`
to make them reproducible.
"""
import torch


def exterior_points(boundary):
    """train_gpis_completion.py:25-38: the 8 corners and 6 face centres of the +-boundary box, in the reference's order."""
    b = float(boundary)
    return torch.tensor([[-b, -b, -b], [b, -b, -b], [-b, b, -b], [b, b, -b], [-b, -b, b], [b, -b, b], [-b, b, b], [b, b, b],
                         [-b, 0.0, 0.0], [0.0, -b, 0.0], [b, 0.0, 0.0], [0.0, b, 0.0], [0.0, 0.0, b], [0.0, 0.0, -b]],
                        dtype=torch.float32).double()   # float32 literal tensor widened by .double() (:38): 0.1 -> 0.10000000149


def assemble_training_set(observed, completed, num_obs=128, num_gen=64, obs_noise=0.005, gen_noise=0.02, num_internals=5,
                          boundary=0.1, weights=None, generator=None):
    """-> X1 [N,3], y1 [N,1], noise [N] (float64, on the device of `observed`), N = 14 + num_obs + num_gen + num_internals.
    `weights` [num_internals, num_obs + num_gen]: the uniform draws of train_gpis_completion.py:40 (before x10 / softmax)."""
    observed = torch.as_tensor(observed, dtype=torch.float64)[:num_obs]
    dev = observed.device
    completed = torch.as_tensor(completed, dtype=torch.float64)[:num_gen].to(dev)
    ext = exterior_points(boundary).to(dev)
    surface = torch.vstack([observed, completed])
    if weights is None:
        weights = torch.rand(num_internals, surface.shape[0], generator=generator)          # :40 (float32 draws, then .double())
    weights = torch.as_tensor(weights).double().to(dev)
    internal = torch.softmax(weights * 10, dim=1) @ surface                                  # :41-42
    y1 = torch.hstack([boundary * torch.ones(len(ext)), torch.zeros(len(surface)),
                       -boundary * torch.ones(len(internal))]).view(-1, 1).double().to(dev)    # :46-48
    X1 = torch.vstack([ext, observed, completed, internal])                                   # :50-53
    noise = torch.tensor([0.3] * len(ext) + [obs_noise] * len(observed) + [gen_noise] * len(completed) +
                         [0.02] * len(internal)).double().to(dev)                            # :55-58
    return X1, y1, noise


def fit_completion(observed, completed, bias=1.0, device="cuda", **kw):
    """train_gpis_completion.py:44-58: GPIS(0.08, bias).fit on the assembled set -> the fitted compliancedex_b200.GPIS."""
    from .gpis import GPIS
    X1, y1, noise = assemble_training_set(observed, completed, **kw)
    gpis = GPIS(0.08, bias)
    gpis.fit(X1.to(device), y1.to(device), noise.to(device))
    return gpis
