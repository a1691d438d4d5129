"""TEST INFRASTRUCTURE ONLY.  Keyed uniform stream for the Kabsch perturbation.

The reference perturbs the 3x3 cross-covariance with `1e-6 * torch.rand_like(H)` before the SVD
(optimize_pregrasp.py:61).  At the standard initialisation (all four targets at the object centre,
optimize_pregrasp.py:940-941) H is exactly rank one, so that perturbation decides the rotation and
the whole Adam trajectory.  For parity the stream is made a pure function of
(seed, iteration, copy p, GLOBAL candidate index b, element 0..8) - independent of batch size and
of the number of ranks (SURVEY.md 8e caveat).  The CUDA kernel implements the same function
(compliancedex_b200/csrc/cdx_noise.cuh); both are checked against fixed values in tests.

u = (mix(mix(mix(seed ^ G*(it+1)) ^ b*C1) ^ (16*p + elem)) >> 11) * 2^-53,  mix = splitmix64 finaliser.
"""
import numpy as np
import torch

_G = np.uint64(0x9E3779B97F4A7C15)
_C1 = np.uint64(0xBF58476D1CE4E5B9)
_C2 = np.uint64(0x94D049BB133111EB)


def _mix(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * _C1
        z = (z ^ (z >> np.uint64(27))) * _C2
        z = z ^ (z >> np.uint64(31))
    return z


def kabsch_noise_np(seed, it, num_copies, num_envs, env_offset=0):
    """-> float64 [num_copies*num_envs, 3, 3] in copy-major row order (row = p*num_envs + b), the
    order of `target_pose.repeat(P,1,1)` (optimize_pregrasp.py:747)."""
    with np.errstate(over="ignore"):
        h0 = _mix(np.uint64(seed) ^ (_G * np.uint64(it + 1)))
        b = (np.arange(num_envs, dtype=np.uint64) + np.uint64(env_offset))
        h1 = _mix(h0 ^ (b * _C1))                                     # [B]
        p = np.arange(num_copies, dtype=np.uint64)
        e = np.arange(9, dtype=np.uint64)
        tag = (p[:, None] * np.uint64(16) + e[None, :])                 # [P,9]
        h2 = _mix(h1[None, :, None] ^ tag[:, None, :])                  # [P,B,9]
    u = (h2 >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
    return u.reshape(num_copies * num_envs, 3, 3)


def kabsch_noise(seed, it, num_rows, env_offset=0, num_copies=3):
    assert num_rows % num_copies == 0
    return torch.from_numpy(kabsch_noise_np(seed, it, num_copies, num_rows // num_copies, env_offset))
