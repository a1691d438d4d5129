"""compliancedex_b200 - B200-native drop-in for ComplianceDex's GPIS + probabilistic pregrasp hot path.

    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, optimizers

mirror `from gpis import GPIS` / `optimize_pregrasp.optimizers` of the reference (optimize_pregrasp.py:5,842-848).
All arithmetic runs in libcdx.so (hand-written sm_100a CUDA behind the C ABI of include/cdx.h); there is no CPU
fallback - constructing the classes works anywhere, computing requires a B200.
"""
import os

from . import _lib
from .gpis import GPIS

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def state_path(name):
    """Path of a shipped fitted state (frozen layout of gpis.py:160), e.g. state_path('mug_state')."""
    return os.path.join(DATA_DIR, "gpis_states", f"{name}.npz")


def hand_path(hand):
    return os.path.join(DATA_DIR, "hands", f"{hand}.json")


from .hand import load_hand
from .pregrasp import (ProbabilisticGraspOptimizer, KinGPISGraspOptimizer, GPISGraspOptimizer, optimizers, EE_OFFSETS, WRIST_OFFSET, FINGERTIP_LB,
                       FINGERTIP_UB)
from .completion import assemble_training_set, exterior_points, fit_completion
