"""TEST INFRASTRUCTURE ONLY: builds tests/hostcheck/libhostcheck.so (g++ compile of the per-candidate math in
compliancedex_b200/csrc/cdx_pg_math.cuh) and binds it with ctypes.  Used by CPU tests to check the analytic
gradients against the oracle's autograd; never imported by the product."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "libhostcheck.so")
MAXD = 6  # CDX_MAX_CHAIN_DOF


class PgConst(ctypes.Structure):
    _fields_ = [("cos_mu", ctypes.c_double), ("com", ctypes.c_double * 3), ("dummy_k", ctypes.c_double),
                ("dummy_M", ctypes.c_double), ("gravity", ctypes.c_int), ("uncertainty", ctypes.c_double),
                ("w_reward", ctypes.c_double), ("w_cm", ctypes.c_double), ("w_tar", ctypes.c_double),
                ("force_clamp", ctypes.c_double), ("w_center", ctypes.c_double), ("var_sum", ctypes.c_int)]


class PgSetup(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int), ("U", ctypes.c_int), ("uniq", ctypes.c_int * 4),
                ("coeff", (ctypes.c_double * 4) * 4), ("weight", ctypes.c_double * 4),
                ("ref_q", ctypes.c_double * 16), ("n_dofs", ctypes.c_int), ("optimize_palm", ctypes.c_int),
                ("variant", ctypes.c_int), ("w_ref", ctypes.c_double), ("w_pre", ctypes.c_double), ("C", PgConst)]


def build():
    srcs = [os.path.join(HERE, "hostcheck.cpp"), os.path.join(ROOT, "compliancedex_b200", "csrc", "cdx_pg_math.cuh"),
            os.path.join(ROOT, "compliancedex_b200", "csrc", "cdx_noise.cuh")]
    if not os.path.isfile(SO) or any(os.path.getmtime(s) > os.path.getmtime(SO) for s in srcs):
        # -ffp-contract=off: no FMA contraction, so float32 FK follows the source expression order
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-x", "c++",
                               srcs[0], "-o", SO])
    lib = ctypes.CDLL(SO)
    assert lib.hc_setup_size() == ctypes.sizeof(PgSetup), (lib.hc_setup_size(), ctypes.sizeof(PgSetup))
    return lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class Host:
    def __init__(self, tree):
        self.lib = build()
        self.hand = ctypes.create_string_buffer(self.lib.hc_hand_size())
        rc = self.lib.hc_build(tree.n_bodies, _p(tree.parent), _p(tree.dof), _p(tree.axis), _p(tree.sign), _p(tree.rot),
                               _p(tree.xyz), tree.n_dofs, _p(tree.tip_body), _p(tree.tip_offset), self.hand)
        assert rc == 0, rc

    def fk(self, q):
        q = np.ascontiguousarray(q, dtype=np.float64)
        tl = np.zeros((4, 3))
        J = np.zeros((4, 3, MAXD), dtype=np.float32)
        self.lib.hc_fk(self.hand, _p(q), _p(tl), _p(J))
        return tl, J

    def svd3(self, A):
        A = np.ascontiguousarray(A, dtype=np.float64)
        U, S, V = np.zeros((3, 3)), np.zeros(3), np.zeros((3, 3))
        self.lib.hc_svd3(_p(A), _p(U), _p(S), _p(V))
        return U, S, V

    def closure(self, setup, q, k, T, palm, dist, sd, gmu, gsd, mean9, gmean9, seed, it, b, tl=None):
        arrs = [np.ascontiguousarray(x, dtype=np.float64) for x in (q, k, T, palm, dist, sd, gmu, gsd, mean9, gmean9)]
        total = np.zeros(1)
        tm, gq, gk, gT, gp = np.zeros(4), np.zeros(16), np.zeros(4), np.zeros((4, 3)), np.zeros(6)
        self.lib.hc_closure(ctypes.byref(setup), self.hand, *[_p(x) for x in arrs], ctypes.c_uint64(seed), int(it),
                            ctypes.c_int64(b), _p(total), _p(tm), _p(gq), _p(gk), _p(gT), _p(gp),
                            _p(np.ascontiguousarray(tl, dtype=np.float64)) if tl is not None else None)
        return total[0], tm, gq, gk, gT, gp


def make_setup(coeffs, weights, ref_q, n_dofs, optimize_palm, cos_mu, com, dummy_k, gravity, uncertainty):
    s = PgSetup()
    coeffs = np.asarray(coeffs, dtype=np.float32)
    s.P = coeffs.shape[0]
    rows = []
    for p in range(s.P):
        key = tuple(coeffs[p].tolist())
        if key not in rows:
            rows.append(key)
        s.uniq[p] = rows.index(key)
        for f in range(4):
            s.coeff[p][f] = float(coeffs[p, f])
        s.weight[p] = float(np.float32(weights[p]))
    s.U = len(rows)
    for i in range(n_dofs):
        s.ref_q[i] = float(np.float32(ref_q[i]))
    s.n_dofs = n_dofs
    s.optimize_palm = int(optimize_palm)
    s.C.cos_mu = float(cos_mu)
    for i in range(3):
        s.C.com[i] = float(np.float32(com[i]))
    s.C.dummy_k = float(np.float32(dummy_k))
    s.C.dummy_M = 2.0
    s.C.gravity = int(gravity)
    s.C.uncertainty = float(uncertainty)
    # ProbabilisticGraspOptimizer weights (optimize_pregrasp.py:719-737, :757-759)
    s.variant, s.w_ref, s.w_pre = 0, 10.0, 5.0
    s.C.w_reward, s.C.w_cm, s.C.w_tar, s.C.force_clamp, s.C.w_center, s.C.var_sum = 200.0, 200.0, 20.0, 10.0, 0.0, 0
    return s
