// Per-candidate arithmetic of the probabilistic pregrasp objective (optimize_pregrasp.py:49-118, 657-669,
// 703-769) and of the fingertip forward kinematics (vendored differentiable-robot-model: robot_model.py:141-264,
// rigid_body.py:130-157, spatial_vector_algebra.py:14-53,98-136, se3_so3_util.py:241-250), written as plain
// scalar functions with an ANALYTIC reverse pass (no autograd).  The functions are __host__ __device__ so that
// the same source can be compiled by g++ into tests/hostcheck (CPU tests check the derivations against the
// oracle's autograd without a GPU); the product only ever runs them inside the kernels of cdx_pregrasp.cu.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define CDX_HD __host__ __device__ __forceinline__
#else
#define CDX_HD inline
#endif

#define CDX_MAX_CHAIN 8      // bodies between the root and a tip link (Allegro: 6, Leap: 4)
#define CDX_MAX_CHAIN_DOF 6  // actuated joints on one finger chain (both hands: 4)

// one finger's kinematic chain, root excluded, in root -> tip order (float32 as the reference stores it)
struct cdx_chain {
  int n;                          // bodies on the chain
  int dof[CDX_MAX_CHAIN];         // column of q, or -1 for a fixed joint
  int axis[CDX_MAX_CHAIN];        // 0,1,2
  float sign[CDX_MAX_CHAIN];      // +-1 (rigid_body.py:149-154)
  float fixed_rot[CDX_MAX_CHAIN][9];  // (Rz(yaw) Ry(pitch)) Rx(roll), rigid_body.py:142
  float xyz[CDX_MAX_CHAIN][3];
  float offset[3];                // tip offset (robot_model.py:260)
  int ndof;                       // actuated joints on the chain
  int dofs[CDX_MAX_CHAIN_DOF];    // their q columns in chain order
};

struct cdx_hand_dev {
  cdx_chain finger[4];
  int n_dofs;
};

// ------------------------------------------------------------------------------------------------ helpers
// float32 product / sum that the compiler may NOT contract into an FMA: the reference's FK runs as separate torch
// float32 multiplies and adds (CPU ATen), so the kernels round after every operation as well.
#ifdef __CUDA_ARCH__
#define CDX_FM(a, b) __fmul_rn((a), (b))
#define CDX_FA(a, b) __fadd_rn((a), (b))
#define CDX_FS(a, b) __fsub_rn((a), (b))
#else   // host build (tests/hostcheck) is compiled with -ffp-contract=off
#define CDX_FM(a, b) ((a) * (b))
#define CDX_FA(a, b) ((a) + (b))
#define CDX_FS(a, b) ((a) - (b))
#endif
CDX_HD void cdx_cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
CDX_HD void cdx_cross3f(const float* a, const float* b, float* o) {
  o[0] = CDX_FS(CDX_FM(a[1], b[2]), CDX_FM(a[2], b[1]));
  o[1] = CDX_FS(CDX_FM(a[2], b[0]), CDX_FM(a[0], b[2]));
  o[2] = CDX_FS(CDX_FM(a[0], b[1]), CDX_FM(a[1], b[0]));
}
CDX_HD double cdx_dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
CDX_HD float cdx_dot3f(const float* a, const float* b) {
  return CDX_FA(CDX_FA(CDX_FM(a[0], b[0]), CDX_FM(a[1], b[1])), CDX_FM(a[2], b[2]));
}
// o = M v, o = M^T v  (row-major 3x3)
CDX_HD void cdx_mv(const double* M, const double* v, double* o) {
  o[0] = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
  o[1] = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
  o[2] = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
}
CDX_HD void cdx_mtv(const double* M, const double* v, double* o) {
  o[0] = M[0] * v[0] + M[3] * v[1] + M[6] * v[2];
  o[1] = M[1] * v[0] + M[4] * v[1] + M[7] * v[2];
  o[2] = M[2] * v[0] + M[5] * v[1] + M[8] * v[2];
}
CDX_HD void cdx_mm(const double* A, const double* B, double* C) {
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) C[r * 3 + c] = A[r * 3] * B[c] + A[r * 3 + 1] * B[3 + c] + A[r * 3 + 2] * B[6 + c];
}
CDX_HD void cdx_mmf(const float* A, const float* B, float* C) {
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++)
      C[r * 3 + c] = CDX_FA(CDX_FA(CDX_FM(A[r * 3], B[c]), CDX_FM(A[r * 3 + 1], B[3 + c])), CDX_FM(A[r * 3 + 2], B[6 + c]));
}

// ------------------------------------------------------------------------------------------------ FK
// un-normalised quaternion (x,y,z,w) of CoordinateTransform.get_quaternion (spatial_vector_algebra.py:108-136):
// branch -1 is `t > M33` (:118-122), branch i in {0,1,2} the i,j,k rotation of :123-133.  The map is affine in R;
// with add_one = 0 it is applied to a differential dR.  Returns the component the normaliser is built from.
template <int I>
CDX_HD float cdx_quat_raw_axis(const float* R, float one, float* q) {
  constexpr int i = I, j = (i + 1) % 3, k = (j + 1) % 3;
  q[i] = R[i * 3 + i] - (R[j * 3 + j] + R[k * 3 + k]) + one;
  q[j] = R[i * 3 + j] + R[j * 3 + i];
  q[k] = R[k * 3 + i] + R[i * 3 + k];
  q[3] = R[k * 3 + j] - R[j * 3 + k];
  return q[i];
}
CDX_HD float cdx_quat_raw(const float* R, int branch, float one, float* q) {   // compile-time subscripts in every branch
  if (branch < 0) {
    q[0] = R[7] - R[5];
    q[1] = R[2] - R[6];
    q[2] = R[3] - R[1];
    q[3] = R[0] + R[4] + R[8] + one;
    return q[3];
  }
  if (branch == 0) return cdx_quat_raw_axis<0>(R, one, q);
  if (branch == 1) return cdx_quat_raw_axis<1>(R, one, q);
  return cdx_quat_raw_axis<2>(R, one, q);
}

// Fingertip position in the hand frame and the gradient the REFERENCE's autograd produces for it
// (the quaternion normaliser 0.5/math.sqrt(.) is a python float, i.e. a constant).
//   q32: joint angles already cast to float32 (optimize_pregrasp.py:665)
//   tip[3], J[a][c] = d tip[a] / d q[ch.dofs[c]]   (float32 arithmetic throughout, like the reference)
//   q64 (optional): the angles as float64 - then cos/sin are evaluated in double and rounded to float32, which is what
//   x_rot/y_rot/z_rot do with a float64 q (their buffer is float32, spatial_vector_algebra.py:19-24): the Kin optimizers
//   pass q without the .float() cast (optimize_pregrasp.py:436)
//   With q64 the chain is also composed the way compute_forward_kinematics(recursive=True) does it (rigid_body.py:85-127:
//   every body applies its joint pose to the poses of its descendants, i.e. tip-to-root association
//   J_0 (J_1 (... (J_{n-1} I)))); the Jacobian keeps using the root-to-tip prefix frames (same quantities up to float32
//   rounding).
template <bool REC = false>   // REC: q64 given (compile-time so that the non-recursive path keeps no per-joint rotation array)
CDX_HD void cdx_fk_finger(const cdx_chain& ch, const float* q32, float* tip, float (*J)[CDX_MAX_CHAIN_DOF],
                          const double* q64 = nullptr) {
  // Every local array below is indexed with compile-time subscripts only (the joint axis and the running dof count select
  // through predicates): on the device they stay in registers.  With run-time subscripts (R[r * 3 + axis], om[nd][r], ...)
  // the whole chain walked through local memory - a load-to-use latency on every step of a serial chain that four lanes
  // per candidate execute alone.
  float R[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f};
  float p[3] = {0.f, 0.f, 0.f};
  float om[CDX_MAX_CHAIN_DOF][3], org[CDX_MAX_CHAIN_DOF][3];
#pragma unroll
  for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++)
    for (int r = 0; r < 3; r++) { om[c][r] = 0.f; org[c][r] = 0.f; }
  float jall[REC ? CDX_MAX_CHAIN : 1][9];
  int nd = 0;
  for (int i = 0; i < ch.n; i++) {
    float jr[9];
    const float* F = ch.fixed_rot[i];
    const int a = ch.axis[i];
    if (ch.dof[i] >= 0) {
      const float ang = CDX_FM(ch.sign[i], q32[ch.dof[i]]);
      float c, s;
      if (REC) {
        const double a64 = (double)ch.sign[i] * q64[ch.dof[i]];
        c = (float)cos(a64); s = (float)sin(a64);
      } else {
        c = cosf(ang); s = sinf(ang);             // torch.cos/sin on a CUDA float32 tensor are these libdevice functions
      }
      // fixed @ R_axis(ang): the two columns orthogonal to the axis mix, the axis column is kept
      const int u = (a + 1) % 3, v = (a + 2) % 3;
#pragma unroll
      for (int r = 0; r < 3; r++) {
        const float fu = F[r * 3 + u], fv = F[r * 3 + v], fa = F[r * 3 + a];
        const float ju = CDX_FA(CDX_FM(fu, c), CDX_FM(fv, s)), jv = CDX_FS(CDX_FM(fv, c), CDX_FM(fu, s));
#pragma unroll
        for (int col = 0; col < 3; col++) jr[r * 3 + col] = col == a ? fa : (col == u ? ju : jv);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 9; e++) jr[e] = F[e];
    }
    if (REC) {
#pragma unroll
      for (int b = 0; b < CDX_MAX_CHAIN; b++)
        if (b == i)
          for (int e = 0; e < 9; e++) jall[REC ? b : 0][e] = jr[e];
    }
    // CoordinateTransform.multiply_transform (spatial_vector_algebra.py:98-103)
    float Rn[9], pn[3];
    cdx_mmf(R, jr, Rn);
    const float x0 = ch.xyz[i][0], x1 = ch.xyz[i][1], x2 = ch.xyz[i][2];
#pragma unroll
    for (int r = 0; r < 3; r++)
      pn[r] = CDX_FA(CDX_FA(CDX_FA(CDX_FM(R[r * 3], x0), CDX_FM(R[r * 3 + 1], x1)), CDX_FM(R[r * 3 + 2], x2)), p[r]);
#pragma unroll
    for (int e = 0; e < 9; e++) R[e] = Rn[e];
#pragma unroll
    for (int r = 0; r < 3; r++) p[r] = pn[r];
    if (ch.dof[i] >= 0 && nd < CDX_MAX_CHAIN_DOF) {
      const float sg = ch.sign[i];
#pragma unroll
      for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++)
        if (c == nd) {
#pragma unroll
          for (int r = 0; r < 3; r++) {
            om[c][r] = sg * (a == 0 ? R[r * 3] : (a == 1 ? R[r * 3 + 1] : R[r * 3 + 2]));
            org[c][r] = p[r];
          }
        }
      nd++;
    }
  }
  if (REC) {   // recursive=True association (rigid_body.py:118-127)
    float Ra[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f}, pa[3] = {0.f, 0.f, 0.f};
    for (int i = ch.n - 1; i >= 0; i--) {
      float jr[9];
#pragma unroll
      for (int b = 0; b < CDX_MAX_CHAIN; b++)
        if (b == i)
          for (int e = 0; e < 9; e++) jr[e] = jall[REC ? b : 0][e];
      float Rn[9], pn[3];
      cdx_mmf(jr, Ra, Rn);
      for (int r = 0; r < 3; r++)
        pn[r] = CDX_FA(CDX_FA(CDX_FA(CDX_FM(jr[r * 3], pa[0]), CDX_FM(jr[r * 3 + 1], pa[1])), CDX_FM(jr[r * 3 + 2], pa[2])), ch.xyz[i][r]);
      for (int e = 0; e < 9; e++) Ra[e] = Rn[e];
      for (int r = 0; r < 3; r++) pa[r] = pn[r];
    }
    for (int e = 0; e < 9; e++) R[e] = Ra[e];
    for (int r = 0; r < 3; r++) p[r] = pa[r];
  }
  // tip = pos + quat_rotate(quat, offset)   (robot_model.py:255-263, se3_so3_util.py:241-250)
  const float t = R[0] + R[4] + R[8] + 1.f;
  int branch = -1;
  if (!(t > 1.f)) {
    branch = 0;
    if (R[4] > R[0]) branch = 1;
    if (R[8] > (branch == 0 ? R[0] : R[4])) branch = 2;
  }
  float qr[4];
  const float tn = cdx_quat_raw(R, branch, 1.f, qr);
  const float s = (float)(0.5 / sqrt((double)tn));  // python-float normaliser (spatial_vector_algebra.py:135)
  const float qv[3] = {CDX_FM(qr[0], s), CDX_FM(qr[1], s), CDX_FM(qr[2], s)}, qw = CDX_FM(qr[3], s);
  const float* o = ch.offset;
  float cvo[3];
  cdx_cross3f(qv, o, cvo);
  const float qvo = cdx_dot3f(qv, o);
  // a = v (2 qw^2 - 1);  b = cross(qv, v) qw 2;  c = qv (qv . v) 2;  a + b + c   (se3_so3_util.py:245-250)
  const float aw = CDX_FS(CDX_FM(2.f, CDX_FM(qw, qw)), 1.f);
  for (int r = 0; r < 3; r++)
    tip[r] = CDX_FA(p[r], CDX_FA(CDX_FA(CDX_FM(o[r], aw), CDX_FM(CDX_FM(cvo[r], qw), 2.f)), CDX_FM(CDX_FM(qv[r], qvo), 2.f)));
  for (int c = 0; c < nd; c++) {
    float omc[3], orc[3];
#pragma unroll
    for (int r = 0; r < 3; r++) { omc[r] = om[0][r]; orc[r] = org[0][r]; }
#pragma unroll
    for (int b = 1; b < CDX_MAX_CHAIN_DOF; b++)
      if (b == c)
        for (int r = 0; r < 3; r++) { omc[r] = om[b][r]; orc[r] = org[b][r]; }
    float d[3] = {p[0] - orc[0], p[1] - orc[1], p[2] - orc[2]}, dp[3];
    cdx_cross3f(omc, d, dp);
    float dR[9];  // [omega]x R
#pragma unroll
    for (int col = 0; col < 3; col++) {
      const float rc[3] = {R[col], R[3 + col], R[6 + col]};
      float x[3];
      cdx_cross3f(omc, rc, x);
      dR[col] = x[0]; dR[3 + col] = x[1]; dR[6 + col] = x[2];
    }
    float dq[4];
    cdx_quat_raw(dR, branch, 0.f, dq);
    const float dqv[3] = {dq[0] * s, dq[1] * s, dq[2] * s}, dqw = dq[3] * s;
    float dcvo[3];
    cdx_cross3f(dqv, o, dcvo);
    const float dqvo = cdx_dot3f(dqv, o);
#pragma unroll
    for (int r = 0; r < 3; r++) {
      const float jv = dp[r] + (o[r] * (4.f * qw * dqw) + 2.f * (dqw * cvo[r] + qw * dcvo[r]) + 2.f * (dqv[r] * qvo + qv[r] * dqvo));
#pragma unroll
      for (int b = 0; b < CDX_MAX_CHAIN_DOF; b++)
        if (b == c) J[r][b] = jv;               // compile-time column subscript: a caller's local J stays in registers
    }
  }
#pragma unroll
  for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++)
    if (c >= nd) J[0][c] = J[1][c] = J[2][c] = 0.f;
}

// euler_angles_to_matrix(e, "XYZ") = Rx(e0) Ry(e1) Rz(e2)  (math_utils.py:97-123); optional partials dR/de_i
CDX_HD void cdx_euler_xyz(const double* e, double* R, double* dR0, double* dR1, double* dR2) {
  // sincos shares the argument reduction of the pair (half the code and latency of separate calls; same values)
  double ca, sa, cb, sb, cc, sc;
  sincos(e[0], &sa, &ca); sincos(e[1], &sb, &cb); sincos(e[2], &sc, &cc);
  const double X[9] = {1, 0, 0, 0, ca, -sa, 0, sa, ca}, Y[9] = {cb, 0, sb, 0, 1, 0, -sb, 0, cb}, Z[9] = {cc, -sc, 0, sc, cc, 0, 0, 0, 1};
  double XY[9];
  cdx_mm(X, Y, XY);
  cdx_mm(XY, Z, R);
  if (dR0) {
    const double dX[9] = {0, 0, 0, 0, -sa, -ca, 0, ca, -sa}, dY[9] = {-sb, 0, cb, 0, 0, 0, -cb, 0, -sb}, dZ[9] = {-sc, -cc, 0, cc, -sc, 0, 0, 0, 0};
    double T[9];
    cdx_mm(dX, Y, T); cdx_mm(T, Z, dR0);
    cdx_mm(X, dY, T); cdx_mm(T, Z, dR1);
    cdx_mm(XY, dZ, dR2);
  }
}

// timing hook for the step kernel's sections (scripts/pg_ticks.py builds with -DCDX_PG_TICKS); a no-op otherwise
#ifndef CDX_PG_TICK
#define CDX_PG_TICK(i)
#endif

// ------------------------------------------------------------------------------------------------ 3x3 SVD
// One-sided (Hestenes) Jacobi: A V = U diag(S), S descending - keeps full RELATIVE accuracy of the tiny singular
// values that decide the Kabsch rotation when the cross-covariance is rank one + 1e-6 noise
// (optimize_pregrasp.py:61 at the standard initialisation).  A, U, V row-major.
CDX_HD void cdx_svd3(const double* A, double* U, double* S, double* V) {
  double g[3][3], v[3][3];  // columns: g[c][r]
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) { g[c][r] = A[r * 3 + c]; v[c][r] = (r == c) ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 12; sweep++) {
    bool rotated = false;
    for (int pr = 0; pr < 3; pr++) {
      const int p = (pr == 2) ? 1 : 0, q = (pr == 0) ? 1 : 2;
      const double al = cdx_dot3(g[p], g[p]), be = cdx_dot3(g[q], g[q]), ga = cdx_dot3(g[p], g[q]);
      if (ga * ga <= 1e-30 * (al * be)) continue;   // columns orthogonal to 1e-15 relative: converged pair
      rotated = true;
      const double zeta = (be - al) / (2.0 * ga);
      const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
      const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
      for (int r = 0; r < 3; r++) {
        const double gp = g[p][r], gq = g[q][r], vp = v[p][r], vq = v[q][r];
        g[p][r] = c * gp - s * gq; g[q][r] = s * gp + c * gq;
        v[p][r] = c * vp - s * vq; v[q][r] = s * vp + c * vq;
      }
    }
    if (!rotated) break;
  }
  double s[3] = {sqrt(cdx_dot3(g[0], g[0])), sqrt(cdx_dot3(g[1], g[1])), sqrt(cdx_dot3(g[2], g[2]))};
  int o0 = 0, o1 = 1, o2 = 2, tmp;
  if (s[o0] < s[o1]) { tmp = o0; o0 = o1; o1 = tmp; }
  if (s[o0] < s[o2]) { tmp = o0; o0 = o2; o2 = tmp; }
  if (s[o1] < s[o2]) { tmp = o1; o1 = o2; o2 = tmp; }
  const int ord[3] = {o0, o1, o2};
  double u[3][3];
  for (int m = 0; m < 3; m++) {
    const int c = ord[m];
    S[m] = s[c];
    const double inv = s[c] > 0.0 ? 1.0 / s[c] : 0.0;
    for (int r = 0; r < 3; r++) { u[m][r] = g[c][r] * inv; V[r * 3 + m] = v[c][r]; }
  }
  // exactly rank-deficient input: complete U to an orthonormal basis (never hit with the noise term present)
  if (!(S[2] > 0.0)) {
    if (!(S[1] > 0.0)) {
      if (!(S[0] > 0.0)) { u[0][0] = 1.0; u[0][1] = 0.0; u[0][2] = 0.0; }
      const int k = fabs(u[0][0]) < 0.6 ? 0 : 1;
      double e[3] = {0.0, 0.0, 0.0};
      e[k] = 1.0;
      cdx_cross3(u[0], e, u[1]);
      const double nn = 1.0 / sqrt(cdx_dot3(u[1], u[1]));
      for (int r = 0; r < 3; r++) u[1][r] *= nn;
    }
    cdx_cross3(u[0], u[1], u[2]);
  }
  for (int m = 0; m < 3; m++)
    for (int r = 0; r < 3; r++) U[r * 3 + m] = u[m][r];
}
CDX_HD double cdx_det3(const double* M) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// ------------------------------------------------------------------------------------------------ objective
struct cdx_pg_const {
  double cos_mu;       // torch.sqrt(1/(1+torch.tensor(mu)**2)) as float32 (optimize_pregrasp.py:111,707)
  double com[3];       // float32-rounded COM (dummy_tip is a float32 tensor, :90,:92)
  double dummy_k;      // float32(gravity * mass / M) (:94)
  double dummy_M;      // M = 2.0 (:73)
  int gravity;         // 0: no dummy finger (:98-99)
  double uncertainty;  // :630
  // weights of the objective's terms; the defaults are ProbabilisticGraspOptimizer.compute_loss (:719-737), the Kin / GPIS
  // optimizers (:468-479, :377-387) use other values and a centre term instead of the contact-margin term
  double w_reward = 200.0, w_cm = 200.0, w_tar = 20.0, force_clamp = 10.0, w_center = 0.0;
  int var_sum = 0;     // 0: uncertainty * max_f log(100 std)   1: uncertainty * sum_f std (:386)
};

// One row (copy p of candidate b) of compute_loss (optimize_pregrasp.py:713-739) without the ref_cost term, and
// its gradient w.r.t. the row's all_tip a[4][3], target T[4][3] and compliance k[4].  GPIS quantities come from the
// query kernels: at a: dist (mean), sd (std), gmu = grad mean, gsd = grad std; at T: tdist, gT.  The contact normal
// n = gmu / (|gmu| + 1e-8) is a constant of the graph (gpis.py:72-73,83).  noise[9] replaces torch.rand_like(H) (:61).
CDX_HD void cdx_pg_row(const double (*a)[3], const double (*T)[3], const double* k, const double* dist, const double* sd,
                       const double (*gmu)[3], const double (*gsd)[3], const double* tdist, const double (*gT)[3],
                       const double* noise, const cdx_pg_const& C, double& loss, double* margin, double (*ga)[3],
                       double (*gTo)[3], double* gk) {
  CDX_PG_TICK(0);
  const int np = C.gravity ? 5 : 4;
  double S1[5][3], S2[5][3], w[5], n[4][3];
  for (int f = 0; f < 4; f++) {
    const double nn = sqrt(cdx_dot3(gmu[f], gmu[f])) + 1e-8;
    for (int r = 0; r < 3; r++) { S1[f][r] = a[f][r]; S2[f][r] = T[f][r]; n[f][r] = gmu[f][r] / nn; }
    w[f] = k[f];
  }
  S1[4][0] = C.com[0]; S1[4][1] = C.com[1]; S1[4][2] = C.com[2];
  S2[4][0] = 0.0; S2[4][1] = 0.0; S2[4][2] = -C.dummy_M;
  w[4] = C.dummy_k;
  // ---- Kabsch forward (optimal_transformation_batch, :49-69)
  double c1[3] = {0, 0, 0}, c2[3] = {0, 0, 0}, W = 0.0;
  for (int i = 0; i < np; i++) {
    for (int r = 0; r < 3; r++) { c1[r] += S1[i][r]; c2[r] += S2[i][r]; }
    W += w[i];
  }
  for (int r = 0; r < 3; r++) { c1[r] /= np; c2[r] /= np; }
  double A[5][3], Bm[5][3], H[9];
  for (int e = 0; e < 9; e++) H[e] = 1e-6 * noise[e];
  for (int i = 0; i < np; i++) {
    for (int r = 0; r < 3; r++) { A[i][r] = w[i] * (S1[i][r] - c1[r]); Bm[i][r] = w[i] * (S2[i][r] - c2[r]); }
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) H[r * 3 + c] += A[i][r] * Bm[i][c];
  }
  double U[9], S[3], V[9];
  CDX_PG_TICK(1);
  cdx_svd3(H, U, S, V);
  CDX_PG_TICK(2);
  const double dsg = (cdx_det3(V) * cdx_det3(U) < 0.0) ? -1.0 : 1.0;   // det(V U^T) < 0 -> flip V[:, -1] (:64-67)
  double R[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) R[r * 3 + c] = V[r * 3] * U[c * 3] + V[r * 3 + 1] * U[c * 3 + 1] + dsg * V[r * 3 + 2] * U[c * 3 + 2];
  const double lam[3] = {S[0], S[1], dsg * S[2]};
  double t[3] = {0, 0, 0}, RS1[5][3];
  for (int i = 0; i < np; i++) {
    cdx_mv(R, S1[i], RS1[i]);
    for (int r = 0; r < 3; r++) t[r] += w[i] * (S2[i][r] - RS1[i][r]);
  }
  for (int r = 0; r < 3; r++) t[r] /= W;
  // ---- force_eq_reward (:101-117) + force cost (:730)
  double diff[4][3], dn[4], dir[4][3], neq[4][3], ang[4], mraw[4], fn[4], reward = 0.0;
  for (int f = 0; f < 4; f++) {
    for (int r = 0; r < 3; r++) diff[f][r] = RS1[f][r] + t[r] - T[f][r];
    dn[f] = sqrt(cdx_dot3(diff[f], diff[f]));
    for (int r = 0; r < 3; r++) dir[f][r] = diff[f][r] / dn[f];
    cdx_mv(R, n[f], neq[f]);
    ang[f] = cdx_dot3(dir[f], neq[f]);
    mraw[f] = ang[f] - C.cos_mu;
    margin[f] = mraw[f] < -0.9999 ? -0.9999 : mraw[f];
    reward += 0.2 * log(ang[f] + 1.0) + 0.8 * log(margin[f] + 1.0);
    const double fx = k[f] * diff[f][0], fy = k[f] * diff[f][1], fz = k[f] * diff[f][2];
    fn[f] = sqrt(fx * fx + fy * fy + fz * fz);
  }
  double fmin = fn[0];
  for (int f = 1; f < 4; f++) fmin = fn[f] < fmin ? fn[f] : fmin;
  double ex[4], esum = 0.0, sm[4], z[4], force_cost = 0.0;
  for (int f = 0; f < 4; f++) { ex[f] = exp(fmin - fn[f]); esum += ex[f]; }
  for (int f = 0; f < 4; f++) {
    sm[f] = ex[f] / esum;
    z[f] = fn[f] * sm[f];
    force_cost -= (z[f] > C.force_clamp ? C.force_clamp : z[f]);
  }
  // ---- compute_contact_margin (:703-710), GPIS terms (:733-736)
  double vv[4][3], rr[4], dd[4][3], ang2[4], m2[4], cm = 0.0, dist_cost = 0.0, tar_cost = 0.0;
  int fstar = 0;
  double lsd[4];   // log(100 std_f): evaluated once per finger, compared and reused below (same values as re-evaluating)
  for (int f = 0; f < 4; f++) lsd[f] = C.var_sum ? sd[f] : log(100.0 * sd[f]);
  for (int f = 0; f < 4; f++) {
    for (int r = 0; r < 3; r++) vv[f][r] = a[f][r] - T[f][r];
    rr[f] = sqrt(cdx_dot3(vv[f], vv[f]));
    for (int r = 0; r < 3; r++) dd[f][r] = vv[f][r] / rr[f];
    ang2[f] = cdx_dot3(dd[f], n[f]);
    m2[f] = ang2[f] - C.cos_mu;
    if (C.w_cm != 0.0) cm += 0.1 * log(ang2[f] + 1.0) + 0.9 * log(m2[f] + 1.0);   // prob only (:703-710)
    dist_cost += fabs(dist[f]);
    tar_cost += tdist[f];
    if (lsd[f] > lsd[fstar]) fstar = f;
  }
  const double var_cost = C.var_sum ? C.uncertainty * (sd[0] + sd[1] + sd[2] + sd[3]) : C.uncertainty * lsd[fstar];
  // centre term of the Kin / GPIS optimizers (:474-477): w |mean_f a_f - mean_f T_f|
  double cd[3] = {0, 0, 0}, cn = 0.0;
  if (C.w_center != 0.0) {
    for (int f = 0; f < 4; f++)
      for (int r = 0; r < 3; r++) cd[r] += a[f][r] - T[f][r];
    for (int r = 0; r < 3; r++) cd[r] *= 0.25;
    cn = sqrt(cdx_dot3(cd, cd));
  }
  loss = -C.w_reward * reward + 1000.0 * dist_cost + C.w_tar * tar_cost + (-C.w_cm * cm) + force_cost + var_cost + C.w_center * cn;

  CDX_PG_TICK(3);
  // ================================================================================ reverse pass
  double gS1[5][3], gS2[5][3], gw[5], gR[9], gt[3] = {0, 0, 0};
  for (int i = 0; i < 5; i++) { gw[i] = 0.0; for (int r = 0; r < 3; r++) { gS1[i][r] = 0.0; gS2[i][r] = 0.0; } }
  for (int e = 0; e < 9; e++) gR[e] = 0.0;
  for (int f = 0; f < 4; f++) { gk[f] = 0.0; for (int r = 0; r < 3; r++) { ga[f][r] = 0.0; gTo[f][r] = 0.0; } }
  // softmin / clamp(max=10) adjoint:  force_cost = -sum_f min(z_f, 10),  z_f = fn_f sm_f
  double gz[4], gfn[4], zs = 0.0;
  for (int f = 0; f < 4; f++) { gz[f] = (z[f] <= C.force_clamp) ? -1.0 : 0.0; zs += gz[f] * z[f]; }
  for (int f = 0; f < 4; f++) gfn[f] = gz[f] * sm[f] * (1.0 - fn[f]) + sm[f] * zs;
  for (int f = 0; f < 4; f++) {
    // reward terms
    const double pass = (mraw[f] >= -0.9999) ? 1.0 : 0.0;
    const double Gang = -C.w_reward * (0.2 / (ang[f] + 1.0) + 0.8 * pass / (margin[f] + 1.0));
    double gdir[3], gdiff[3];
    for (int r = 0; r < 3; r++) gdir[r] = Gang * neq[f][r];
    // neq = R n
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) gR[r * 3 + c] += Gang * dir[f][r] * n[f][c];
    const double gd = cdx_dot3(gdir, dir[f]);
    for (int r = 0; r < 3; r++) gdiff[r] = (gdir[r] - gd * dir[f][r]) / dn[f];
    // force norm: fn = |k diff|
    const double Fv[3] = {-k[f] * diff[f][0], -k[f] * diff[f][1], -k[f] * diff[f][2]};
    double gF[3];
    for (int r = 0; r < 3; r++) gF[r] = gfn[f] * Fv[r] / fn[f];
    gk[f] += -(gF[0] * diff[f][0] + gF[1] * diff[f][1] + gF[2] * diff[f][2]);
    for (int r = 0; r < 3; r++) gdiff[r] += -k[f] * gF[r];
    // diff = R a + t - T
    for (int r = 0; r < 3; r++) {
      gTo[f][r] -= gdiff[r];
      gt[r] += gdiff[r];
      for (int c = 0; c < 3; c++) gR[r * 3 + c] += gdiff[r] * a[f][c];
    }
    double rt[3];
    cdx_mtv(R, gdiff, rt);
    for (int r = 0; r < 3; r++) gS1[f][r] += rt[r];
    // contact margin (static pose)
    if (C.w_cm != 0.0) {
      const double G2 = -C.w_cm * (0.1 / (ang2[f] + 1.0) + 0.9 / (m2[f] + 1.0));
      for (int r = 0; r < 3; r++) {
        const double gv = G2 * (n[f][r] - ang2[f] * dd[f][r]) / rr[f];
        ga[f][r] += gv;
        gTo[f][r] -= gv;
      }
    }
    // GPIS terms
    const double sg = dist[f] > 0.0 ? 1.0 : (dist[f] < 0.0 ? -1.0 : 0.0);
    for (int r = 0; r < 3; r++) {
      ga[f][r] += 1000.0 * sg * gmu[f][r];
      gTo[f][r] += C.w_tar * gT[f][r];
      if (C.w_center != 0.0 && cn > 0.0) {
        const double gc = C.w_center * 0.25 * cd[r] / cn;
        ga[f][r] += gc;
        gTo[f][r] -= gc;
      }
      if (C.var_sum) ga[f][r] += C.uncertainty * gsd[f][r];
    }
  }
  if (!C.var_sum)
    for (int r = 0; r < 3; r++) ga[fstar][r] += C.uncertainty / sd[fstar] * gsd[fstar][r];
  CDX_PG_TICK(4);
  // t = sum_i w_i (S2_i - R S1_i) / W
  double rgt[3];
  cdx_mtv(R, gt, rgt);
  for (int i = 0; i < np; i++) {
    double e[3];
    for (int r = 0; r < 3; r++) e[r] = S2[i][r] - RS1[i][r] - t[r];
    gw[i] += cdx_dot3(gt, e) / W;
    const double wi = w[i] / W;
    for (int r = 0; r < 3; r++) {
      gS2[i][r] += wi * gt[r];
      gS1[i][r] -= wi * rgt[r];
      for (int c = 0; c < 3; c++) gR[r * 3 + c] -= wi * gt[r] * S1[i][c];
    }
  }
  // R(H'): R H' = V diag(lam) V^T is symmetric  =>  dR = Om R,  (V^T Om V)_ij = -(X_ij - X_ji)/(lam_i+lam_j),
  // X = V^T R dH' V   =>   G_H = R^T V Z V^T,  Z_ij = -(Y_ij - Y_ji)/(lam_i + lam_j),  Y = V^T (gR R^T) V
  double Y0[9], Y1[9], Y[9], Z[9], GH[9], Vt[9], Rt[9], Tm[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) { Vt[r * 3 + c] = V[c * 3 + r]; Rt[r * 3 + c] = R[c * 3 + r]; }
  cdx_mm(gR, Rt, Y0);
  cdx_mm(Vt, Y0, Y1);
  cdx_mm(Y1, V, Y);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) Z[i * 3 + j] = (i == j) ? 0.0 : -(Y[i * 3 + j] - Y[j * 3 + i]) / (lam[i] + lam[j]);
  cdx_mm(V, Z, Tm);
  cdx_mm(Tm, Vt, Y0);
  cdx_mm(Rt, Y0, GH);
  CDX_PG_TICK(5);
  // H = sum_i A_i B_i^T,  A_i = w_i (S1_i - c1),  B_i = w_i (S2_i - c2)
  double sA[3] = {0, 0, 0}, sB[3] = {0, 0, 0}, gA[5][3], gB[5][3];
  for (int i = 0; i < np; i++) {
    cdx_mv(GH, Bm[i], gA[i]);
    cdx_mtv(GH, A[i], gB[i]);
    for (int r = 0; r < 3; r++) {
      gw[i] += gA[i][r] * (S1[i][r] - c1[r]) + gB[i][r] * (S2[i][r] - c2[r]);
      sA[r] += w[i] * gA[i][r];
      sB[r] += w[i] * gB[i][r];
    }
  }
  for (int i = 0; i < np; i++)
    for (int r = 0; r < 3; r++) {
      gS1[i][r] += w[i] * gA[i][r] - sA[r] / np;
      gS2[i][r] += w[i] * gB[i][r] - sB[r] / np;
    }
  for (int f = 0; f < 4; f++) {
    gk[f] += gw[f];
    for (int r = 0; r < 3; r++) { ga[f][r] += gS1[f][r]; gTo[f][r] += gS2[f][r]; }
  }
  CDX_PG_TICK(6);
}

// ------------------------------------------------------------------------------------------------ closure
#define CDX_MAX_COPIES 4
struct cdx_pg_setup {
  int P, U;                          // pregrasp copies (:623-625) and how many of them have distinct coefficients
  int uniq[CDX_MAX_COPIES];          // copy -> index of its GPIS query block (identical coefficient rows share one)
  double coeff[CDX_MAX_COPIES][4];   // float32 values (:644) widened
  double weight[CDX_MAX_COPIES];     // float32 values widened (:645)
  double ref_q[16];                  // float32 values widened (:634)
  int n_dofs;
  int optimize_palm;
  // optimizer variant (cdx_pregrasp_params.variant): 0 = ProbabilisticGraspOptimizer, 1 = KinGPISGraspOptimizer,
  // 2 = GPISGraspOptimizer (the "q" slot holds tip_pose [4,3] directly: no hand, no FK)
  int variant = 0;
  double w_ref = 10.0;               // weight of |q - ref_q| (:731 10, :478 20, GPIS variant 0)
  double w_pre = 5.0;                // weight of the pregrasp tips' own distance (:757-759; Kin / GPIS variants 0)
  cdx_pg_const C;
};

// GPIS query layout per candidate (written by the points kernel, read back here):
//   extended block u (U of them): 4 points all_tip[u][f]           -> dist, sd, gmu, gsd
//   mean block: 4 targets, 4 pregrasp tips, 1 palm position        -> mean, gmean
struct cdx_pg_gpis {
  const double* dist;   // [U*4]
  const double* sd;     // [U*4]
  const double* gmu;    // [U*4*3]
  const double* gsd;    // [U*4*3]
  const double* mean9;  // [9]
  const double* gmean9; // [9*3]
};

// closure() of ProbabilisticGraspOptimizer (optimize_pregrasp.py:741-769) for ONE candidate: total_loss, the
// weighted margin and d total_loss / d (q, compliance, target, palm) - what total_loss.sum().backward() leaves in
// the leaves' .grad.  tl = fingertips in the hand frame (float32 FK widened), J = cdx_fk_finger's Jacobians.
template <class NoiseFn>
CDX_HD void cdx_pg_closure(const cdx_pg_setup& P, const cdx_hand_dev& hand, const double* q, const double* k,
                           const double (*T)[3], const double* palm, const double (*tl)[3],
                           const float (*J)[3][CDX_MAX_CHAIN_DOF], const cdx_pg_gpis& G, NoiseFn noise_fn,
                           double& total, double* tmargin, double* gq, double* gk, double (*gT)[3], double* gpalm) {
  double Rw[9], dR0[9], dR1[9], dR2[9], tips[4][3];
  cdx_euler_xyz(palm + 3, Rw, dR0, dR1, dR2);
  for (int f = 0; f < 4; f++) {
    cdx_mv(Rw, tl[f], tips[f]);
    for (int r = 0; r < 3; r++) tips[f][r] += palm[r];
  }
  double refn = 0.0;
  for (int i = 0; i < P.n_dofs; i++) refn += (q[i] - P.ref_q[i]) * (q[i] - P.ref_q[i]);
  refn = sqrt(refn);
  const double ref_cost = P.w_ref * refn;
  total = 0.0;
  double gtips[4][3], wsum = 0.0;
  for (int f = 0; f < 4; f++) {
    tmargin[f] = 0.0; gk[f] = 0.0;
    for (int r = 0; r < 3; r++) { gtips[f][r] = 0.0; gT[f][r] = 0.0; }
  }
  double tdist[4], gTq[4][3];
  for (int f = 0; f < 4; f++) {
    tdist[f] = G.mean9[f];
    for (int r = 0; r < 3; r++) gTq[f][r] = G.gmean9[f * 3 + r];
  }
  for (int p = 0; p < P.P; p++) {
    const int u = P.uniq[p];
    double a[4][3], noise[9], loss, mg[4], ga[4][3], gTr[4][3], gkr[4];
    for (int f = 0; f < 4; f++)
      for (int r = 0; r < 3; r++) a[f][r] = T[f][r] + P.coeff[p][f] * (tips[f][r] - T[f][r]);
    for (int e = 0; e < 9; e++) noise[e] = noise_fn(p, e);
    cdx_pg_row(a, T, k, G.dist + u * 4, G.sd + u * 4, reinterpret_cast<const double (*)[3]>(G.gmu + u * 12),
               reinterpret_cast<const double (*)[3]>(G.gsd + u * 12), tdist, gTq, noise, P.C, loss, mg, ga, gTr, gkr);
    const double wp = P.weight[p];
    total += wp * (loss + ref_cost);
    wsum += wp;
    for (int f = 0; f < 4; f++) {
      tmargin[f] += wp * mg[f];
      gk[f] += wp * gkr[f];
      const double c = P.coeff[p][f];
      for (int r = 0; r < 3; r++) {
        gtips[f][r] += wp * c * ga[f][r];
        gT[f][r] += wp * ((1.0 - c) * ga[f][r] + gTr[f][r]);
      }
    }
  }
  // :757-763
  for (int f = 0; f < 4; f++) {
    total -= P.w_pre * G.mean9[4 + f];
    for (int r = 0; r < 3; r++) gtips[f][r] -= P.w_pre * G.gmean9[(4 + f) * 3 + r];
  }
  for (int r = 0; r < 6; r++) gpalm[r] = 0.0;
  if (P.optimize_palm) {
    const double pm = G.mean9[8];
    total += 1.0 / pm;
    for (int r = 0; r < 3; r++) gpalm[r] -= G.gmean9[24 + r] / (pm * pm);
  }
  // tips = Rw tl + palm_pos (:667-668)
  double gRw[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < P.n_dofs; i++) gq[i] = refn > 0.0 ? wsum * P.w_ref * (q[i] - P.ref_q[i]) / refn : 0.0;  // norm'(0) = 0 in torch
  for (int f = 0; f < 4; f++) {
    double gl[3];
    cdx_mtv(Rw, gtips[f], gl);
    for (int r = 0; r < 3; r++) {
      gpalm[r] += gtips[f][r];
      for (int c = 0; c < 3; c++) gRw[r * 3 + c] += gtips[f][r] * tl[f][c];
    }
    // back through .double() (:666) into the float32 FK graph
    const float g32[3] = {(float)gl[0], (float)gl[1], (float)gl[2]};
    const cdx_chain& ch = hand.finger[f];
    for (int c = 0; c < ch.ndof; c++) {
      const float gj = J[f][0][c] * g32[0] + J[f][1][c] * g32[1] + J[f][2][c] * g32[2];
      gq[ch.dofs[c]] += (double)gj;
    }
  }
  for (int e = 0; e < 9; e++) {
    gpalm[3] += gRw[e] * dR0[e];
    gpalm[4] += gRw[e] * dR1[e];
    gpalm[5] += gRw[e] * dR2[e];
  }
}

// ------------------------------------------------------------------------------------------------ host: tree -> chains
// Kinematic tree in URDF link order (robot_model.py:115-127; body i has parent[i] < i, the root has -1) to the four
// root->tip chains the kernels walk.  Returns 0 or a negative code (-1 bad tip, -2 chain too long, -3 too many dofs).
inline int cdx_build_chains(int n_bodies, const int* parent, const int* dof, const int* axis, const float* axis_sign,
                            const float* fixed_rot, const float* xyz, int n_dofs, const int* tip_body,
                            const float* tip_offset, cdx_hand_dev* out) {
  out->n_dofs = n_dofs;
  for (int f = 0; f < 4; f++) {
    cdx_chain& ch = out->finger[f];
    int ids[CDX_MAX_CHAIN], n = 0, b = tip_body[f];
    if (b < 0 || b >= n_bodies) return -1;
    while (b > 0) {
      if (n >= CDX_MAX_CHAIN) return -2;
      ids[n++] = b;
      if (parent[b] < 0 || parent[b] >= b) return -1;
      b = parent[b];
    }
    ch.n = n;
    ch.ndof = 0;
    for (int i = 0; i < n; i++) {
      const int id = ids[n - 1 - i];
      ch.dof[i] = dof[id];
      ch.axis[i] = axis[id];
      ch.sign[i] = axis_sign[id];
      for (int e = 0; e < 9; e++) ch.fixed_rot[i][e] = fixed_rot[id * 9 + e];
      for (int e = 0; e < 3; e++) ch.xyz[i][e] = xyz[id * 3 + e];
      if (dof[id] >= 0) {
        if (ch.ndof >= CDX_MAX_CHAIN_DOF || dof[id] >= n_dofs) return -3;
        ch.dofs[ch.ndof++] = dof[id];
      }
    }
    for (int i = n; i < CDX_MAX_CHAIN; i++) { ch.dof[i] = -1; ch.axis[i] = 0; ch.sign[i] = 1.f; }
    for (int e = 0; e < 3; e++) ch.offset[e] = tip_offset[f * 3 + e];
  }
  return 0;
}
