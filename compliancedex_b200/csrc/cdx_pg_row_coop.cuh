// cdx_pg_row (cdx_pg_math.cuh) evaluated by FOUR lanes, lane f owning finger f.
//
// The row's cost is its long dependent chains of FP64 divisions, square roots and logarithms (about 20 log, 12 sqrt and
// 40 divide per row), almost all of them per finger.  Here each lane evaluates those for its own finger only; the cheap
// 3x3 algebra that needs every finger (centroids, cross-covariance, Jacobi SVD, the Kabsch adjoint) is replicated on the
// four lanes, and per-finger scalars that feed a sum over fingers are gathered with width-4 shuffles and added in finger
// order, i.e. in the order cdx_pg_row adds them.  Same arguments and meaning as cdx_pg_row, except that margin / ga / gTo /
// gk come back for finger f only.  Device only.
#pragma once
#include "cdx_pg_math.cuh"

// Every shuffle of the step kernel names the FULL warp: the kernel keeps all 32 lanes on one path up to its last shuffle
// (idle copy quads and out-of-range candidate groups evaluate a duplicate row whose results are dropped).  With a run-time
// partial mask each double shuffle compiled into a WARPSYNC collective region (~13 instructions, 30 % of the kernel's SASS).
constexpr unsigned PG_FULL = 0xffffffffu;
// The four lanes of a row exchange per-finger scalars either by width-4 shuffles (a double shuffle is two SHFL plus lane
// arithmetic: the ~130 double shuffles of the step kernel were 2,600 of its 11,000 SASS instructions, and the kernel is
// instruction-FETCH bound - ncu: `no_instruction` is 53 % of its stall samples at 1,024 candidates, every warp walks 177 KB
// of straight-line code once) or, SMX, through a 448-byte shared-memory scratch per lane quad: one STS.64, one
// __syncwarp and two LDS.128 per gather.  Two single-value slots alternate (the next gather's barrier orders the reuse),
// followed by a [12][4] area for the batched gather of the reverse pass.  Values and summation order are the same.
constexpr int PG_XQ = 60;   // doubles per quad: 2 x 4 + 12 x 4, padded to 480 B (quads of a warp then start 96 B apart mod 128)
struct pg_xchg {
  double* q;   // this quad's scratch (SMX) or nullptr
  int n;       // gathers done so far (selects the slot)
};
template <bool SMX>
__device__ __forceinline__ void pg_gather4(pg_xchg& x, int f, double v, double* out) {
  if (SMX) {
    double* slot = x.q + (x.n & 1) * 4;
    x.n++;
    slot[f] = v;
    __syncwarp();
    const double2 a = *reinterpret_cast<const double2*>(slot), b = *reinterpret_cast<const double2*>(slot + 2);
    out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++) out[j] = __shfl_sync(PG_FULL, v, j, 4);
  }
}
template <bool SMX>
__device__ __forceinline__ double pg_sum4(pg_xchg& x, int f, double v) {   // ((v0 + v1) + v2) + v3, identical on the four lanes
  double g[4];
  pg_gather4<SMX>(x, f, v, g);
  return ((g[0] + g[1]) + g[2]) + g[3];
}
// NV finger-ordered sums at once (NV <= 12)
template <bool SMX, int NV>
__device__ __forceinline__ void pg_sum4_batch(pg_xchg& x, int f, const double* v, double* out) {
  if (SMX) {
    double* area = x.q + 8;
#pragma unroll
    for (int e = 0; e < NV; e++) area[e * 4 + f] = v[e];
    __syncwarp();
#pragma unroll
    for (int e = 0; e < NV; e++) {
      const double2 a = *reinterpret_cast<const double2*>(area + e * 4), b = *reinterpret_cast<const double2*>(area + e * 4 + 2);
      out[e] = ((a.x + a.y) + b.x) + b.y;
    }
  } else {
#pragma unroll
    for (int e = 0; e < NV; e++) out[e] = pg_sum4<false>(x, f, v[e]);
  }
}

template <bool SMX>
__device__ __forceinline__ void cdx_pg_row_coop(int f, pg_xchg& xc, const double (*a)[3], const double (*T)[3],
                                                const double* k, const double* dist, const double* sd,
                                                const double (*gmu)[3], const double (*gsd)[3], const double* tdist,
                                                const double (*gT)[3], const double* noise, const cdx_pg_const& C,
                                                double& loss, double& margin_f, double* ga, double* gTo, double& gk_f) {
  const int np = C.gravity ? 5 : 4;
  // ---- replicated: point sets, centroids, cross-covariance, Kabsch rotation (:49-69)
  double S1[5][3], S2[5][3], w[5];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    for (int r = 0; r < 3; r++) { S1[i][r] = a[i][r]; S2[i][r] = T[i][r]; }
    w[i] = k[i];
  }
  S1[4][0] = C.com[0]; S1[4][1] = C.com[1]; S1[4][2] = C.com[2];
  S2[4][0] = 0.0; S2[4][1] = 0.0; S2[4][2] = -C.dummy_M;
  w[4] = C.dummy_k;
  double c1[3] = {0, 0, 0}, c2[3] = {0, 0, 0}, W = 0.0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    if (i >= np) continue;   // compile-time subscripts: the point sets stay in registers
    for (int r = 0; r < 3; r++) { c1[r] += S1[i][r]; c2[r] += S2[i][r]; }
    W += w[i];
  }
  for (int r = 0; r < 3; r++) { c1[r] /= np; c2[r] /= np; }
  double A[5][3], Bm[5][3], H[9];
  for (int e = 0; e < 9; e++) H[e] = 1e-6 * noise[e];
#pragma unroll
  for (int i = 0; i < 5; i++) {
    if (i >= np) continue;   // compile-time subscripts: the point sets stay in registers
    for (int r = 0; r < 3; r++) { A[i][r] = w[i] * (S1[i][r] - c1[r]); Bm[i][r] = w[i] * (S2[i][r] - c2[r]); }
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) H[r * 3 + c] += A[i][r] * Bm[i][c];
  }
  // my finger's contact normal (gpis.py:72-73,83)
  double n[3];
  {
    const double nn = sqrt(cdx_dot3(gmu[f], gmu[f])) + 1e-8;
    for (int r = 0; r < 3; r++) n[r] = gmu[f][r] / nn;
  }
  double U[9], S[3], V[9];
  cdx_svd3(H, U, S, V);
  const double dsg = (cdx_det3(V) * cdx_det3(U) < 0.0) ? -1.0 : 1.0;
  double R[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) R[r * 3 + c] = V[r * 3] * U[c * 3] + V[r * 3 + 1] * U[c * 3 + 1] + dsg * V[r * 3 + 2] * U[c * 3 + 2];
  const double lam[3] = {S[0], S[1], dsg * S[2]};
  double t[3] = {0, 0, 0}, RS1[5][3];
#pragma unroll
  for (int i = 0; i < 5; i++) {
    if (i >= np) continue;   // compile-time subscripts: the point sets stay in registers
    cdx_mv(R, S1[i], RS1[i]);
    for (int r = 0; r < 3; r++) t[r] += w[i] * (S2[i][r] - RS1[i][r]);
  }
  for (int r = 0; r < 3; r++) t[r] /= W;
  // my finger's row of the point sets (dynamic f: selected once, kept in registers)
  double af[3], Tf[3], RS1f[3], Af[3], Bf[3];
#pragma unroll
  for (int i = 0; i < 4; i++)
    if (i == f)
      for (int r = 0; r < 3; r++) { af[r] = S1[i][r]; Tf[r] = S2[i][r]; RS1f[r] = RS1[i][r]; Af[r] = A[i][r]; Bf[r] = Bm[i][r]; }
  const double kf = k[f];
  // ---- force_eq_reward (:101-117) + force cost (:730), finger f
  double diff[3], dir[3], neq[3];
  for (int r = 0; r < 3; r++) diff[r] = RS1f[r] + t[r] - Tf[r];
  const double dn = sqrt(cdx_dot3(diff, diff));
  for (int r = 0; r < 3; r++) dir[r] = diff[r] / dn;
  cdx_mv(R, n, neq);
  const double ang = cdx_dot3(dir, neq);
  const double mraw = ang - C.cos_mu;
  const double margin = mraw < -0.9999 ? -0.9999 : mraw;
  margin_f = margin;
  const double reward = pg_sum4<SMX>(xc, f, 0.2 * log(ang + 1.0) + 0.8 * log(margin + 1.0));
  double fnf;
  {
    const double fx = kf * diff[0], fy = kf * diff[1], fz = kf * diff[2];
    fnf = sqrt(fx * fx + fy * fy + fz * fz);
  }
  double fn4[4];
  pg_gather4<SMX>(xc, f, fnf, fn4);
  double fmin = fn4[0];
  for (int i = 1; i < 4; i++) fmin = fn4[i] < fmin ? fn4[i] : fmin;
  const double exf = exp(fmin - fnf);
  const double esum = pg_sum4<SMX>(xc, f, exf);
  const double smf = exf / esum;
  const double zf = fnf * smf;
  const double force_cost = -pg_sum4<SMX>(xc, f, zf > C.force_clamp ? C.force_clamp : zf);
  // ---- compute_contact_margin (:703-710), GPIS terms (:733-736)
  double vv[3], dd[3];
  for (int r = 0; r < 3; r++) vv[r] = af[r] - Tf[r];
  const double rr = sqrt(cdx_dot3(vv, vv));
  for (int r = 0; r < 3; r++) dd[r] = vv[r] / rr;
  const double ang2 = cdx_dot3(dd, n);
  const double m2 = ang2 - C.cos_mu;
  double cm = 0.0;
  if (C.w_cm != 0.0) cm = pg_sum4<SMX>(xc, f, 0.1 * log(ang2 + 1.0) + 0.9 * log(m2 + 1.0));
  double dist_cost = 0.0, tar_cost = 0.0;
  for (int i = 0; i < 4; i++) { dist_cost += fabs(dist[i]); tar_cost += tdist[i]; }
  const double sdf = sd[f];
  double lsd[4];
  pg_gather4<SMX>(xc, f, C.var_sum ? sdf : log(100.0 * sdf), lsd);
  int fstar = 0;
  for (int i = 1; i < 4; i++)
    if (lsd[i] > lsd[fstar]) fstar = i;
  const double var_cost = C.var_sum ? C.uncertainty * (sd[0] + sd[1] + sd[2] + sd[3]) : C.uncertainty * lsd[fstar];
  double cd[3] = {0, 0, 0}, cn = 0.0;
  if (C.w_center != 0.0) {
    for (int i = 0; i < 4; i++)
      for (int r = 0; r < 3; r++) cd[r] += S1[i][r] - S2[i][r];
    for (int r = 0; r < 3; r++) cd[r] *= 0.25;
    cn = sqrt(cdx_dot3(cd, cd));
  }
  loss = -C.w_reward * reward + 1000.0 * dist_cost + C.w_tar * tar_cost + (-C.w_cm * cm) + force_cost + var_cost + C.w_center * cn;

  // ================================================================================ reverse pass, finger f
  const double gzf = (zf <= C.force_clamp) ? -1.0 : 0.0;
  const double zs = pg_sum4<SMX>(xc, f, gzf * zf);
  const double gfn = gzf * smf * (1.0 - fnf) + smf * zs;
  double gS1f[3], gS2f[3], Zr[9], gdiff[3];
  for (int r = 0; r < 3; r++) { ga[r] = 0.0; gTo[r] = 0.0; }
  {
    const double pass = (mraw >= -0.9999) ? 1.0 : 0.0;
    const double Gang = -C.w_reward * (0.2 / (ang + 1.0) + 0.8 * pass / (margin + 1.0));
    double gdir[3];
    for (int r = 0; r < 3; r++) gdir[r] = Gang * neq[r];
    const double gd = cdx_dot3(gdir, dir);
    for (int r = 0; r < 3; r++) gdiff[r] = (gdir[r] - gd * dir[r]) / dn;
    const double Fv[3] = {-kf * diff[0], -kf * diff[1], -kf * diff[2]};
    double gF[3];
    for (int r = 0; r < 3; r++) gF[r] = gfn * Fv[r] / fnf;
    gk_f = -(gF[0] * diff[0] + gF[1] * diff[1] + gF[2] * diff[2]);
    for (int r = 0; r < 3; r++) gdiff[r] += -kf * gF[r];
    // neq = R n and diff = R a + t - T: this finger's share of dL/dR
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) Zr[r * 3 + c] = Gang * dir[r] * n[c] + gdiff[r] * af[c];
    for (int r = 0; r < 3; r++) gTo[r] -= gdiff[r];
    cdx_mtv(R, gdiff, gS1f);
    if (C.w_cm != 0.0) {
      const double G2 = -C.w_cm * (0.1 / (ang2 + 1.0) + 0.9 / (m2 + 1.0));
      for (int r = 0; r < 3; r++) {
        const double gv = G2 * (n[r] - ang2 * dd[r]) / rr;
        ga[r] += gv;
        gTo[r] -= gv;
      }
    }
    const double df = dist[f];
    const double sg = df > 0.0 ? 1.0 : (df < 0.0 ? -1.0 : 0.0);
    for (int r = 0; r < 3; r++) {
      ga[r] += 1000.0 * sg * gmu[f][r];
      gTo[r] += C.w_tar * gT[f][r];
      if (C.w_center != 0.0 && cn > 0.0) {
        const double gc = C.w_center * 0.25 * cd[r] / cn;
        ga[r] += gc;
        gTo[r] -= gc;
      }
      if (C.var_sum) ga[r] += C.uncertainty * gsd[f][r];
    }
    if (!C.var_sum && f == fstar)
      for (int r = 0; r < 3; r++) ga[r] += C.uncertainty / sdf * gsd[f][r];
  }
  double gR[9], gt[3];
  {
    double zin[12], zout[12];
    for (int e = 0; e < 9; e++) zin[e] = Zr[e];
    for (int r = 0; r < 3; r++) zin[9 + r] = gdiff[r];
    pg_sum4_batch<SMX, 12>(xc, f, zin, zout);
    for (int e = 0; e < 9; e++) gR[e] = zout[e];
    for (int r = 0; r < 3; r++) gt[r] = zout[9 + r];
  }
  // t = sum_i w_i (S2_i - R S1_i) / W
  double rgt[3], gwf;
  cdx_mtv(R, gt, rgt);
  {
    double e[3];
    for (int r = 0; r < 3; r++) e[r] = Tf[r] - RS1f[r] - t[r];
    gwf = cdx_dot3(gt, e) / W;
    const double wi = kf / W;
    for (int r = 0; r < 3; r++) { gS2f[r] = wi * gt[r]; gS1f[r] -= wi * rgt[r]; }
  }
#pragma unroll
  for (int i = 0; i < 5; i++) {
    if (i >= np) continue;   // compile-time subscripts: the point sets stay in registers
    const double wi = w[i] / W;
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) gR[r * 3 + c] -= wi * gt[r] * S1[i][c];
  }
  // R(H'): G_H = R^T V Z V^T,  Z_ij = -(Y_ij - Y_ji)/(lam_i + lam_j),  Y = V^T (gR R^T) V   (see cdx_pg_row)
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
  // H = sum_i A_i B_i^T: the centroid terms need every point (replicated), the rest is this finger's
  double sA[3] = {0, 0, 0}, sB[3] = {0, 0, 0};
#pragma unroll
  for (int i = 0; i < 5; i++) {
    if (i >= np) continue;   // compile-time subscripts: the point sets stay in registers
    double gAi[3], gBi[3];
    cdx_mv(GH, Bm[i], gAi);
    cdx_mtv(GH, A[i], gBi);
    for (int r = 0; r < 3; r++) { sA[r] += w[i] * gAi[r]; sB[r] += w[i] * gBi[r]; }
  }
  double gAf[3], gBf[3];
  cdx_mv(GH, Bf, gAf);
  cdx_mtv(GH, Af, gBf);
  for (int r = 0; r < 3; r++) {
    gwf += gAf[r] * (af[r] - c1[r]) + gBf[r] * (Tf[r] - c2[r]);
    gS1f[r] += kf * gAf[r] - sA[r] / np;
    gS2f[r] += kf * gBf[r] - sB[r] / np;
  }
  gk_f += gwf;
  for (int r = 0; r < 3; r++) { ga[r] += gS1f[r]; gTo[r] += gS2f[r]; }
}
