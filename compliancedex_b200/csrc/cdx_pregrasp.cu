// Batched probabilistic pregrasp optimisation (ProbabilisticGraspOptimizer, optimize_pregrasp.py:614-839) without
// autograd and without host synchronisation.  One optimiser iteration is three launches:
//   gpis_var_kernel<EXT> mean, grad mean, std, grad std at the blended contact points (cdx_query.cu, FP64 DMMA)
//   gpis_mean_kernel     mean, grad mean at targets / tips / palm (cdx_query.cu, TMA-staged training set), on a forked
//                        stream so that it overlaps the extended query
//   pg_step_kernel       4 lanes per candidate: Kabsch (3x3 Jacobi SVD) equilibrium per pregrasp copy, all cost terms,
//                        ANALYTIC reverse pass (cdx_pg_math.cuh), best-iterate tracking (:822-829), Adam (:831), the
//                        clamps (:833-834), then float32 FK + reference-compatible Jacobian of the NEW iterate and its
//                        GPIS query points (pregrasp blend :747-750, targets, tips, palm position)
// pg_points_kernel does the FK / query points once for the initial iterate.
#include <math.h>
#include <stdlib.h>
#include <vector>
#include "cdx_internal.cuh"
#include "cdx_noise.cuh"
#ifdef CDX_PG_TICKS   // section timing of the step kernel (scripts/pg_ticks.py): thread 0 of CTA 0 records clock64()
__device__ long long cdx_pg_ticks[16];
#ifdef __CUDA_ARCH__
#define CDX_PG_TICK(i) do { if (threadIdx.x == 0 && blockIdx.x == 0) cdx_pg_ticks[i] = clock64(); } while (0)
#endif
#endif
#include "cdx_pg_math.cuh"
#include "cdx_pg_row_coop.cuh"
#ifdef CDX_PG_TICKS
extern "C" int cdx_debug_pg_ticks(long long* out16) {
  return cudaMemcpyFromSymbol(out16, cdx_pg_ticks, sizeof(long long) * 16) == cudaSuccess ? 0 : -1;
}
#endif

struct cdx_pg_state {
  int64_t B = 0, cap = 0;
  cdx_pregrasp_params prm;
  cdx_pg_setup setup;
  int U = 1;
  // iterate, Adam moments, best iterate (all [B, .] row-major f64)
  double *q = nullptr, *k = nullptr, *T = nullptr, *palm = nullptr;
  double *m1 = nullptr, *m2 = nullptr;          // [B,38] each: q16 | k4 | T12 | palm6
  double *oq = nullptr, *ok = nullptr, *oT = nullptr, *opalm = nullptr, *omargin = nullptr, *ovalue = nullptr;
  // per-iteration scratch
  double* tl = nullptr;      // [B,4,3] fingertips in the hand frame
  float* J = nullptr;        // [B,4,3,CDX_MAX_CHAIN_DOF]
  double* Xext = nullptr;    // [B,U,4,3]
  double* Xmean = nullptr;   // [B,9,3]
  double *dist = nullptr, *sd = nullptr, *gmu = nullptr, *gsd = nullptr;   // [B*U*4], [.. ,3]
  double *mean9 = nullptr, *gmean9 = nullptr;                               // [B*9], [B*9,3]
  double *loss = nullptr, *margin = nullptr, *grad = nullptr;               // [B], [B,4], [B,38]
  // Kin / GPIS variants: `opt_margin = margin` of the WHOLE batch whenever any candidate improved (:491-494)
  double *mring = nullptr, *latched = nullptr;   // [2][B,4] margins of the last two iterations, [B,4] latched copy
  int* latch = nullptr;                          // [0],[1]: iteration that improved (by parity), [2]: latched iteration
  int it_host = 0;           // absolute iteration index (keys the noise stream, Adam bias correction, s > 20)
  bool points_valid = false; // tl / J / Xext / Xmean describe the current iterate
  cudaStream_t side = nullptr;              // forked stream for the mean-only GPIS query
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // candidate groups 1.. of cdx_pregrasp_run (group 0 uses the caller's stream and the members above): each group walks
  // its own iterations on its own pair of streams, so one group's step kernel runs beside another group's GPIS queries
  static constexpr int MAX_GROUPS = 4;
  cudaStream_t gmain[MAX_GROUPS] = {}, gside[MAX_GROUPS] = {};
  cudaEvent_t gfork[MAX_GROUPS] = {}, gjoin[MAX_GROUPS] = {}, gend[MAX_GROUPS] = {};
  cudaEvent_t ev_begin = nullptr;
  double* adam_tab = nullptr;  // [tab_cap][2]: 1 - beta1^t, sqrt(1 - beta2^t) for t = index + 1
  int tab_cap = 0;
  cdx_hand_dev* hand_dev = nullptr;
  int hand_n_dofs = 0;
  bool prof = false;
  std::vector<cudaEvent_t> prof_ev;   // 6 events per profiled (iteration, group)
  uint64_t hand_id = 0;               // hash of the loaded kinematic chains
  uint64_t generation = 0;            // bumped by every cdx_pregrasp_init: lets a caller detect that another optimizer took the state over
  int* it_dev = nullptr;              // device copy of it_host: the base iteration the nodes of a captured graph add their offset to
  long long* nan_dev = nullptr;       // [0] candidate-iterations whose total_loss was NaN, [1] first such iteration (INT_MAX: none)
  int tune_groups = 0, tune_occ = 0;  // cdx_pregrasp_set_tuning: 0 = automatic
  bool use_graph = true;
  // the iteration loop captured once as a CUDA graph of g_K iterations (all groups, all streams) and re-launched
  cudaGraphExec_t gexec = nullptr;
  cudaStream_t cap_stream = nullptr;
  int g_K = 0;
  uint64_t g_key = 0;
};

namespace {

constexpr int NPAR = 38;  // q16 | k4 | T12 | palm6

struct pg_lr { double v[5]; };

// FK of finger f for candidate b at joint angles q (float64, n_dofs entries), palm pose palm6, target Tf; writes the
// per-iteration scratch (tl, J) and the GPIS query points of this finger (pregrasp blend :747-750, target, tip, palm).
template <int V>   // V = optimizer variant (cdx_pregrasp_params.variant), compile-time so that the prob path carries no variant code
__device__ __forceinline__ void pg_emit_points(const cdx_hand_dev* __restrict__ hand, const cdx_pg_setup& S, int U, int64_t b,
                                               int f, const double* q, const double* Tf, const double* palm6,
                                               double* __restrict__ tl, float* __restrict__ J, double* __restrict__ Xext,
                                               double* __restrict__ Xmean, double* __restrict__ tips_out,
                                               const float* q32s = nullptr) {
  const int64_t gid = b * 4 + f;
  double tlf[3], tw[3];
  float Jf[3][CDX_MAX_CHAIN_DOF];
  if (V == 2) {   // GPISGraspOptimizer: tip_pose itself is the parameter (:345,:361)
    for (int r = 0; r < 3; r++) { tlf[r] = q[f * 3 + r]; tw[r] = tlf[r]; }
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++) Jf[r][c] = 0.f;
  } else {
    float q32l[16], tip[3];
    if (!q32s) {
#pragma unroll
      for (int i = 0; i < 16; i++) q32l[i] = i < S.n_dofs ? (float)q[i] : 0.f;               // q.float(), :665
    }
    const float* q32 = q32s ? q32s : q32l;   // step kernel: the float32 angles are staged in shared memory (run-time subscripts)
    // Kin variant: q stays float64 and the chain is composed recursively (:436, robot_model.py:243-251)
    cdx_fk_finger<V == 1>(hand->finger[f], q32, tip, Jf, V == 1 ? q : nullptr);
    for (int r = 0; r < 3; r++) tlf[r] = (double)tip[r];                                     // .double(), :666
    double Rw[9];
    cdx_euler_xyz(palm6 + 3, Rw, nullptr, nullptr, nullptr);
    cdx_mv(Rw, tlf, tw);
    for (int r = 0; r < 3; r++) tw[r] += palm6[r];
  }
  if (tips_out) {
    for (int r = 0; r < 3; r++) tips_out[gid * 3 + r] = tw[r];
    return;
  }
  for (int r = 0; r < 3; r++) {
    tl[gid * 3 + r] = tlf[r];
    for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++) J[(gid * 3 + r) * CDX_MAX_CHAIN_DOF + c] = Jf[r][c];
  }
  int seen = 0;
  for (int p = 0; p < S.P; p++) {
    const int u = S.uniq[p];
    if (u < seen) continue;   // this coefficient row already has its query block
    seen = u + 1;
    double* x = Xext + ((b * U + u) * 4 + f) * 3;
    for (int r = 0; r < 3; r++) x[r] = V ? tw[r] : Tf[r] + S.coeff[p][f] * (tw[r] - Tf[r]);   // :750 / tips (:462)
  }
  double* xm = Xmean + b * 27;
  for (int r = 0; r < 3; r++) {
    xm[f * 3 + r] = Tf[r];
    xm[(4 + f) * 3 + r] = tw[r];
    if (f == 0) xm[24 + r] = palm6[r];
  }
}

template <int V>
__global__ void __launch_bounds__(128)
pg_points_kernel(const cdx_hand_dev* __restrict__ hand, int64_t B, int U, cdx_pg_setup S, const double* __restrict__ q,
                 const double* __restrict__ T, const double* __restrict__ palm, double* __restrict__ tl,
                 float* __restrict__ J, double* __restrict__ Xext, double* __restrict__ Xmean,
                 double* __restrict__ tips_out) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * 4) return;
  const int64_t b = gid >> 2;
  pg_emit_points<V>(hand, S, U, b, (int)(gid & 3), q + b * S.n_dofs, T ? T + gid * 3 : nullptr, palm + b * 6, tl, J, Xext,
                 Xmean, tips_out);
}

__global__ void pg_fill_kernel(double* p, int64_t n, double v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// NaN-propagating clamps (torch.clamp keeps NaN; fmax/fmin would not)
__device__ __forceinline__ double clamp_lo(double x, double lo) { return x < lo ? lo : x; }
__device__ __forceinline__ double clamp_hi(double x, double hi) { return x > hi ? hi : x; }

#ifndef CDX_PG_CPB
#define CDX_PG_CPB 8
#endif
constexpr int CPB = CDX_PG_CPB;   // candidates per CTA: 16 lanes each
constexpr int NRED = 33;    // total | margin 4 | gk 4 | gtips 12 | gT 12
// the step kernel runs one or two warps per SM, so every global load's latency is exposed: everything it will need (GPIS
// outputs, moments, its finger's Jacobian) is requested with cp.async at entry and lands in shared memory while the
// prologue (Euler angles, noise keys) runs
constexpr int NGP = 16 + 16 + 48 + 48 + 9 + 27;   // dist | sd | gmu | gsd (up to 4 copy blocks) | mean9 | gmean9
__device__ __forceinline__ void pg_cp8(void* smem, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void pg_cp4(void* smem, const void* g) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void pg_cp_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// One closure (+ optional optimiser step) per candidate, spread over a group of 16 lanes (p, f) = (lane >> 2, lane & 3):
//   the four lanes of copy p < P evaluate pregrasp copy p together (cdx_pg_row_coop: Kabsch equilibrium, costs, analytic
//   reverse pass, lane f owning finger f) - the copies are independent given the GPIS outputs - and the weighted sum over
//   copies (:755) is a 2-step shuffle reduction over p;
//   lane (0, f) then owns finger f: chain rule through the palm transform and the float32 FK Jacobian, and - when stepping -
//   the FK and query points of the NEW iterate, so the next iteration needs no separate points launch;
//   the 38 parameters are spread over the 16 lanes for best-iterate tracking (:822-829), Adam (:831) and the clamps (:833-834).
// OCC = resident CTAs per SM the register allocation is sized for: 2 (255 registers) for launches of at most one wave, where
// only a warp's own latency counts; 3 (168 registers) for larger launches, where the kernel is paced by how many of its
// latency-bound warps an SM can hold.  Same source, same arithmetic (tests compare the two bit for bit).
template <int V, int OCC>
__global__ void __launch_bounds__(CPB * 16, OCC)
pg_step_kernel(const cdx_hand_dev* __restrict__ hand, int64_t B, int64_t b0, int64_t b1, int U, cdx_pg_setup S,
               cdx_pregrasp_params prm,
               double* __restrict__ q, double* __restrict__ k, double* __restrict__ T, double* __restrict__ palm,
               double* __restrict__ tl, float* __restrict__ J, const double* __restrict__ dist,
               const double* __restrict__ sd, const double* __restrict__ gmu, const double* __restrict__ gsd,
               const double* __restrict__ mean9, const double* __restrict__ gmean9, double* __restrict__ loss,
               double* __restrict__ margin, double* __restrict__ grad, int do_step, const int* __restrict__ it_base, int it_off,
               long long* __restrict__ nan_rec, const double* __restrict__ adam_tab, double* __restrict__ m1, double* __restrict__ m2,
               double* __restrict__ oq, double* __restrict__ ok, double* __restrict__ oT, double* __restrict__ opalm,
               double* __restrict__ omargin, double* __restrict__ ovalue, double* __restrict__ Xext,
               double* __restrict__ Xmean, double* __restrict__ mring, double* __restrict__ latched,
               int* __restrict__ latch) {
  __shared__ double sp[CPB][NPAR];      // the candidate's parameters: q16 | k4 | T12 | palm6
  __shared__ double sr[CPB][NRED];      // results reduced over the copies
  __shared__ double sg[CPB][NPAR];      // gradient
  __shared__ double sq[CPB][NGP];       // GPIS outputs of this candidate
  __shared__ double sm1[CPB][NPAR], sm2[CPB][NPAR];
  extern __shared__ __align__(16) double sx[];         // [CPB * 4][PG_XQ] quad exchange scratch of the row evaluation (prob variant; dynamic:
                                                       // the static arrays above and below already fill 41 KB of the 48 KB static limit)
  __shared__ __align__(8) float sJ[CPB][4][3 * CDX_MAX_CHAIN_DOF];
  __shared__ float sq32[V != 2 ? CPB : 1][16];        // float32 joint angles of the new iterate
  constexpr int HW = (int)((sizeof(cdx_hand_dev) + 3) / 4);
  __shared__ __align__(16) unsigned shand[CPB][HW];   // the group's own copy of the finger chains (FK walks them body by body)
  CDX_PG_TICK(7);
  const int sub = threadIdx.x & 15, p = sub >> 2, l = sub & 3, cl = threadIdx.x >> 4;
  // absolute iteration index: inside a captured graph the base lives in device memory, so one graph serves every chunk
  const int it = (it_base ? *it_base : 0) + it_off;
  const int64_t braw = b0 + (int64_t)blockIdx.x * CPB + cl;   // this launch covers candidates [b0, b1) of the B
  if (braw - (cl & 1) >= b1) return;    // whole WARPS (two candidates) leave together
  // a warp's second 16-lane group may lie past the end: it stays on the warp's path until the last shuffle (all shuffles
  // name the full warp, see PG_FULL) evaluating a duplicate of the last candidate, writes nothing and leaves after that
  const bool live = braw < b1;
  const int64_t b = live ? braw : b1 - 1;
  const unsigned gmask = 0xFFFFu << ((threadIdx.x & 31) & 16);
  const unsigned fmask = 0xFu << ((threadIdx.x & 31) & ~3);
  const int nd = S.n_dofs;
  {
    double* g = sq[cl];
    for (int i = sub; i < U * 4; i += 16) { pg_cp8(g + i, dist + b * U * 4 + i); pg_cp8(g + 16 + i, sd + b * U * 4 + i); }
    for (int i = sub; i < U * 12; i += 16) { pg_cp8(g + 32 + i, gmu + b * U * 12 + i); pg_cp8(g + 80 + i, gsd + b * U * 12 + i); }
    if (sub < 9) pg_cp8(g + 128 + sub, mean9 + b * 9 + sub);
    for (int i = sub; i < 27; i += 16) pg_cp8(g + 137 + i, gmean9 + b * 27 + i);
    if (V != 2 && p == 1) {
      const float* Jf = J + (b * 4 + l) * 3 * CDX_MAX_CHAIN_DOF;
      for (int i = 0; i < 3 * CDX_MAX_CHAIN_DOF / 2; i++) pg_cp8(&sJ[cl][l][2 * i], Jf + 2 * i);
    }
    if (do_step)
      for (int i = sub; i < NPAR; i += 16) { pg_cp8(&sm1[cl][i], m1 + b * NPAR + i); pg_cp8(&sm2[cl][i], m2 + b * NPAR + i); }
    if (V != 2)
      for (int i = sub; i < HW; i += 16) pg_cp4(&shand[cl][i], reinterpret_cast<const unsigned*>(hand) + i);
  }
  const cdx_hand_dev* hs = V != 2 ? reinterpret_cast<const cdx_hand_dev*>(shand[cl]) : hand;
  double oval = 0.0, bc1 = 0.0, bc2s = 0.0;
  if (do_step) { oval = ovalue[b]; bc1 = adam_tab[2 * it]; bc2s = adam_tab[2 * it + 1]; }
  double* par = sp[cl];
  for (int i = sub; i < NPAR; i += 16)
    par[i] = i < 16 ? (i < nd ? q[b * nd + i] : 0.0) : (i < 20 ? k[b * 4 + i - 16] : (i < 32 ? T[b * 12 + i - 20] : palm[b * 6 + i - 32]));
  sg[cl][sub] = 0.0;
  __syncwarp(gmask);
  const double (*Tp)[3] = reinterpret_cast<const double (*)[3]>(par + 20);
  const double (*tlp)[3] = reinterpret_cast<const double (*)[3]>(tl + b * 12);
  double Rw[9], dR0[9], dR1[9], dR2[9], tips[4][3];
  cdx_euler_xyz(par + 35, Rw, dR0, dR1, dR2);
#pragma unroll
  for (int f = 0; f < 4; f++) {
    cdx_mv(Rw, tlp[f], tips[f]);
    for (int r = 0; r < 3; r++) tips[f][r] += par[32 + r];
    if (V == 2)
      for (int r = 0; r < 3; r++) tips[f][r] = par[f * 3 + r];
  }
  double refn = 0.0;
  if (S.w_ref != 0.0) {
    if (sub < nd) refn = (par[sub] - S.ref_q[sub]) * (par[sub] - S.ref_q[sub]);
    refn += __shfl_xor_sync(PG_FULL, refn, 4);
    refn += __shfl_xor_sync(PG_FULL, refn, 8);
    refn += __shfl_xor_sync(PG_FULL, refn, 1);
    refn += __shfl_xor_sync(PG_FULL, refn, 2);
    refn = sqrt(refn);
  }
  // Kin / GPIS variants: the margin of the previous iteration becomes the latched opt_margin if anything improved then
  if (V != 0 && live && do_step && it > 0 && latch[(it - 1) & 1] == it - 1 && p == 0) {
    latched[b * 4 + l] = mring[((it - 1) & 1) * B * 4 + b * 4 + l];
    if (b == 0 && l == 0) latch[2] = it - 1;
  }
  const double* m9 = sq[cl] + 128;
  const double* g9 = sq[cl] + 137;
  double wsum = 0.0;
  for (int c = 0; c < S.P; c++) wsum += S.weight[c];
  // noise[9] of copy p (replaces torch.rand_like(H), :61): lane f draws elements f, f+4, f+8 and the four lanes exchange
  // lane quads p >= P have no copy of their own: they evaluate copy P-1 again (results dropped) so that the warp stays converged
  const int pc = p < S.P ? p : S.P - 1;
  double noise[9];
  {
    const int64_t gb = prm.env_offset + b;
    double mine[3];
#pragma unroll
    for (int j = 0; j < 3; j++) mine[j] = (l + 4 * j < 9) ? cdx_kabsch_noise(prm.noise_seed, it, pc, gb, l + 4 * j) : 0.0;
#pragma unroll
    for (int e = 0; e < 9; e++) noise[e] = __shfl_sync(PG_FULL, mine[e >> 2], e & 3, 4);
  }
  pg_cp_wait();
  __syncwarp(gmask);                    // the group's staged GPIS outputs are visible to its sixteen lanes
  // ---- copy p, finger l
  double red[9];                        // total | margin_f | gk_f | gtips_f 3 | gT_f 3
#pragma unroll
  for (int i = 0; i < 9; i++) red[i] = 0.0;
  {
    const int u = S.uniq[pc];
    double a[4][3], lossp, mg, ga[3], gTr[3], gkr;
#pragma unroll
    for (int f = 0; f < 4; f++)
      for (int r = 0; r < 3; r++) a[f][r] = V ? tips[f][r] : Tp[f][r] + S.coeff[pc][f] * (tips[f][r] - Tp[f][r]);
    const double* gq = sq[cl];
    CDX_PG_TICK(8);
    pg_xchg xc = {V == 0 ? &sx[(cl * 4 + p) * PG_XQ] : nullptr, 0};
    cdx_pg_row_coop<V == 0>(l, xc, a, Tp, par + 16, gq + u * 4, gq + 16 + u * 4, reinterpret_cast<const double (*)[3]>(gq + 32 + u * 12),
                    reinterpret_cast<const double (*)[3]>(gq + 80 + u * 12), m9, reinterpret_cast<const double (*)[3]>(g9), noise,
                    S.C, lossp, mg, ga, gTr, gkr);
    CDX_PG_TICK(6);
    const double wp = S.weight[pc];
    const double c = S.coeff[pc][l];
    red[0] = wp * (lossp + S.w_ref * refn);
    red[1] = wp * mg;
    red[2] = wp * gkr;
    for (int r = 0; r < 3; r++) {
      red[3 + r] = wp * c * ga[r];
      red[6 + r] = wp * ((1.0 - c) * ga[r] + gTr[r]);
    }
    if (p >= S.P)
      for (int i = 0; i < 9; i++) red[i] = 0.0;
  }
#pragma unroll
  for (int i = 0; i < 9; i++) {         // over the copies: (p0 + p1) + (p2 + p3)
    double v = red[i];
    v += __shfl_xor_sync(PG_FULL, v, 4);
    v += __shfl_xor_sync(PG_FULL, v, 8);
    red[i] = v;
  }
  if (p == 0) {
    if (l == 0) sr[cl][0] = red[0];
    sr[cl][1 + l] = red[1];
    sr[cl][5 + l] = red[2];
    for (int r = 0; r < 3; r++) { sr[cl][9 + l * 3 + r] = red[3 + r]; sr[cl][21 + l * 3 + r] = red[6 + r]; }
  }
  __syncwarp(gmask);
  CDX_PG_TICK(9);
  // ---- :757-763 and the chain rule of finger f = l (lanes of p == 0)
  double total = red[0];
  for (int f = 0; f < 4; f++) total -= S.w_pre * m9[4 + f];
  double gpalm_pm[3] = {0.0, 0.0, 0.0};
  if (S.optimize_palm) {
    const double pm = m9[8];
    total += 1.0 / pm;
    for (int r = 0; r < 3; r++) gpalm_pm[r] = -g9[24 + r] / (pm * pm);
  }
  const int f = l;
  double* g = sg[cl];
  // the reduced values are identical on the four copy quads, so every lane forms the palm sums (the last shuffles of the
  // kernel: all 32 lanes take part); only the lanes of p == 0 go on
  double gt[3], gl[3], acc[12];
  for (int r = 0; r < 3; r++) gt[r] = red[3 + r] - S.w_pre * g9[(4 + f) * 3 + r];
  cdx_mtv(Rw, gt, gl);
  for (int r = 0; r < 3; r++) {
    acc[r] = gt[r];
    for (int c = 0; c < 3; c++) acc[3 + r * 3 + c] = gt[r] * tlp[f][c];
  }
#pragma unroll
  for (int i = 0; i < 12; i++) {
    acc[i] += __shfl_xor_sync(PG_FULL, acc[i], 1);
    acc[i] += __shfl_xor_sync(PG_FULL, acc[i], 2);
  }
  if (!live) return;                    // a group past the end has served the warp's shuffles and writes nothing
  if (p == 0) {
    // back through .double() (:666) into the float32 FK graph; fingers take turns so shared joints would accumulate
    const float g32[3] = {(float)gl[0], (float)gl[1], (float)gl[2]};
    if (V == 2) {
      for (int r = 0; r < 3; r++) g[f * 3 + r] = gt[r];       // tip_pose is the leaf
      __syncwarp(fmask);
    } else {
      const cdx_chain& ch = hs->finger[f];
      const float* Jf = sJ[cl][f];
      for (int turn = 0; turn < 4; turn++) {
        if (turn == f)
          for (int c = 0; c < ch.ndof; c++) {
            const float gj = Jf[c] * g32[0] + Jf[CDX_MAX_CHAIN_DOF + c] * g32[1] + Jf[2 * CDX_MAX_CHAIN_DOF + c] * g32[2];
            g[ch.dofs[c]] += (double)gj;
          }
        __syncwarp(fmask);
      }
    }
    g[16 + f] = red[2];
    for (int r = 0; r < 3; r++) g[20 + f * 3 + r] = red[6 + r];
    for (int i = l; i < 6; i += 4) {
      double v;
      if (i < 3) v = acc[i] + gpalm_pm[i];
      else {
        const double* dR = i == 3 ? dR0 : (i == 4 ? dR1 : dR2);
        v = 0.0;
        for (int e = 0; e < 9; e++) v += acc[3 + e] * dR[e];
      }
      g[32 + i] = v;
    }
  }
  __syncwarp(gmask);
  if (S.w_ref != 0.0 && sub < nd) g[sub] += refn > 0.0 ? wsum * S.w_ref * (par[sub] - S.ref_q[sub]) / refn : 0.0;
  __syncwarp(gmask);
  if (sub == 0 && loss) loss[b] = total;
  // the reference prints "NaN detected" and carries on (optimize_pregrasp.py:819): counted on the device, read after the run
  if (sub == 0 && do_step && total != total) {
    atomicAdd(reinterpret_cast<unsigned long long*>(nan_rec), 1ull);
    atomicMin(nan_rec + 1, (long long)it);
  }
  if (sub < 4 && margin) margin[b * 4 + sub] = sr[cl][1 + sub];
  if (grad) for (int i = sub; i < NPAR; i += 16) grad[b * NPAR + i] = g[i];
  CDX_PG_TICK(10);
  if (!do_step) return;
  // ---- best-iterate tracking (:822-829): BEFORE the optimiser step, only for s > track_after; NaN never updates
  const bool better = it > prm.track_after && total < oval;
  if (V != 0 && p == 0) {
    mring[(it & 1) * B * 4 + b * 4 + l] = sr[cl][1 + l];
    if (better && l == 0) latch[it & 1] = it;          // benign race: every writer stores the same value
  }
  if (better) {
    if (sub == 0) ovalue[b] = total;
    if (sub < 4) omargin[b * 4 + sub] = sr[cl][1 + sub];
    for (int i = sub; i < NPAR; i += 16) {
      const double v = par[i];
      if (i < 16) { if (i < nd) oq[b * nd + i] = v; }
      else if (i < 20) ok[b * 4 + i - 16] = v;
      else if (i < 32) oT[b * 12 + i - 20] = v;
      else opalm[b * 6 + i - 32] = v;
    }
  }
  // ---- torch.optim.Adam (:796,:831): exp_avg.lerp_(g, 1-b1); exp_avg_sq = b2 v + (1-b2) g^2;
  //      p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps);  then the clamps (:833-834)
  const double omb1 = 1.0 - prm.beta1, omb2 = 1.0 - prm.beta2;
#pragma unroll
  for (int j = 0; j < (NPAR + 15) / 16; j++) {
    const int i = sub + 16 * j;
    if (i >= NPAR) break;
    int grp;
    bool upd = true;
    if (i < 16) { upd = i < nd; grp = 0; }
    else if (i < 20) grp = 1;
    else if (i < 32) { upd = prm.optimize_target != 0; grp = 2; }
    else { upd = prm.optimize_palm != 0; grp = i < 35 ? 3 : 4; }
    double pv = par[i];
    if (upd && V == 0) {
      const double gi = g[i];
      double m = sm1[cl][i], v = sm2[cl][i];
      m = m + omb1 * (gi - m);
      v = v * prm.beta2 + omb2 * gi * gi;
      m1[b * NPAR + i] = m;
      m2[b * NPAR + i] = v;
      const double step = prm.lr[grp] / bc1;
      const double denom = sqrt(v) / bc2s + prm.eps;
      pv = pv - step * (m / denom);
    } else if (upd) {
      // torch.optim.RMSprop (:451-456, :347-352; alpha 0.99, eps 1e-8, no momentum, not centered):
      //   square_avg = alpha square_avg + (1 - alpha) g^2;  p -= lr * g / (sqrt(square_avg) + eps)
      const double gi = g[i];
      double v = sm2[cl][i];
      v = v * prm.rms_alpha + (1.0 - prm.rms_alpha) * gi * gi;
      m2[b * NPAR + i] = v;
      pv = pv + (-prm.lr[grp] * gi) / (sqrt(v) + prm.eps);   // addcdiv_(g, avg, value=-lr): (value * g) / avg
    }
    if (V == 2 && i < 12)   // tip_pose.clamp_(box) (:402)
      pv = clamp_hi(clamp_lo(pv, (double)prm.box_lo[i / 3][i % 3]), (double)prm.box_hi[i / 3][i % 3]);
    if (i >= 16 && i < 20) pv = clamp_lo(pv, prm.compliance_min);
    if (i >= 20 && i < 32 && V != 2) {   // target clamp: prob (:834) and Kin (:505); commented out at :404
      const int e = i - 20;
      pv = clamp_hi(clamp_lo(pv, (double)prm.box_lo[e / 3][e % 3]), (double)prm.box_hi[e / 3][e % 3]);
    }
    par[i] = pv;
    if (i < 16) { if (i < nd) q[b * nd + i] = pv; }
    else if (i < 20) k[b * 4 + i - 16] = pv;
    else if (i < 32) T[b * 12 + i - 20] = pv;
    else palm[b * 6 + i - 32] = pv;
  }
  if (V != 2) sq32[cl][sub] = sub < nd ? (float)par[sub] : 0.f;       // q.float() (:665) of the new iterate, for the FK below
  __syncwarp(gmask);
  CDX_PG_TICK(11);
  // ---- FK + GPIS query points of the NEW iterate (what pg_points_kernel does for the first iteration)
  if (p == 0) pg_emit_points<V>(hs, S, U, b, f, par, par + 20 + f * 3, par + 32, tl, J, Xext, Xmean, nullptr, V != 2 ? sq32[cl] : nullptr);
  CDX_PG_TICK(12);
}

// ProbabilisticGraspOptimizer.compute_collision_loss (optimize_pregrasp.py:684-701) on anchor poses [B,A,3]: thread per
// candidate; masked-out terms are multiplied by 0.0 exactly as the reference does (so NaN / inf propagate the same way)
__global__ void __launch_bounds__(128) pg_collision_kernel(const double* __restrict__ anchor, int64_t B, int A,
                                                           const int* __restrict__ left, const int* __restrict__ right, int npairs,
                                                           double threshold, const double* __restrict__ palm,
                                                           double* __restrict__ cost) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double* a = anchor + b * A * 3;
  double c = 0.0;
  for (int p = 0; p < npairs; p++) {
    const double* l = a + left[p] * 3;
    const double* r = a + right[p] * 3;
    const double dx = l[0] - r[0], dy = l[1] - r[1], dz = l[2] - r[2];
    const double dist = sqrt(dx * dx + dy * dy + dz * dz);
    double inv = 1.0 / dist;
    if (!(dist < threshold)) inv *= 0.0;
    c += inv;
  }
  double zc = 0.0;
  for (int i = 0; i < A; i++) {
    const double z = a[i * 3 + 2];
    double v = 1.0 / z * 0.1;
    if (!(z < 0.02)) v *= 0.0;
    zc += v;
  }
  c += zc;
  if (palm) {
    const double z = palm[b * 6 + 2];
    double v = 1.0 / z;
    if (!(z < 0.02)) v *= 0.0;
    c += v;
  }
  cost[b] = c;
}

// arg-min with lowest-index tie-break over opt_value (NaN / inf never win unless everything is)
__global__ void __launch_bounds__(1024) pg_argmin_kernel(const double* __restrict__ v, int64_t B, int64_t offset,
                                                         double* __restrict__ out) {
  __shared__ double sv[1024];
  __shared__ long long si[1024];
  double best = INFINITY;
  long long bi = -1;
  for (int64_t i = threadIdx.x; i < B; i += blockDim.x) {
    const double x = v[i];
    if (!(x != x) && (bi < 0 || x < best)) { best = x; bi = i; }
  }
  sv[threadIdx.x] = best; si[threadIdx.x] = bi;
  __syncthreads();
  for (int o = blockDim.x / 2; o; o >>= 1) {
    if ((int)threadIdx.x < o) {
      const double x = sv[threadIdx.x + o];
      const long long j = si[threadIdx.x + o];
      const bool take = j >= 0 && (si[threadIdx.x] < 0 || x < sv[threadIdx.x] || (x == sv[threadIdx.x] && j < si[threadIdx.x]));
      if (take) { sv[threadIdx.x] = x; si[threadIdx.x] = j; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out[0] = si[0] >= 0 ? sv[0] : NAN;
    out[1] = (double)(si[0] >= 0 ? si[0] + offset : -1);
  }
}

// opt_margin of the Kin / GPIS variants after the last iteration: the step kernel moves an improving iteration's margins into
// `latched` one iteration later, so the final iteration (index last) is still pending in its ring slot
__global__ void pg_latch_read_kernel(const double* __restrict__ mring, const double* __restrict__ latched,
                                     const int* __restrict__ latch, int64_t B, int last, double* __restrict__ out,
                                     int* __restrict__ iter_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool pending = last >= 0 && latch[last & 1] == last;
  if (i < B * 4 && out) out[i] = pending ? mring[(int64_t)(last & 1) * B * 4 + i] : latched[i];
  if (i == 0 && iter_out) iter_out[0] = pending ? last : latch[2];
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
static void pg_free_buffers(cdx_pg_state* s) {
  void* ptrs[] = {s->q, s->k, s->T, s->palm, s->m1, s->m2, s->oq, s->ok, s->oT, s->opalm, s->omargin, s->ovalue, s->tl,
                  s->J, s->Xext, s->Xmean, s->dist, s->sd, s->gmu, s->gsd, s->mean9, s->gmean9, s->loss, s->margin,
                  s->grad, s->mring, s->latched, s->latch};
  for (void* p : ptrs) cudaFree(p);
  s->mring = s->latched = nullptr;
  s->latch = nullptr;
  s->q = s->k = s->T = s->palm = s->m1 = s->m2 = s->oq = s->ok = s->oT = s->opalm = s->omargin = s->ovalue = nullptr;
  s->tl = s->Xext = s->Xmean = s->dist = s->sd = s->gmu = s->gsd = s->mean9 = s->gmean9 = s->loss = s->margin = s->grad = nullptr;
  s->J = nullptr;
  s->cap = 0;
}

void cdx_pregrasp_free(cdx_ctx* ctx) {
  if (!ctx || !ctx->pg) return;
  pg_free_buffers(ctx->pg);
  if (ctx->pg->side) { cudaStreamDestroy(ctx->pg->side); cudaEventDestroy(ctx->pg->ev_fork); cudaEventDestroy(ctx->pg->ev_join); }
  for (int g = 1; g < cdx_pg_state::MAX_GROUPS; g++)
    if (ctx->pg->gmain[g]) {
      cudaStreamDestroy(ctx->pg->gmain[g]); cudaStreamDestroy(ctx->pg->gside[g]);
      cudaEventDestroy(ctx->pg->gfork[g]); cudaEventDestroy(ctx->pg->gjoin[g]); cudaEventDestroy(ctx->pg->gend[g]);
    }
  if (ctx->pg->ev_begin) cudaEventDestroy(ctx->pg->ev_begin);
  for (cudaEvent_t e : ctx->pg->prof_ev) cudaEventDestroy(e);
  if (ctx->pg->gexec) cudaGraphExecDestroy(ctx->pg->gexec);
  if (ctx->pg->cap_stream) cudaStreamDestroy(ctx->pg->cap_stream);
  cudaFree(ctx->pg->it_dev);
  cudaFree(ctx->pg->nan_dev);
  cudaFree(ctx->pg->adam_tab);
  cudaFree(ctx->pg->hand_dev);
  delete ctx->pg;
  ctx->pg = nullptr;
}

static uint64_t pg_hash(uint64_t h, const void* p, size_t n) {   // FNV-1a
  const unsigned char* c = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) { h ^= c[i]; h *= 1099511628211ull; }
  return h;
}

static cdx_pg_state* pg_get(cdx_ctx* ctx) {
  if (!ctx->pg) ctx->pg = new cdx_pg_state();
  return ctx->pg;
}

extern "C" int cdx_hand_load(cdx_ctx* ctx, int n_bodies, const int* parent, const int* dof, const int* axis,
                             const float* axis_sign, const float* fixed_rot, const float* xyz, int n_dofs,
                             const int* tip_body, const float* tip_offset) {
  CDX_NVTX("cdx_hand_load");
  if (!ctx || !parent || !dof || !axis || !axis_sign || !fixed_rot || !xyz || !tip_body || !tip_offset || n_bodies < 1 ||
      n_dofs < 1 || n_dofs > 16) {
    cdx_set_error("cdx_hand_load: bad argument (n_dofs must be 1..16)");
    return CDX_ERR_ARG;
  }
  cdx_hand_dev h;
  memset(&h, 0, sizeof(h));
  int rc = cdx_build_chains(n_bodies, parent, dof, axis, axis_sign, fixed_rot, xyz, n_dofs, tip_body, tip_offset, &h);
  if (rc) {
    cdx_set_error("cdx_hand_load: invalid kinematic tree (code %d: -1 tip/parent, -2 chain > %d bodies, -3 > %d dofs per finger)",
                  rc, CDX_MAX_CHAIN, CDX_MAX_CHAIN_DOF);
    return CDX_ERR_ARG;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_pg_state* s = pg_get(ctx);
  // the loaded hand's identity lives on the context: callers (one optimizer object per hand, several sharing one GPIS
  // context) load unconditionally and an identical hand costs nothing
  const uint64_t id = pg_hash(1469598103934665603ull, &h, sizeof(h));
  if (s->hand_dev && ctx->hand_ready && s->hand_id == id && s->hand_n_dofs == n_dofs) return CDX_OK;
  if (!s->hand_dev) CDX_CUDA(cudaMalloc((void**)&s->hand_dev, sizeof(cdx_hand_dev)));
  CDX_CUDA(cudaDeviceSynchronize());   // non-blocking side / group streams may still be reading the previous hand
  CDX_CUDA(cudaMemcpy(s->hand_dev, &h, sizeof(h), cudaMemcpyHostToDevice));
  s->hand_id = id;
  s->setup.n_dofs = s->hand_n_dofs = n_dofs;
  ctx->hand_ready = true;
  return CDX_OK;
}

extern "C" int cdx_hand_fk(cdx_ctx* ctx, const double* q, const double* palm, int64_t B, double* tips, void* stream) {
  CDX_NVTX("cdx_hand_fk");
  if (!ctx || !ctx->hand_ready) { cdx_set_error("cdx_hand_fk: no hand loaded"); return CDX_ERR_STATE; }
  if (B < 0 || (B > 0 && (!q || !palm || !tips))) { cdx_set_error("cdx_hand_fk: bad argument"); return CDX_ERR_ARG; }
  if (B == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_pg_state* s = ctx->pg;
  cdx_pg_setup S;
  memset(&S, 0, sizeof(S));
  S.n_dofs = s->hand_n_dofs;
  pg_points_kernel<0><<<(unsigned)((B * 4 + 127) / 128), 128, 0, (cudaStream_t)stream>>>(s->hand_dev, B, 1, S, q, nullptr, palm, nullptr,
                                                                                 nullptr, nullptr, nullptr, tips);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

// KinGPISGraspOptimizer.forward_kinematics (optimize_pregrasp.py:429-436): recursive FK on float64 angles; palm [B,6] =
// the constant palm offset (xyz, zero Euler angles) repeated per candidate
extern "C" int cdx_hand_fk_recursive(cdx_ctx* ctx, const double* q, const double* palm_offset3, int64_t B, double* tips,
                                     void* stream) {
  CDX_NVTX("cdx_hand_fk_recursive");
  if (!ctx || !ctx->hand_ready) { cdx_set_error("cdx_hand_fk_recursive: no hand loaded"); return CDX_ERR_STATE; }
  if (B < 0 || (B > 0 && (!q || !palm_offset3 || !tips))) { cdx_set_error("cdx_hand_fk_recursive: bad argument"); return CDX_ERR_ARG; }
  if (B == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_pg_state* s = ctx->pg;
  cdx_pg_setup S;
  memset(&S, 0, sizeof(S));
  S.n_dofs = s->hand_n_dofs;
  S.variant = 1;
  pg_points_kernel<1><<<(unsigned)((B * 4 + 127) / 128), 128, 0, (cudaStream_t)stream>>>(s->hand_dev, B, 1, S, q, nullptr, palm_offset3,
                                                                                 nullptr, nullptr, nullptr, nullptr, tips);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

extern "C" int cdx_collision_cost(cdx_ctx* ctx, const double* anchor, int64_t B, int A, const int* left, const int* right,
                                  int npairs, double threshold, const double* palm, double* cost, void* stream) {
  CDX_NVTX("cdx_collision_cost");
  if (!ctx || B < 0 || A < 1 || npairs < 0 || (B > 0 && (!anchor || !cost)) || (npairs > 0 && (!left || !right))) {
    cdx_set_error("cdx_collision_cost: bad argument");
    return CDX_ERR_ARG;
  }
  if (B == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  pg_collision_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(anchor, B, A, left, right, npairs, threshold,
                                                                                 palm, cost);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

#define PG_ALLOC(ptr, n) CDX_CUDA(cudaMalloc((void**)&(ptr), sizeof(*(ptr)) * (size_t)(n)))

extern "C" int cdx_pregrasp_init(cdx_ctx* ctx, const cdx_pregrasp_params* p, int64_t B, const double* q0,
                                 const double* k0, const double* T0, const double* palm0, void* stream) {
  CDX_NVTX("cdx_pregrasp_init");
  if (!ctx || !p) { cdx_set_error("cdx_pregrasp_init: NULL argument"); return CDX_ERR_ARG; }
  if (p->variant < 0 || p->variant > 2) { cdx_set_error("cdx_pregrasp_init: variant must be 0, 1 or 2"); return CDX_ERR_ARG; }
  if (p->variant != 2 && !ctx->hand_ready) { cdx_set_error("cdx_pregrasp_init: cdx_hand_load first"); return CDX_ERR_STATE; }
  if (B < 1 || !q0 || !k0 || !T0 || !palm0 || p->num_copies < 1 || p->num_copies > CDX_MAX_COPIES ||
      (p->variant != 0 && p->num_copies != 1)) {
    cdx_set_error("cdx_pregrasp_init: bad argument (1 <= num_copies <= %d, exactly 1 for the Kin / GPIS variants, B >= 1)",
                  CDX_MAX_COPIES);
    return CDX_ERR_ARG;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = pg_get(ctx);
  s->prm = *p;
  cdx_pg_setup& S = s->setup;
  if (p->variant == 2) {   // GPISGraspOptimizer: the 12 tip coordinates take the place of the joint angles; no hand
    S.n_dofs = 12;
    if (!s->hand_dev) {
      CDX_CUDA(cudaMalloc((void**)&s->hand_dev, sizeof(cdx_hand_dev)));
      CDX_CUDA(cudaMemset(s->hand_dev, 0, sizeof(cdx_hand_dev)));
    }
  }
  else S.n_dofs = s->hand_n_dofs;
  const int nd = S.n_dofs;
  S.variant = p->variant;
  S.C = cdx_pg_const();
  if (p->variant == 0) { S.w_ref = 10.0; S.w_pre = 5.0; }
  else {                  // :468-479 (Kin) / :377-387 (GPIS)
    S.w_ref = p->variant == 1 ? 20.0 : 0.0;
    S.w_pre = 0.0;
    S.C.w_reward = p->variant == 1 ? 5.0 : 25.0;
    S.C.w_cm = 0.0;
    S.C.w_tar = 10.0;
    S.C.force_clamp = 1.0;
    S.C.w_center = 10.0;
    S.C.var_sum = p->variant == 2 ? 1 : 0;
  }
  S.P = p->num_copies;
  S.U = 0;
  for (int a = 0; a < S.P; a++) {       // identical coefficient rows share one block of GPIS queries
    int found = -1;
    for (int b = 0; b < a && found < 0; b++)
      if (memcmp(p->coeffs[a], p->coeffs[b], sizeof(float) * 4) == 0) found = S.uniq[b];
    S.uniq[a] = found >= 0 ? found : S.U++;
    for (int f = 0; f < 4; f++) S.coeff[a][f] = (double)p->coeffs[a][f];
    S.weight[a] = (double)p->weights[a];
  }
  for (int i = 0; i < 16; i++) S.ref_q[i] = i < nd ? (double)p->ref_q[i] : 0.0;
  S.optimize_palm = p->optimize_palm;
  S.C.cos_mu = (double)p->cos_mu;
  for (int r = 0; r < 3; r++) S.C.com[r] = (double)p->com[r];
  S.C.dummy_k = (double)p->dummy_k;
  S.C.dummy_M = (double)p->dummy_M;
  S.C.gravity = p->gravity;
  S.C.uncertainty = p->uncertainty;
  const int U = S.U;
  if (B > s->cap || U != s->U) {
    pg_free_buffers(s);
    PG_ALLOC(s->q, B * 16); PG_ALLOC(s->k, B * 4); PG_ALLOC(s->T, B * 12); PG_ALLOC(s->palm, B * 6);
    PG_ALLOC(s->m1, B * NPAR); PG_ALLOC(s->m2, B * NPAR);
    PG_ALLOC(s->oq, B * 16); PG_ALLOC(s->ok, B * 4); PG_ALLOC(s->oT, B * 12); PG_ALLOC(s->opalm, B * 6);
    PG_ALLOC(s->omargin, B * 4); PG_ALLOC(s->ovalue, B);
    PG_ALLOC(s->tl, B * 12); PG_ALLOC(s->J, B * 12 * CDX_MAX_CHAIN_DOF);
    PG_ALLOC(s->Xext, B * U * 12); PG_ALLOC(s->Xmean, B * 27);
    PG_ALLOC(s->dist, B * U * 4); PG_ALLOC(s->sd, B * U * 4); PG_ALLOC(s->gmu, B * U * 12); PG_ALLOC(s->gsd, B * U * 12);
    PG_ALLOC(s->mean9, B * 9); PG_ALLOC(s->gmean9, B * 27);
    PG_ALLOC(s->loss, B); PG_ALLOC(s->margin, B * 4); PG_ALLOC(s->grad, B * NPAR);
    PG_ALLOC(s->mring, B * 8); PG_ALLOC(s->latched, B * 4); PG_ALLOC(s->latch, 4);
    s->cap = B;
    s->U = U;
  }
  if (!s->side) {
    // highest priority: the short mean-only query's CTAs are placed ahead of the extended query's whenever SMs free up
    int prio_lo = 0, prio_hi = 0;
    CDX_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CDX_CUDA(cudaStreamCreateWithPriority(&s->side, cudaStreamNonBlocking, prio_hi));
    CDX_CUDA(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    CDX_CUDA(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
  }
  if (!s->it_dev) {
    CDX_CUDA(cudaMalloc((void**)&s->it_dev, sizeof(int)));
    CDX_CUDA(cudaMalloc((void**)&s->nan_dev, sizeof(long long) * 2));
  }
  {
    const long long nan0[2] = {0, 0x7fffffffll};
    CDX_CUDA(cudaMemcpyAsync(s->nan_dev, nan0, sizeof(nan0), cudaMemcpyHostToDevice, st));
    CDX_CUDA(cudaMemsetAsync(s->it_dev, 0, sizeof(int), st));
  }
  s->B = B;
  s->generation++;
  CDX_CUDA(cudaMemcpyAsync(s->q, q0, sizeof(double) * B * nd, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->k, k0, sizeof(double) * B * 4, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->T, T0, sizeof(double) * B * 12, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->palm, palm0, sizeof(double) * B * 6, cudaMemcpyDeviceToDevice, st));
  // :797-804: best iterate starts at the initial iterate, opt_value = inf, opt_margin = 0
  CDX_CUDA(cudaMemcpyAsync(s->oq, q0, sizeof(double) * B * nd, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->ok, k0, sizeof(double) * B * 4, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->oT, T0, sizeof(double) * B * 12, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemcpyAsync(s->opalm, palm0, sizeof(double) * B * 6, cudaMemcpyDeviceToDevice, st));
  CDX_CUDA(cudaMemsetAsync(s->omargin, 0, sizeof(double) * B * 4, st));
  CDX_CUDA(cudaMemsetAsync(s->m1, 0, sizeof(double) * B * NPAR, st));
  CDX_CUDA(cudaMemsetAsync(s->m2, 0, sizeof(double) * B * NPAR, st));
  CDX_CUDA(cudaMemsetAsync(s->latched, 0, sizeof(double) * B * 4, st));
  CDX_CUDA(cudaMemsetAsync(s->latch, 0xFF, sizeof(int) * 4, st));     // -1: nothing improved yet
  pg_fill_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(s->ovalue, B, INFINITY);
  CDX_CHECK_LAUNCH(ctx);
  s->it_host = 0;
  s->points_valid = false;
  return CDX_OK;
}

static int pg_points(cdx_ctx* ctx, cdx_pg_state* s, cudaStream_t st) {
  if (s->points_valid) return CDX_OK;
  const int64_t B = s->B;
  const int U = s->U;
#define PG_POINTS(V) pg_points_kernel<V><<<(unsigned)((B * 4 + 127) / 128), 128, 0, st>>>(s->hand_dev, B, U, s->setup, s->q, s->T, \
                                                                                        s->palm, s->tl, s->J, s->Xext, s->Xmean, nullptr)
  if (s->setup.variant == 0) PG_POINTS(0); else if (s->setup.variant == 1) PG_POINTS(1); else PG_POINTS(2);
#undef PG_POINTS
  CDX_CHECK_LAUNCH(ctx);
  s->points_valid = true;
  return CDX_OK;
}

// one closure evaluation of candidates [b0, b1): (points of the initial iterate) -> GPIS queries (extended on `st`, mean-only
// overlapped on the side stream) -> step kernel (objective + gradient, optionally Adam + points of the new iterate).
// The iteration index the step kernel sees is *it_base + it_off (it_base == nullptr: it_off alone).
static int pg_iteration(cdx_ctx* ctx, cdx_pg_state* s, int do_step, double* loss, double* margin, double* grad,
                        cudaStream_t st, const int* it_base, int it_off, int64_t b0 = 0, int64_t b1 = -1, int group = 0) {
  const int64_t B = s->B;
  if (b1 < 0) b1 = B;
  const int64_t nb = b1 - b0;
  const int U = s->U;
  const bool prof = s->prof && do_step;
  cudaStream_t side = group ? s->gside[group] : s->side;
  cudaEvent_t ev_fork = group ? s->gfork[group] : s->ev_fork, ev_join = group ? s->gjoin[group] : s->ev_join;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  auto mark = [&](int slot, cudaStream_t on) {
    if (!prof) return;
    cudaEventCreate(&ev[slot]);
    cudaEventRecord(ev[slot], on);
  };
  int rc = pg_points(ctx, s, st);
  if (rc) return rc;
  CDX_CUDA(cudaEventRecord(ev_fork, st));
  CDX_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
  // the short mean-only query is launched FIRST (forked stream): its CTAs take their SMs for ~10 us and are gone long before
  // the extended query ends, instead of queueing behind the extended query's CTAs and finishing after it
  mark(2, side);
  rc = cdx_gpis_query(ctx, s->Xmean + b0 * 27, nb * 9, s->mean9 + b0 * 9, s->gmean9 + b0 * 27, nullptr, nullptr, (void*)side);
  if (rc) return rc;
  mark(3, side);
  CDX_CUDA(cudaEventRecord(ev_join, side));
  mark(0, st);
  rc = cdx_gpis_query(ctx, s->Xext + b0 * U * 12, nb * U * 4, s->dist + b0 * U * 4, s->gmu + b0 * U * 12, s->sd + b0 * U * 4,
                      s->gsd + b0 * U * 12, (void*)st);
  if (rc) return rc;
  mark(1, st);
  CDX_CUDA(cudaStreamWaitEvent(st, ev_join, 0));
  mark(4, st);
  constexpr size_t PG_SX_BYTES = sizeof(double) * CPB * 4 * PG_XQ;
  if (s->setup.variant == 0) {
    if (int rc2 = cdx_smem_attr(ctx, pg_step_kernel<0, 2>, PG_SX_BYTES)) return rc2;
    if (int rc2 = cdx_smem_attr(ctx, pg_step_kernel<0, 3>, PG_SX_BYTES)) return rc2;
  }
  int occ = (nb + CPB - 1) / CPB > 2 * (int64_t)ctx->sms ? 3 : 2;
  if (s->tune_occ == 2 || s->tune_occ == 3) occ = s->tune_occ;
#define PG_STEP(V) PG_STEP_O(V, 2); else PG_STEP_O(V, 3)
#define PG_STEP_O(V, O) if (occ == O) pg_step_kernel<V, O><<<(unsigned)((nb + CPB - 1) / CPB), CPB * 16, V == 0 ? PG_SX_BYTES : 0, st>>>(                                  \
      s->hand_dev, B, b0, b1, U, s->setup, s->prm, s->q, s->k, s->T, s->palm, s->tl, s->J, s->dist, s->sd, s->gmu, s->gsd,      \
      s->mean9, s->gmean9, loss, margin, grad, do_step, it_base, it_off, s->nan_dev, s->adam_tab, s->m1, s->m2, s->oq, s->ok,   \
      s->oT, s->opalm, s->omargin, s->ovalue, s->Xext, s->Xmean, s->mring, s->latched, s->latch)
  if (s->setup.variant == 0) { PG_STEP(0); } else if (s->setup.variant == 1) { PG_STEP(1); } else { PG_STEP(2); }
#undef PG_STEP
#undef PG_STEP_O
  CDX_CHECK_LAUNCH(ctx);
  mark(5, st);
  if (prof)
    for (int i = 0; i < 6; i++) s->prof_ev.push_back(ev[i]);
  return CDX_OK;
}

// How many candidate groups cdx_pregrasp_run pipelines.  The iteration is a chain of latency-bound kernels (queries ->
// step -> queries ...) that leaves most of the chip idle at C2's 1,024 candidates, and the candidates are independent
// (ProbabilisticGraspOptimizer has no cross-candidate term; the noise stream is keyed by candidate), so two or more groups
// walking the same iterations on their own streams fill each other's gaps: results are bit-identical to one group.
// Measured at 1,024 candidates: 76.3 us per iteration in one group, 72.2 in two, 73 in three or four (a group's step kernel
// slows down beside another group's FP64 tensor work, so the gain stays well under what the gaps suggest).
// The Kin / GPIS variants latch the WHOLE batch's margin when any candidate improves (:491-494): one group.
static int pg_groups(const cdx_pg_state* s) {
  if (s->setup.variant != 0) return 1;
  int G = s->B >= 768 ? 2 : 1;
  if (s->tune_groups > 0) G = s->tune_groups;
  if (G > cdx_pg_state::MAX_GROUPS) G = cdx_pg_state::MAX_GROUPS;
  while (G > 1 && s->B / G < CPB) G--;
  return G;
}

static int pg_ready(cdx_ctx* ctx, const char* who) {
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("%s: cdx_pregrasp_init first", who); return CDX_ERR_STATE; }
  if (!ctx->g.ready) { cdx_set_error("%s: no fitted GPIS state in this context", who); return CDX_ERR_STATE; }
  return CDX_OK;
}

static int pg_adam_table(cdx_pg_state* s, int upto, cudaStream_t st) {
  if (upto <= s->tab_cap) return CDX_OK;
  int cap = upto < 1024 ? 1024 : upto * 2;
  std::vector<double> tab((size_t)cap * 2);
  for (int i = 0; i < cap; i++) {   // python: 1 - beta ** step ; bias_correction2 ** 0.5
    tab[2 * i] = 1.0 - pow(s->prm.beta1, (double)(i + 1));
    tab[2 * i + 1] = sqrt(1.0 - pow(s->prm.beta2, (double)(i + 1)));
  }
  CDX_CUDA(cudaDeviceSynchronize());   // group / side streams may still read the old table
  cudaFree(s->adam_tab);
  s->adam_tab = nullptr;
  s->tab_cap = 0;
  s->g_key = 0;                        // a captured graph holds the old pointer
  CDX_CUDA(cudaMalloc((void**)&s->adam_tab, sizeof(double) * tab.size()));
  CDX_CUDA(cudaMemcpy(s->adam_tab, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice));
  s->tab_cap = cap;
  return CDX_OK;
}

__global__ void pg_set_int_kernel(int* p, int v) { *p = v; }

// iterations [0, n) of all G candidate groups on their stream pairs, forked from and joined back into `st`; the step
// kernels see iteration *it_base + it0 + i
static int pg_enqueue(cdx_ctx* ctx, cdx_pg_state* s, int G, int n, cudaStream_t st, const int* it_base, int it0) {
  if (G == 1) {
    for (int i = 0; i < n; i++) {
      int rc = pg_iteration(ctx, s, 1, s->loss, s->margin, nullptr, st, it_base, it0 + i);
      if (rc) return rc;
    }
    return CDX_OK;
  }
  // group boundaries on whole CTAs
  const int64_t per = ((s->B + G - 1) / G + CPB - 1) / CPB * CPB;
  CDX_CUDA(cudaEventRecord(s->ev_begin, st));
  for (int g = 1; g < G; g++) CDX_CUDA(cudaStreamWaitEvent(s->gmain[g], s->ev_begin, 0));
  for (int i = 0; i < n; i++)
    for (int g = 0; g < G; g++) {
      const int64_t b0 = g * per, b1 = b0 + per < s->B ? b0 + per : s->B;
      if (b0 >= b1) continue;
      int rc = pg_iteration(ctx, s, 1, s->loss, s->margin, nullptr, g ? s->gmain[g] : st, it_base, it0 + i, b0, b1, g);
      if (rc) return rc;
    }
  for (int g = 1; g < G; g++) {
    CDX_CUDA(cudaEventRecord(s->gend[g], s->gmain[g]));
    CDX_CUDA(cudaStreamWaitEvent(st, s->gend[g], 0));
  }
  return CDX_OK;
}

// everything a captured iteration bakes into its kernel nodes
static uint64_t pg_graph_key(const cdx_ctx* ctx, const cdx_pg_state* s, int G, int K) {
  uint64_t h = 1469598103934665603ull;
  const void* ptrs[] = {s->q, s->m1, s->ovalue, s->Xext, s->adam_tab, s->hand_dev, ctx->g.Xa, ctx->g.P, ctx->g.U, s->gmain[1]};
  h = pg_hash(h, ptrs, sizeof(ptrs));
  h = pg_hash(h, &s->B, sizeof(s->B));
  h = pg_hash(h, &s->prm, sizeof(s->prm));
  h = pg_hash(h, &s->setup, sizeof(s->setup));
  const int64_t misc[] = {G, K, s->U, s->tune_occ, ctx->g.N, ctx->g.Np, ctx->g.kernel, ctx->g.var_mode};
  h = pg_hash(h, misc, sizeof(misc));
  const double dm[] = {ctx->g.R, ctx->g.bias, ctx->g.sigma, ctx->g.prior};
  h = pg_hash(h, dm, sizeof(dm));
  return h ? h : 1;
}

extern "C" int cdx_pregrasp_run(cdx_ctx* ctx, int n_iters, void* stream) {
  CDX_NVTX("cdx_pregrasp_run");
  int rc = pg_ready(ctx, "cdx_pregrasp_run");
  if (rc) return rc;
  if (n_iters < 0) { cdx_set_error("cdx_pregrasp_run: n_iters < 0"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  rc = pg_adam_table(s, s->it_host + n_iters, st);
  if (rc) return rc;
  if (n_iters == 0) return CDX_OK;
  const int G = pg_groups(s);
  if (G > 1) {
    int prio_lo = 0, prio_hi = 0;
    CDX_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    if (!s->ev_begin) CDX_CUDA(cudaEventCreateWithFlags(&s->ev_begin, cudaEventDisableTiming));
    for (int g = 1; g < G; g++)
      if (!s->gmain[g]) {
        CDX_CUDA(cudaStreamCreateWithFlags(&s->gmain[g], cudaStreamNonBlocking));
        CDX_CUDA(cudaStreamCreateWithPriority(&s->gside[g], cudaStreamNonBlocking, prio_hi));
        CDX_CUDA(cudaEventCreateWithFlags(&s->gfork[g], cudaEventDisableTiming));
        CDX_CUDA(cudaEventCreateWithFlags(&s->gjoin[g], cudaEventDisableTiming));
        CDX_CUDA(cudaEventCreateWithFlags(&s->gend[g], cudaEventDisableTiming));
      }
  }
  rc = pg_points(ctx, s, st);
  if (rc) return rc;
  // The loop as a CUDA graph: K iterations - every group's three launches, forks and joins - are captured
  // once and re-launched n_iters / K times; the nodes read the chunk's base iteration from device memory, so the same
  // executable graph serves every chunk and every later run with the same batch.  Without it the host issues ~14 driver
  // calls per iteration.  Profiling (per-kernel events) and the remainder iterations use direct launches.
  int done = 0;
  if (s->use_graph && !s->prof) {
    int K = n_iters < 50 ? n_iters : 50;
    if (K >= 4) {
      const uint64_t key = pg_graph_key(ctx, s, G, K);
      if (s->gexec == nullptr || s->g_key != key) {
        if (s->gexec) { cudaGraphExecDestroy(s->gexec); s->gexec = nullptr; }
        // one direct iteration first: every kernel's attributes (dynamic shared memory) are set outside the capture
        rc = pg_enqueue(ctx, s, G, 1, st, nullptr, s->it_host);
        if (rc) return rc;
        s->it_host++;
        done = 1;
        cudaGraph_t graph = nullptr;
        const int64_t launches0 = ctx->launches;
        // captured on a stream of our own (the caller's may be the legacy default stream, which cannot capture); the
        // executable graph is then launched into the caller's stream
        if (!s->cap_stream) CDX_CUDA(cudaStreamCreateWithFlags(&s->cap_stream, cudaStreamNonBlocking));
        CDX_CUDA(cudaStreamBeginCapture(s->cap_stream, cudaStreamCaptureModeRelaxed));
        rc = pg_enqueue(ctx, s, G, K, s->cap_stream, s->it_dev, 0);
        cudaError_t ce = cudaStreamEndCapture(s->cap_stream, &graph);
        ctx->launches = launches0;                                   // captured, not launched
        if (rc || ce != cudaSuccess || !graph) {
          if (graph) cudaGraphDestroy(graph);
          if (!rc) { cdx_set_error("cdx_pregrasp_run: graph capture failed: %s", cudaGetErrorString(ce)); rc = CDX_ERR_CUDA; }
          cudaGetLastError();
          return rc;
        }
        ce = cudaGraphInstantiate(&s->gexec, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) {
          s->gexec = nullptr;
          cdx_set_error("cdx_pregrasp_run: cudaGraphInstantiate: %s", cudaGetErrorString(ce));
          return CDX_ERR_CUDA;
        }
        s->g_key = key;
        s->g_K = K;
      }
      const int per_iter = 3 * G;   // kernel nodes per captured iteration (mean query, extended query, step) per group
      while (n_iters - done >= s->g_K) {
        pg_set_int_kernel<<<1, 1, 0, st>>>(s->it_dev, s->it_host);
        CDX_CHECK_LAUNCH(ctx);
        CDX_CUDA(cudaGraphLaunch(s->gexec, st));
        ctx->launches += (int64_t)per_iter * s->g_K;
        s->it_host += s->g_K;
        done += s->g_K;
      }
    }
  }
  if (done < n_iters) {
    rc = pg_enqueue(ctx, s, G, n_iters - done, st, nullptr, s->it_host);
    if (rc) return rc;
    s->it_host += n_iters - done;
  }
  return CDX_OK;
}

extern "C" int cdx_pregrasp_set_tuning(cdx_ctx* ctx, int groups, int occupancy, int use_graph) {
  if (!ctx || groups < 0 || groups > cdx_pg_state::MAX_GROUPS || !(occupancy == 0 || occupancy == 2 || occupancy == 3)) {
    cdx_set_error("cdx_pregrasp_set_tuning: groups in 0..%d (0 = auto), occupancy 0 (auto), 2 or 3", cdx_pg_state::MAX_GROUPS);
    return CDX_ERR_ARG;
  }
  cdx_pg_state* s = pg_get(ctx);
  s->tune_groups = groups;
  s->tune_occ = occupancy;
  s->use_graph = use_graph != 0;
  return CDX_OK;
}

extern "C" int cdx_pregrasp_nan_count(cdx_ctx* ctx, int64_t* count_host, int* first_iter_host, void* stream) {
  if (!ctx || !ctx->pg || ctx->pg->B == 0 || !count_host) { cdx_set_error("cdx_pregrasp_nan_count: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  long long h[2] = {0, 0};
  CDX_CUDA(cudaMemcpyAsync(h, ctx->pg->nan_dev, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CDX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  *count_host = (int64_t)h[0];
  if (first_iter_host) *first_iter_host = h[0] > 0 ? (int)h[1] : -1;
  return CDX_OK;
}

extern "C" int64_t cdx_pregrasp_generation(cdx_ctx* ctx, int64_t* B_out) {
  if (!ctx || !ctx->pg) return 0;
  if (B_out) *B_out = ctx->pg->B;
  return (int64_t)ctx->pg->generation;
}

// the caller sized its buffers for B candidates: they must be the batch of the last cdx_pregrasp_init on this context
static int pg_batch_check(const cdx_pg_state* s, int64_t B, const char* who) {
  if (B == s->B) return CDX_OK;
  cdx_set_error("%s: buffers for %lld candidates, but the context holds a batch of %lld (another optimizer re-initialised it?)",
                who, (long long)B, (long long)s->B);
  return CDX_ERR_ARG;
}

extern "C" int cdx_pregrasp_eval(cdx_ctx* ctx, int64_t Bcap, double* total_loss, double* margin, double* gq, double* gc,
                                 double* gt, double* gp, void* stream) {
  CDX_NVTX("cdx_pregrasp_eval");
  int rc = pg_ready(ctx, "cdx_pregrasp_eval");
  if (rc) return rc;
  if ((rc = pg_batch_check(ctx->pg, Bcap, "cdx_pregrasp_eval"))) return rc;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  rc = pg_adam_table(s, s->it_host + 1, st);
  if (rc) return rc;
  rc = pg_iteration(ctx, s, 0, s->loss, s->margin, s->grad, st, nullptr, s->it_host);
  if (rc) return rc;
  const int64_t B = s->B;
  const int nd = s->setup.n_dofs;
  if (total_loss) CDX_CUDA(cudaMemcpyAsync(total_loss, s->loss, sizeof(double) * B, cudaMemcpyDeviceToDevice, st));
  if (margin) CDX_CUDA(cudaMemcpyAsync(margin, s->margin, sizeof(double) * B * 4, cudaMemcpyDeviceToDevice, st));
  const size_t pitch = sizeof(double) * NPAR;
  if (gq) CDX_CUDA(cudaMemcpy2DAsync(gq, sizeof(double) * nd, s->grad, pitch, sizeof(double) * nd, B, cudaMemcpyDeviceToDevice, st));
  if (gc) CDX_CUDA(cudaMemcpy2DAsync(gc, sizeof(double) * 4, s->grad + 16, pitch, sizeof(double) * 4, B, cudaMemcpyDeviceToDevice, st));
  if (gt) CDX_CUDA(cudaMemcpy2DAsync(gt, sizeof(double) * 12, s->grad + 20, pitch, sizeof(double) * 12, B, cudaMemcpyDeviceToDevice, st));
  if (gp) CDX_CUDA(cudaMemcpy2DAsync(gp, sizeof(double) * 6, s->grad + 32, pitch, sizeof(double) * 6, B, cudaMemcpyDeviceToDevice, st));
  return CDX_OK;
}

#define PG_COPY(dst, src, n) if (dst) CDX_CUDA(cudaMemcpyAsync(dst, src, sizeof(double) * (size_t)(n), cudaMemcpyDeviceToDevice, st))

extern "C" int cdx_pregrasp_results(cdx_ctx* ctx, int64_t Bcap, double* oq, double* ok, double* oT, double* opalm,
                                    double* omargin, double* ovalue, void* stream) {
  CDX_NVTX("cdx_pregrasp_results");
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_results: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  if (int rcb = pg_batch_check(ctx->pg, Bcap, "cdx_pregrasp_results")) return rcb;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  const int64_t B = s->B;
  PG_COPY(oq, s->oq, B * s->setup.n_dofs); PG_COPY(ok, s->ok, B * 4); PG_COPY(oT, s->oT, B * 12);
  PG_COPY(opalm, s->opalm, B * 6); PG_COPY(omargin, s->omargin, B * 4); PG_COPY(ovalue, s->ovalue, B);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_state(cdx_ctx* ctx, int64_t Bcap, double* q, double* k, double* T, double* palm, void* stream) {
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_state: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  if (int rcb = pg_batch_check(ctx->pg, Bcap, "cdx_pregrasp_state")) return rcb;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  const int64_t B = s->B;
  PG_COPY(q, s->q, B * s->setup.n_dofs); PG_COPY(k, s->k, B * 4); PG_COPY(T, s->T, B * 12); PG_COPY(palm, s->palm, B * 6);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_latched_margin(cdx_ctx* ctx, int64_t Bcap, double* margin, int* iter, void* stream) {
  if (ctx && ctx->pg && ctx->pg->B)
    if (int rcb = pg_batch_check(ctx->pg, Bcap, "cdx_pregrasp_latched_margin")) return rcb;
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_latched_margin: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  cdx_pg_state* s = ctx->pg;
  if (s->setup.variant == 0) {
    cdx_set_error("cdx_pregrasp_latched_margin: only the Kin / GPIS variants latch a whole-batch margin");
    return CDX_ERR_STATE;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  const int64_t B = s->B;
  pg_latch_read_kernel<<<(unsigned)((B * 4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(s->mring, s->latched, s->latch, B,
                                                                                   s->it_host - 1, margin, iter);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_argmin(cdx_ctx* ctx, double* out, void* stream) {
  CDX_NVTX("cdx_pregrasp_argmin");
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_argmin: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  if (!out) { cdx_set_error("cdx_pregrasp_argmin: out is NULL"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_pg_state* s = ctx->pg;
  pg_argmin_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(s->ovalue, s->B, s->prm.env_offset, out);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_profile(cdx_ctx* ctx, int enable) {
  if (!ctx) { cdx_set_error("cdx_pregrasp_profile: NULL context"); return CDX_ERR_ARG; }
  pg_get(ctx)->prof = enable != 0;
  return CDX_OK;
}

extern "C" int cdx_pregrasp_profile_read(cdx_ctx* ctx, double* ms_out4, int* n_iters, void* stream) {
  if (!ctx || !ctx->pg || !ms_out4 || !n_iters) { cdx_set_error("cdx_pregrasp_profile_read: bad argument"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  CDX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  cdx_pg_state* s = ctx->pg;
  for (int k = 0; k < 4; k++) ms_out4[k] = 0.0;
  // per iteration: [0] before ext, [1] after ext (main) | [2] before mean, [3] after mean (side) | [4] before step, [5] after
  const size_t n = s->prof_ev.size() / 6;
  const int pairs[4][2] = {{1, 4}, {0, 1}, {2, 3}, {4, 5}};   // wait for the mean query after ext | ext | mean | step
  for (size_t i = 0; i < n; i++)
    for (int k = 0; k < 4; k++) {
      float ms = 0.f;
      CDX_CUDA(cudaEventElapsedTime(&ms, s->prof_ev[i * 6 + pairs[k][0]], s->prof_ev[i * 6 + pairs[k][1]]));
      ms_out4[k] += (double)ms;
    }
  for (cudaEvent_t e : s->prof_ev) cudaEventDestroy(e);
  s->prof_ev.clear();
  *n_iters = (int)n;
  return CDX_OK;
}
