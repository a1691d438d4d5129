// Batched probabilistic pregrasp optimisation (ProbabilisticGraspOptimizer, optimize_pregrasp.py:614-839) without
// autograd and without host synchronisation.  One optimiser iteration is four launches on the caller's stream:
//   pg_points_kernel     float32 FK + reference-compatible Jacobian per (candidate, finger); palm transform; emits the
//                        GPIS query points (pregrasp blend :747-750, targets, tips, palm position)
//   gpis_var_kernel<EXT> mean, grad mean, std, grad std at the blended contact points (cdx_query.cu, FP64 DMMA)
//   gpis_mean_kernel     mean, grad mean at targets / tips / palm (cdx_query.cu, TMA-staged training set)
//   pg_objective_kernel  per candidate: Kabsch (3x3 Jacobi SVD) equilibrium, all cost terms, ANALYTIC reverse pass
//                        (cdx_pg_math.cuh), best-iterate tracking (:822-829), Adam (:831) and the clamps (:833-834)
#include <math.h>
#include <vector>
#include "cdx_internal.cuh"
#include "cdx_noise.cuh"
#include "cdx_pg_math.cuh"

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
  int* it_dev = nullptr;     // absolute iteration index (keys the noise stream, Adam bias correction, s > 20)
  int it_host = 0;
  double* adam_tab = nullptr;  // [tab_cap][2]: 1 - beta1^t, sqrt(1 - beta2^t) for t = index + 1
  int tab_cap = 0;
  cdx_hand_dev* hand_dev = nullptr;
  bool prof = false;
  std::vector<cudaEvent_t> prof_ev;   // 5 events per profiled iteration
};

namespace {

constexpr int NPAR = 38;  // q16 | k4 | T12 | palm6

struct pg_lr { double v[5]; };

__global__ void __launch_bounds__(128)
pg_points_kernel(const cdx_hand_dev* __restrict__ hand, int64_t B, int U, cdx_pg_setup S, const double* __restrict__ q,
                 const double* __restrict__ T, const double* __restrict__ palm, double* __restrict__ tl,
                 float* __restrict__ J, double* __restrict__ Xext, double* __restrict__ Xmean,
                 double* __restrict__ tips_out) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= B * 4) return;
  const int64_t b = gid >> 2;
  const int f = (int)(gid & 3);
  float q32[16];
#pragma unroll
  for (int i = 0; i < 16; i++) q32[i] = i < S.n_dofs ? (float)q[b * S.n_dofs + i] : 0.f;   // q.float(), :665
  float tip[3], Jf[3][CDX_MAX_CHAIN_DOF];
  cdx_fk_finger(hand->finger[f], q32, tip, Jf);
  const double tlf[3] = {(double)tip[0], (double)tip[1], (double)tip[2]};                   // .double(), :666
  double Rw[9], tw[3];
  cdx_euler_xyz(palm + b * 6 + 3, Rw, nullptr, nullptr, nullptr);
  cdx_mv(Rw, tlf, tw);
  for (int r = 0; r < 3; r++) tw[r] += palm[b * 6 + r];
  if (tips_out) {
    for (int r = 0; r < 3; r++) tips_out[gid * 3 + r] = tw[r];
    return;
  }
  for (int r = 0; r < 3; r++) {
    tl[gid * 3 + r] = tlf[r];
    for (int c = 0; c < CDX_MAX_CHAIN_DOF; c++) J[(gid * 3 + r) * CDX_MAX_CHAIN_DOF + c] = Jf[r][c];
  }
  const double* Tf = T + gid * 3;
  int seen = 0;
  for (int p = 0; p < S.P; p++) {
    const int u = S.uniq[p];
    if (u < seen) continue;   // this coefficient row already has its query block
    seen = u + 1;
    double* x = Xext + ((b * U + u) * 4 + f) * 3;
    for (int r = 0; r < 3; r++) x[r] = Tf[r] + S.coeff[p][f] * (tw[r] - Tf[r]);             // :750
  }
  double* xm = Xmean + b * 27;
  for (int r = 0; r < 3; r++) {
    xm[f * 3 + r] = Tf[r];
    xm[(4 + f) * 3 + r] = tw[r];
    if (f == 0) xm[24 + r] = palm[b * 6 + r];
  }
}

__global__ void pg_tick_kernel(int* it) { *it += 1; }
__global__ void pg_fill_kernel(double* p, int64_t n, double v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// NaN-propagating clamps (torch.clamp keeps NaN; fmax/fmin would not)
__device__ __forceinline__ double clamp_lo(double x, double lo) { return x < lo ? lo : x; }
__device__ __forceinline__ double clamp_hi(double x, double hi) { return x > hi ? hi : x; }

__global__ void __launch_bounds__(64)
pg_objective_kernel(const cdx_hand_dev* __restrict__ hand, int64_t B, int U, cdx_pg_setup S, cdx_pregrasp_params prm,
                    double* __restrict__ q, double* __restrict__ k, double* __restrict__ T, double* __restrict__ palm,
                    const double* __restrict__ tl, const float* __restrict__ J, const double* __restrict__ dist,
                    const double* __restrict__ sd, const double* __restrict__ gmu, const double* __restrict__ gsd,
                    const double* __restrict__ mean9, const double* __restrict__ gmean9, double* __restrict__ loss,
                    double* __restrict__ margin, double* __restrict__ grad, int do_step, const int* __restrict__ it_dev,
                    const double* __restrict__ adam_tab, double* __restrict__ m1, double* __restrict__ m2,
                    double* __restrict__ oq, double* __restrict__ ok, double* __restrict__ oT,
                    double* __restrict__ opalm, double* __restrict__ omargin, double* __restrict__ ovalue) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int it = *it_dev;
  const int nd = S.n_dofs;
  double par[NPAR], g[NPAR];
  for (int i = 0; i < 16; i++) par[i] = i < nd ? q[b * nd + i] : 0.0;
  for (int i = 0; i < 4; i++) par[16 + i] = k[b * 4 + i];
  for (int i = 0; i < 12; i++) par[20 + i] = T[b * 12 + i];
  for (int i = 0; i < 6; i++) par[32 + i] = palm[b * 6 + i];
  cdx_pg_gpis G = {dist + b * U * 4, sd + b * U * 4, gmu + b * U * 12, gsd + b * U * 12, mean9 + b * 9, gmean9 + b * 27};
  const uint64_t seed = prm.noise_seed;
  const int64_t gb = prm.env_offset + b;
  auto nf = [=](int p, int e) { return cdx_kabsch_noise(seed, it, p, gb, e); };
  double total, tm[4];
  for (int i = 0; i < 16; i++) g[i] = 0.0;
  cdx_pg_closure(S, *hand, par, par + 16, reinterpret_cast<const double (*)[3]>(par + 20), par + 32,
                 reinterpret_cast<const double (*)[3]>(tl + b * 12),
                 reinterpret_cast<const float (*)[3][CDX_MAX_CHAIN_DOF]>(J + b * 12 * CDX_MAX_CHAIN_DOF), G, nf, total, tm,
                 g, g + 16, reinterpret_cast<double (*)[3]>(g + 20), g + 32);
  if (loss) loss[b] = total;
  if (margin) for (int i = 0; i < 4; i++) margin[b * 4 + i] = tm[i];
  if (grad) for (int i = 0; i < NPAR; i++) grad[b * NPAR + i] = g[i];
  if (!do_step) return;
  // best-iterate tracking (:822-829): BEFORE the optimiser step, only for s > track_after; NaN never updates
  if (it > prm.track_after && total < ovalue[b]) {
    ovalue[b] = total;
    for (int i = 0; i < 4; i++) omargin[b * 4 + i] = tm[i];
    for (int i = 0; i < nd; i++) oq[b * nd + i] = par[i];
    for (int i = 0; i < 4; i++) ok[b * 4 + i] = par[16 + i];
    for (int i = 0; i < 12; i++) oT[b * 12 + i] = par[20 + i];
    for (int i = 0; i < 6; i++) opalm[b * 6 + i] = par[32 + i];
  }
  // torch.optim.Adam (:796,:831): exp_avg.lerp_(g, 1-b1); exp_avg_sq = b2 v + (1-b2) g^2;
  // p -= (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
  const double bc1 = adam_tab[2 * it], bc2s = adam_tab[2 * it + 1];
  const double omb1 = 1.0 - prm.beta1, omb2 = 1.0 - prm.beta2;
  for (int i = 0; i < NPAR; i++) {
    int grp;
    if (i < 16) { if (i >= nd) continue; grp = 0; }
    else if (i < 20) grp = 1;
    else if (i < 32) { if (!prm.optimize_target) continue; grp = 2; }
    else { if (!prm.optimize_palm) continue; grp = i < 35 ? 3 : 4; }
    const double gi = g[i];
    double m = m1[b * NPAR + i], v = m2[b * NPAR + i];
    m = m + omb1 * (gi - m);
    v = v * prm.beta2 + omb2 * gi * gi;
    m1[b * NPAR + i] = m;
    m2[b * NPAR + i] = v;
    const double step = prm.lr[grp] / bc1;
    const double denom = sqrt(v) / bc2s + prm.eps;
    par[i] = par[i] - step * (m / denom);
  }
  for (int i = 0; i < 4; i++) par[16 + i] = clamp_lo(par[16 + i], prm.compliance_min);           // :833
  for (int f = 0; f < 4; f++)
    for (int r = 0; r < 3; r++)                                                                 // :834 (always applied)
      par[20 + f * 3 + r] = clamp_hi(clamp_lo(par[20 + f * 3 + r], (double)prm.box_lo[f][r]), (double)prm.box_hi[f][r]);
  for (int i = 0; i < nd; i++) q[b * nd + i] = par[i];
  for (int i = 0; i < 4; i++) k[b * 4 + i] = par[16 + i];
  for (int i = 0; i < 12; i++) T[b * 12 + i] = par[20 + i];
  for (int i = 0; i < 6; i++) palm[b * 6 + i] = par[32 + i];
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

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
static void pg_free_buffers(cdx_pg_state* s) {
  void* ptrs[] = {s->q, s->k, s->T, s->palm, s->m1, s->m2, s->oq, s->ok, s->oT, s->opalm, s->omargin, s->ovalue, s->tl,
                  s->J, s->Xext, s->Xmean, s->dist, s->sd, s->gmu, s->gsd, s->mean9, s->gmean9, s->loss, s->margin,
                  s->grad};
  for (void* p : ptrs) cudaFree(p);
  s->q = s->k = s->T = s->palm = s->m1 = s->m2 = s->oq = s->ok = s->oT = s->opalm = s->omargin = s->ovalue = nullptr;
  s->tl = s->Xext = s->Xmean = s->dist = s->sd = s->gmu = s->gsd = s->mean9 = s->gmean9 = s->loss = s->margin = s->grad = nullptr;
  s->J = nullptr;
  s->cap = 0;
}

void cdx_pregrasp_free(cdx_ctx* ctx) {
  if (!ctx || !ctx->pg) return;
  pg_free_buffers(ctx->pg);
  cudaFree(ctx->pg->it_dev);
  cudaFree(ctx->pg->adam_tab);
  cudaFree(ctx->pg->hand_dev);
  delete ctx->pg;
  ctx->pg = nullptr;
}

static cdx_pg_state* pg_get(cdx_ctx* ctx) {
  if (!ctx->pg) ctx->pg = new cdx_pg_state();
  return ctx->pg;
}

extern "C" int cdx_hand_load(cdx_ctx* ctx, int n_bodies, const int* parent, const int* dof, const int* axis,
                             const float* axis_sign, const float* fixed_rot, const float* xyz, int n_dofs,
                             const int* tip_body, const float* tip_offset) {
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
  if (!s->hand_dev) CDX_CUDA(cudaMalloc((void**)&s->hand_dev, sizeof(cdx_hand_dev)));
  CDX_CUDA(cudaMemcpy(s->hand_dev, &h, sizeof(h), cudaMemcpyHostToDevice));
  s->setup.n_dofs = n_dofs;
  ctx->hand_ready = true;
  return CDX_OK;
}

extern "C" int cdx_hand_fk(cdx_ctx* ctx, const double* q, const double* palm, int64_t B, double* tips, void* stream) {
  if (!ctx || !ctx->hand_ready) { cdx_set_error("cdx_hand_fk: no hand loaded"); return CDX_ERR_STATE; }
  if (B < 0 || (B > 0 && (!q || !palm || !tips))) { cdx_set_error("cdx_hand_fk: bad argument"); return CDX_ERR_ARG; }
  if (B == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_pg_state* s = ctx->pg;
  cdx_pg_setup S;
  memset(&S, 0, sizeof(S));
  S.n_dofs = s->setup.n_dofs;
  pg_points_kernel<<<(unsigned)((B * 4 + 127) / 128), 128, 0, (cudaStream_t)stream>>>(s->hand_dev, B, 1, S, q, nullptr, palm, nullptr,
                                                                              nullptr, nullptr, nullptr, tips);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

#define PG_ALLOC(ptr, n) CDX_CUDA(cudaMalloc((void**)&(ptr), sizeof(*(ptr)) * (size_t)(n)))

extern "C" int cdx_pregrasp_init(cdx_ctx* ctx, const cdx_pregrasp_params* p, int64_t B, const double* q0,
                                 const double* k0, const double* T0, const double* palm0, void* stream) {
  if (!ctx || !ctx->hand_ready) { cdx_set_error("cdx_pregrasp_init: cdx_hand_load first"); return CDX_ERR_STATE; }
  if (!p || B < 1 || !q0 || !k0 || !T0 || !palm0 || p->num_copies < 1 || p->num_copies > CDX_MAX_COPIES) {
    cdx_set_error("cdx_pregrasp_init: bad argument (1 <= num_copies <= %d, B >= 1)", CDX_MAX_COPIES);
    return CDX_ERR_ARG;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  s->prm = *p;
  cdx_pg_setup& S = s->setup;
  const int nd = S.n_dofs;
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
    s->cap = B;
    s->U = U;
  }
  if (!s->it_dev) PG_ALLOC(s->it_dev, 1);
  s->B = B;
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
  CDX_CUDA(cudaMemsetAsync(s->it_dev, 0, sizeof(int), st));
  pg_fill_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(s->ovalue, B, INFINITY);
  CDX_CHECK_LAUNCH(ctx);
  s->it_host = 0;
  return CDX_OK;
}

// one closure evaluation: points -> GPIS queries -> objective (+ optional optimiser step)
static int pg_iteration(cdx_ctx* ctx, cdx_pg_state* s, int do_step, double* loss, double* margin, double* grad,
                        cudaStream_t st) {
  const int64_t B = s->B;
  const int U = s->U;
  const bool prof = s->prof && do_step;
  auto mark = [&]() {
    if (!prof) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    s->prof_ev.push_back(e);
  };
  mark();
  pg_points_kernel<<<(unsigned)((B * 4 + 127) / 128), 128, 0, st>>>(s->hand_dev, B, U, s->setup, s->q, s->T, s->palm, s->tl,
                                                                  s->J, s->Xext, s->Xmean, nullptr);
  CDX_CHECK_LAUNCH(ctx);
  mark();
  int rc = cdx_gpis_query(ctx, s->Xext, B * U * 4, s->dist, s->gmu, s->sd, s->gsd, (void*)st);
  if (rc) return rc;
  mark();
  rc = cdx_gpis_query(ctx, s->Xmean, B * 9, s->mean9, s->gmean9, nullptr, nullptr, (void*)st);
  if (rc) return rc;
  mark();
  pg_objective_kernel<<<(unsigned)((B + 63) / 64), 64, 0, st>>>(
      s->hand_dev, B, U, s->setup, s->prm, s->q, s->k, s->T, s->palm, s->tl, s->J, s->dist, s->sd, s->gmu, s->gsd, s->mean9,
      s->gmean9, loss, margin, grad, do_step, s->it_dev, s->adam_tab, s->m1, s->m2, s->oq, s->ok, s->oT, s->opalm,
      s->omargin, s->ovalue);
  CDX_CHECK_LAUNCH(ctx);
  mark();
  return CDX_OK;
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
  CDX_CUDA(cudaStreamSynchronize(st));
  cudaFree(s->adam_tab);
  s->adam_tab = nullptr;
  s->tab_cap = 0;
  CDX_CUDA(cudaMalloc((void**)&s->adam_tab, sizeof(double) * tab.size()));
  CDX_CUDA(cudaMemcpy(s->adam_tab, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice));
  s->tab_cap = cap;
  return CDX_OK;
}

extern "C" int cdx_pregrasp_run(cdx_ctx* ctx, int n_iters, void* stream) {
  int rc = pg_ready(ctx, "cdx_pregrasp_run");
  if (rc) return rc;
  if (n_iters < 0) { cdx_set_error("cdx_pregrasp_run: n_iters < 0"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  rc = pg_adam_table(s, s->it_host + n_iters, st);
  if (rc) return rc;
  for (int i = 0; i < n_iters; i++) {
    rc = pg_iteration(ctx, s, 1, s->loss, s->margin, nullptr, st);
    if (rc) return rc;
    pg_tick_kernel<<<1, 1, 0, st>>>(s->it_dev);
    CDX_CHECK_LAUNCH(ctx);
  }
  s->it_host += n_iters;
  return CDX_OK;
}

extern "C" int cdx_pregrasp_eval(cdx_ctx* ctx, double* total_loss, double* margin, double* gq, double* gc, double* gt,
                                 double* gp, void* stream) {
  int rc = pg_ready(ctx, "cdx_pregrasp_eval");
  if (rc) return rc;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  rc = pg_adam_table(s, s->it_host + 1, st);
  if (rc) return rc;
  rc = pg_iteration(ctx, s, 0, s->loss, s->margin, s->grad, st);
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

extern "C" int cdx_pregrasp_results(cdx_ctx* ctx, double* oq, double* ok, double* oT, double* opalm, double* omargin,
                                    double* ovalue, void* stream) {
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_results: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  const int64_t B = s->B;
  PG_COPY(oq, s->oq, B * s->setup.n_dofs); PG_COPY(ok, s->ok, B * 4); PG_COPY(oT, s->oT, B * 12);
  PG_COPY(opalm, s->opalm, B * 6); PG_COPY(omargin, s->omargin, B * 4); PG_COPY(ovalue, s->ovalue, B);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_state(cdx_ctx* ctx, double* q, double* k, double* T, double* palm, void* stream) {
  if (!ctx || !ctx->pg || ctx->pg->B == 0) { cdx_set_error("cdx_pregrasp_state: cdx_pregrasp_init first"); return CDX_ERR_STATE; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_pg_state* s = ctx->pg;
  const int64_t B = s->B;
  PG_COPY(q, s->q, B * s->setup.n_dofs); PG_COPY(k, s->k, B * 4); PG_COPY(T, s->T, B * 12); PG_COPY(palm, s->palm, B * 6);
  return CDX_OK;
}

extern "C" int cdx_pregrasp_argmin(cdx_ctx* ctx, double* out, void* stream) {
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
  const size_t n = s->prof_ev.size() / 5;
  for (size_t i = 0; i < n; i++)
    for (int k = 0; k < 4; k++) {
      float ms = 0.f;
      CDX_CUDA(cudaEventElapsedTime(&ms, s->prof_ev[i * 5 + k], s->prof_ev[i * 5 + k + 1]));
      ms_out4[k] += (double)ms;
    }
  for (cudaEvent_t e : s->prof_ev) cudaEventDestroy(e);
  s->prof_ev.clear();
  *n_iters = (int)n;
  return CDX_OK;
}
