// Internal definitions shared by the cdx translation units (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>
#include "../../include/cdx.h"

#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op unless a profiler injects its library

// NVTX range covering a C-ABI entry point (tracing hooks at the boundary)
struct cdx_nvtx_range {
  explicit cdx_nvtx_range(const char* name) { nvtxRangePushA(name); }
  ~cdx_nvtx_range() { nvtxRangePop(); }
};
#define CDX_NVTX(name) cdx_nvtx_range cdx_nvtx_range_##__LINE__(name)

#define CDX_MAX_BODIES 32
#define CDX_NPAD 64  // matrices are padded to a multiple of this (identity on the padded diagonal)

void cdx_set_error(const char* fmt, ...);

#define CDX_CUDA(call)                                                                      \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      cdx_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));  \
      return CDX_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)

#define CDX_CHECK_LAUNCH(ctx)                                                               \
  do {                                                                                      \
    (ctx)->launches++;                                                                      \
    cudaError_t e__ = cudaGetLastError();                                                   \
    if (e__ != cudaSuccess) {                                                               \
      cdx_set_error("%s:%d kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return CDX_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)

struct cdx_hand {
  int n_bodies = 0, n_dofs = 0;
  int parent[CDX_MAX_BODIES], dof[CDX_MAX_BODIES], axis[CDX_MAX_BODIES];
  float sign[CDX_MAX_BODIES];
  float fixed_rot[CDX_MAX_BODIES][9];  // (Rz(yaw) Ry(pitch)) Rx(roll) in float32, rigid_body.py:142
  float xyz[CDX_MAX_BODIES][3];
  int tip_body[4];
  float tip_offset[4][3];
};

struct cdx_gpis_state {
  int64_t N = 0, Np = 0;  // Np = N rounded up to CDX_NPAD
  int kernel = 0;
  double sigma = 0, R = 0, bias = 0, prior = 0;  // prior = k(x,x)
  double* Xa = nullptr;     // [Np,4] (x,y,z,alpha); padded rows are zero
  double* y1 = nullptr;     // [Np]
  double* E11 = nullptr;    // [Np,Np] padded with identity
  double* L = nullptr;      // [Np,Np] lower Cholesky factor (upper part zero)
  double* U = nullptr;      // [Np,Np] L^-1 (lower, upper part zero)
  double* P = nullptr;      // [Np,Np] E11^-1 = U^T U (full symmetric)
  double* Pslab = nullptr;  // states of <= 448 padded points: E11^-1 in slab order [Np/16][Np][16], 16-byte pieces XOR-swizzled by the
                            // column index (piece ^ 2 (n & 3)) - the bulk-TMA stream of gpis_extw_kernel (cdx_query.cu)
  double* Dinv = nullptr;   // [Np/64, 64, 64] inverses of the diagonal blocks of L
  double* tmp = nullptr;    // [Np] scratch vector
  double* dR2 = nullptr;    // device scalar: max pairwise squared distance
  int* flag = nullptr;      // device int: first non-SPD column + 1 (0 = ok)
  int64_t cap = 0;          // allocated Np
  bool ready = false;
  // INT8-sliced copy of U for the tcgen05 std path (cdx_query_i8x.cu), built lazily at the first eligible query: the digits
  // tiled in the shared-memory image of the MMA's A operand
  int8_t* i8x_A = nullptr;  // [row tiles 128][Np/32][S][2][128][16]
  double* i8x_sb = nullptr; // [rows] 2^e_n (0 on padded rows)
  int* i8x_eb = nullptr;
  double* i8x_pts = nullptr;
  double* i8x_bb = nullptr;
  int i8x_S = 0;
  size_t i8x_cap = 0;
  int64_t i8x_rows = 0;
  int8_t* i8x_kq = nullptr;    // scratch: k* byte slices of one chunk of queries in the B-operand image (CTA-pair path)
  double* i8x_scq = nullptr;   // per-query fixed-point scales of the chunk
  double* i8x_part = nullptr;  // [parts][queries][4] partial sums of the mean and its gradient
  size_t i8x_kq_cap = 0;
  int i8_variant = 3;       // cdx_gpis_set_i8_variant: 3 = CTA pair (cta_group::2), 1 = 2-CTA cluster sharing k* over DSMEM, 2 = no cluster
  cudaStream_t fit_side = nullptr;                      // second stream of the look-ahead Cholesky
  cudaEvent_t fit_ev_trsm = nullptr, fit_ev_rest = nullptr;
  // optional phase timing of the fit (cdx_gpis_fit_profile): events at the phase boundaries
  bool fit_prof = false;
  cudaEvent_t fit_ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t i8_ready = nullptr;  // recorded after the slices were built: queries on any other stream wait for it
  int ext_mode = 0;         // cdx_gpis_set_ext_mode: 0 = warp-specialised all-columns extended kernel for states <= 448 padded points, 2 = its single-role predecessor, 1 = column-tiled kernel always
  int var_mode = 0;         // cdx_gpis_set_var_mode: 0 auto, -1 always FP64 DMMA, 4/5/6 = INT8 slices whenever eligible
};

struct cdx_pg_state;  // cdx_pregrasp.cu
struct cdx_sdf_state; // cdx_sdf.cu

struct cdx_ctx {
  int device = 0;
  int sms = 148;
  int64_t launches = 0;
  cdx_gpis_state g;
  cdx_hand hand;
  bool hand_ready = false;
  cdx_pg_state* pg = nullptr;
  cdx_sdf_state* sdf = nullptr;
  // scratch for host-pointer entry points
  double* hx = nullptr; double* hout = nullptr; int64_t hcap = 0;
  // kernels whose dynamic shared-memory limit has been raised ON THIS DEVICE (the attribute is per device, one ctx per device)
  std::vector<const void*> smem_attr_done;
};

template <class F>
static inline int cdx_smem_attr(cdx_ctx* ctx, F* func, size_t bytes) {
  const void* key = reinterpret_cast<const void*>(func);
  for (const void* k : ctx->smem_attr_done)
    if (k == key) return CDX_OK;
  CDX_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  ctx->smem_attr_done.push_back(key);
  return CDX_OK;
}

// covariance function and radial derivative factor:  k(d), c(d) with grad_x k = c * (x - xi)
struct cdx_kparams {
  int kernel;
  double R, R3, inv_s2;  // inv_s2 = 1/sigma^2
};

__device__ __forceinline__ double cdx_fast_dist(double s) {
  // d = sqrt(s) for s >= 0: float rsqrt seed + two Newton steps on the residual (each: exact residual by FMA).
  float sf = fmaxf((float)s, 1e-30f);
  double y = (double)rsqrtf(sf);
  double d = s * y;
  double hy = 0.5 * y;
  d = fma(fma(-d, d, s), hy, d);
  d = fma(fma(-d, d, s), hy, d);
  return d;
}

template <int KERNEL>
__device__ __forceinline__ void cdx_kernel_eval(double s /* squared distance */, const cdx_kparams& kp, double& k,
                                                double& c) {
  if (KERNEL == CDX_KERNEL_TPS) {
    double d = cdx_fast_dist(s);
    k = fma(s, fma(2.0, d, -3.0 * kp.R), kp.R3);  // 2d^3 - 3R d^2 + R^3   (gpis.py:26)
    c = 6.0 * (d - kp.R);
  } else if (KERNEL == CDX_KERNEL_RBF) {
    k = exp(-0.5 * s * kp.inv_s2);                 // gpis.py:18-19
    c = -k * kp.inv_s2;
  } else {
    double d = cdx_fast_dist(s);
    double kt = fma(s, fma(2.0, d, -3.0 * kp.R), kp.R3);
    double kr = exp(-0.5 * s * kp.inv_s2);
    k = 0.3 * kr + 0.7 * kt;                       // gpis.py:28-29
    c = 0.3 * (-kr * kp.inv_s2) + 0.7 * (6.0 * (d - kp.R));
  }
}

// k only, for NQ squared distances at once, written stage by stage so that the NQ dependent chains (conversion, rsqrt seed,
// two Newton steps, cubic) overlap; each value is the one cdx_kernel_eval returns, bit for bit
template <int KERNEL, int NQ>
__device__ __forceinline__ void cdx_kernel_k_batch(const double (&s)[NQ], const cdx_kparams& kp, double (&k)[NQ]) {
  if (KERNEL == CDX_KERNEL_RBF) {
#pragma unroll
    for (int i = 0; i < NQ; i++) k[i] = exp(-0.5 * s[i] * kp.inv_s2);
    return;
  }
  float sf[NQ];
  double y[NQ], d[NQ], hy[NQ];
#pragma unroll
  for (int i = 0; i < NQ; i++) sf[i] = fmaxf((float)s[i], 1e-30f);
#pragma unroll
  for (int i = 0; i < NQ; i++) y[i] = (double)rsqrtf(sf[i]);
#pragma unroll
  for (int i = 0; i < NQ; i++) { d[i] = s[i] * y[i]; hy[i] = 0.5 * y[i]; }
#pragma unroll
  for (int i = 0; i < NQ; i++) d[i] = fma(fma(-d[i], d[i], s[i]), hy[i], d[i]);
#pragma unroll
  for (int i = 0; i < NQ; i++) d[i] = fma(fma(-d[i], d[i], s[i]), hy[i], d[i]);
#pragma unroll
  for (int i = 0; i < NQ; i++) {
    const double kt = fma(s[i], fma(2.0, d[i], -3.0 * kp.R), kp.R3);
    k[i] = KERNEL == CDX_KERNEL_TPS ? kt : 0.3 * exp(-0.5 * s[i] * kp.inv_s2) + 0.7 * kt;
  }
}

static inline cdx_kparams cdx_make_kparams(const cdx_gpis_state& g) {
  cdx_kparams kp;
  kp.kernel = g.kernel;
  kp.R = g.R;
  kp.R3 = g.R * g.R * g.R;
  kp.inv_s2 = g.sigma != 0 ? 1.0 / (g.sigma * g.sigma) : 0.0;
  return kp;
}

// FP64 tensor-core MMA (the only FP64 tensor path on sm_100a; tcgen05 has no f64 kind): D(8x8) += A(8x4) B(4x8).
// fragment ownership (lane T): a = A[T/4][T%4], b = B[T%4][T/4], d0,d1 = D[T/4][2*(T%4) + {0,1}]
__device__ __forceinline__ void cdx_dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ void cdx_cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cdx_cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cdx_cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// generic FP64 GEMM used by the fit (cdx_gemm.cu):  C[M,N] = alpha * op(A) op(B) + beta * C, row-major.
// ta: A is stored [K,M] (else [M,K]);  tb: B is stored [K,N] (else [N,K], i.e. C = A B^T).
// ktri: skip k ranges known to be zero (1: k >= n0, 2: k < m0+64, 3: k >= m0; 0: none).
// lower_only: skip tiles strictly above the diagonal (requires M == N tiles aligned).  batch via strides.
// tcgen05 INT8-sliced std query (cdx_query_i8x.cu); S slices in {4,5,6}
bool cdx_i8_eligible(const cdx_ctx* ctx);
int cdx_i8x_std(cdx_ctx* ctx, int S, int cluster, const double* X, int64_t M, double* stdv, double* mean, double* gmean,
                bool* mean_done, cudaStream_t st);
void cdx_i8x_free(cdx_ctx* ctx);
void cdx_sdf_free(cdx_ctx* ctx);

int cdx_dgemm(cdx_ctx* ctx, bool ta, bool tb, int M, int N, int K, double alpha, const double* A, int64_t lda,
              const double* B, int64_t ldb, double beta, double* C, int64_t ldc, bool lower_only, int ktri, int batch,
              int64_t sA, int64_t sB, int64_t sC, cudaStream_t st);
