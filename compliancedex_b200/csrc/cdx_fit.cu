// GPIS fit: covariance assembly (gpis.py:33-40) and the in-house blocked FP64 factorisation that replaces the
// reference's per-call LU (gpis.py:53) / explicit inverse (gpis.py:77):
//   E11 = L L^T (right-looking: 256 / 512-column outer panels of 64-column sub-steps, one DMMA SYRK trailing update per panel)
//   U = L^-1 (block-recursive doubling, DMMA GEMMs),  P = U^T U = E11^-1,  alpha = U^T (U y1).
#include <stdarg.h>
#include "cdx_internal.cuh"

static thread_local char g_err[512] = "";
void cdx_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* cdx_last_error(void) { return g_err; }
extern "C" int cdx_version(void) { return 100; }

namespace {

__global__ void pairmax_kernel(const double* __restrict__ X, int N, int jlen, unsigned long long* out) {
  // thread = point i, blockIdx.y = a range of jlen partners (the maximum does not depend on the order: same R bit for bit)
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j0 = blockIdx.y * jlen, j1 = min(N, j0 + jlen);
  double best = 0.0;
  if (i < N) {
    double x = X[3 * i], y = X[3 * i + 1], z = X[3 * i + 2];
    for (int j = j0; j < j1; j++) {
      double dx = x - X[3 * j], dy = y - X[3 * j + 1], dz = z - X[3 * j + 2];
      best = fmax(best, fma(dx, dx, fma(dy, dy, dz * dz)));
    }
  }
  for (int o = 16; o; o >>= 1) best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(best));  // >= 0: ordered as u64
}

template <int KERNEL>
__global__ void assemble_kernel(const double* __restrict__ X, const double* __restrict__ noise, double noise_scalar,
                                int N, int Np, cdx_kparams kp, double* __restrict__ E, double* __restrict__ Lcopy) {
  int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= Np) return;
  double v;
  if (i < N && j < N) {
    double dx = X[3 * i] - X[3 * j], dy = X[3 * i + 1] - X[3 * j + 1], dz = X[3 * i + 2] - X[3 * j + 2];
    double k, c;
    cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, c);
    v = k;
    if (i == j) {
      double n = noise ? noise[i] : noise_scalar;
      v += n * n;
    }
  } else {
    v = (i == j) ? 1.0 : 0.0;
  }
  E[(int64_t)i * Np + j] = v;
  Lcopy[(int64_t)i * Np + j] = v;      // the factorisation works in place on this copy (saves a 2 x Np^2 x 8 B device copy)
}

__global__ void pad_copy_kernel(const double* __restrict__ src, int N, int Np, double* __restrict__ dst) {
  int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= Np) return;
  dst[(int64_t)i * Np + j] = (i < N && j < N) ? src[(int64_t)i * N + j] : ((i == j) ? 1.0 : 0.0);
}

__global__ void unpad_copy_kernel(const double* __restrict__ src, int N, int Np, double* __restrict__ dst) {
  int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j < N) dst[(int64_t)i * N + j] = src[(int64_t)i * Np + j];
}

// One level of the recursive-doubling inverse of a 64x64 lower-triangular block (a = L, x = the inverse so far: the diagonal
// S x S blocks are done): X21 = -X22 (L21 X11) for the 32 / S block pairs.  S is a compile-time constant and the dot products
// run over the FULL block width with four independent accumulators - the triangular halves they could skip are exact zeros of
// x, and a fixed trip count lets the loads pipeline; the run-time version (trip counts j..S, one dependent FMA chain per
// output, integer divisions for the indices) took 32 k of the kernel's 70 k cycles.
template <int S>
__device__ __forceinline__ void potrf_inv_level(double (*a)[65], double (*x)[65], int t) {
  constexpr int NOUT = 32 * S, SS = S * S;
  for (int e = t; e < NOUT; e += 256) {          // T = L21 X11 -> a[upper mirror block] (dead storage)
    const int p = e / SS, i = (e / S) % S, j = e % S, base = 2 * p * S;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int k = 0; k < S; k++) acc[k & 3] = fma(a[base + S + i][base + k], x[base + k][base + j], acc[k & 3]);
    a[base + i][base + S + j] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
  }
  __syncthreads();
  for (int e = t; e < NOUT; e += 256) {          // X21 = -X22 T
    const int p = e / SS, i = (e / S) % S, j = e % S, base = 2 * p * S;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int k = 0; k < S; k++) acc[k & 3] = fma(x[base + S + i][base + S + k], a[base + k][base + S + j], acc[k & 3]);
    x[base + S + i][base + j] = -((acc[0] + acc[1]) + (acc[2] + acc[3]));
  }
  __syncthreads();
}

// Cholesky of one 64x64 diagonal block + its triangular inverse.  This kernel is the head of the serial chain of every
// 64-column sub-step, so its latency counts in full (N = 4,096: the whole Cholesky phase is this chain):
//   * factorisation in REGISTERS: thread (row r = t % 64, column group t / 64) holds the 16 entries a[r][16 cg .. 16 cg + 15];
//     per column j the owners publish the raw column through a double-buffered 64-entry line of shared memory (ONE barrier per
//     column), every thread forms 1 / sqrt(a[j][j]) itself (rsqrt + the diagonal by one Newton step: no sqrt / division chain,
//     no wait for a broadcast) and updates its own entries right of j with broadcast reads; the 16 columns of a group are unrolled, so all
//     register subscripts are compile-time.  Entries above the diagonal carry dead values and are zeroed at the write-back.
//     (The shared-memory version - two barriers, an IEEE sqrt and a division per column, 16 predicated passes over the whole
//     block per column - took ~50 us of the ~100 us sub-step.)
//   * inverse: recursive doubling, inv([L11 0; L21 L22]) = [X11 0; -X22 L21 X11, X22], all 256 threads at every level (six
//     levels, two barriers each, potrf_inv_level above) instead of 64 serial forward substitutions; the s x s temporaries
//     T = L21 X11 live in the mirror-image block of `a`'s upper triangle, which is dead (zeroed before the write-back).
// Shared memory: 2 x 64 x 65 doubles (+ the 1 KB column line).
__global__ void __launch_bounds__(256) potrf_diag_kernel(double* __restrict__ A, int64_t ld, double* __restrict__ Dinv,
                                                         double* __restrict__ Udiag, int* flag, int col0) {
  extern __shared__ double potrf_smem[];
  double (*a)[65] = reinterpret_cast<double (*)[65]>(potrf_smem);
  double (*x)[65] = a + 64;
  __shared__ __align__(16) double colbuf[2][64];
  const int t = threadIdx.x;
  const int r = t & 63, cg = t >> 6;
  double ar[16];
  {
    const double2* src = reinterpret_cast<const double2*>(A + (int64_t)r * ld + cg * 16);   // ld and the block origin are even
#pragma unroll
    for (int k = 0; k < 8; k++) { const double2 v = src[k]; ar[2 * k] = v.x; ar[2 * k + 1] = v.y; }
  }
  // the 16 columns of a column group are unrolled (compile-time register subscripts), the four groups are a real loop: the
  // body is fetched once and re-run from the instruction cache (fully unrolled, the kernel was 7.5 k instructions walked once,
  // cold, per launch: ncu put half of its stall cycles on instruction fetch)
#pragma unroll 1
  for (int cgj = 0; cgj < 4; cgj++) {
#pragma unroll
  for (int jj = 0; jj < 16; jj++) {
    const int j = cgj * 16 + jj;
    double* cb = colbuf[jj & 1];
    if (cg == cgj) cb[r] = ar[jj];
    __syncthreads();          // one barrier per column: the line written for column j + 1 is the one column j - 1 used, and every
                              // thread has left column j - 1 before any thread passes this barrier
    const double d = cb[j];
    const double rinv = rsqrt(d);
    double root = d * rinv;
    root = fma(0.5 * rinv, fma(-root, root, d), root);     // sqrt(d) to the last bit or two
    if (t == cgj * 64 + j && !(d > 0.0)) atomicCAS(flag, 0, col0 + j + 1);
    const double lr = cb[r] * rinv;
    if (cg > cgj) {
#pragma unroll
      for (int k = 0; k < 16; k += 2) {
        const double2 c2 = *reinterpret_cast<const double2*>(cb + cg * 16 + k);
        ar[k] = fma(-lr, c2.x * rinv, ar[k]);
        ar[k + 1] = fma(-lr, c2.y * rinv, ar[k + 1]);
      }
    } else if (cg == cgj) {
#pragma unroll
      for (int k = 0; k < 16; k++)
        if (k > jj) ar[k] = fma(-lr, cb[cgj * 16 + k] * rinv, ar[k]);
      ar[jj] = (r == j) ? root : lr;
    }
  }
  }
#pragma unroll
  for (int k = 0; k < 16; k++) a[r][cg * 16 + k] = ar[k];
  __syncthreads();
  if (t < 64) colbuf[0][t] = 1.0 / a[t][t];       // (all threads left the last column's line at the barrier above)
  __syncthreads();
  for (int e = t; e < 4096; e += 256) {
    int rr = e >> 6, c = e & 63;
    x[rr][c] = rr == c ? colbuf[0][rr] : 0.0;
  }
  __syncthreads();
  potrf_inv_level<1>(a, x, t);  potrf_inv_level<2>(a, x, t);  potrf_inv_level<4>(a, x, t);
  potrf_inv_level<8>(a, x, t);  potrf_inv_level<16>(a, x, t); potrf_inv_level<32>(a, x, t);
  for (int e = t; e < 4096; e += 256) {
    int r = e >> 6, c = e & 63;
    const double v = c > r ? 0.0 : a[r][c];
    A[(int64_t)r * ld + c] = v;
    Dinv[e] = x[r][c];
    Udiag[(int64_t)r * ld + c] = x[r][c];
  }
}

// E11^-1 in the slab order of gpis_extw_kernel: [chunk c][column n][16 k], the 16-byte piece index XORed with 2 (n & 3)
__global__ void pack_slab_kernel(const double* __restrict__ P, int Np, double* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x, n = blockIdx.y;
  if (k >= Np) return;
  const int c = k >> 4, kk = k & 15;
  out[((size_t)c * Np + n) * 16 + ((((kk >> 1) ^ ((n & 3) << 1)) << 1) | (kk & 1))] = P[(size_t)n * Np + k];   // P is symmetric
}
__global__ void zero_upper_kernel(double* __restrict__ A, int Np) {
  int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j < Np && j > i) A[(int64_t)i * Np + j] = 0.0;
}
__global__ void mirror_lower_kernel(double* __restrict__ A, int Np) {
  int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j < Np && j > i) A[(int64_t)i * Np + j] = A[(int64_t)j * Np + i];
}
__global__ void fill_kernel(double* __restrict__ A, int64_t n, double v) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) A[i] = v;
}

// out[i] = sum_{j<=i} U[i][j] v[j]   (one warp per row)
__global__ void trmv_lower_kernel(const double* __restrict__ U, int Np, const double* __restrict__ v,
                                  double* __restrict__ out) {
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= Np) return;
  double s = 0.0;
  for (int j = lane; j <= row; j += 32) s = fma(U[(int64_t)row * Np + j], v[j], s);
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[row] = s;
}
// out[j] = sum_{i>=j} U[i][j] v[i]: a CTA owns 32 columns; thread (r, c) of its 8 x 32 threads walks the rows j0 + r, j0 + r + 8, ...
// of column j0 + c (256-byte row segments, four independent chains per thread), the 8 x 4 partial sums of a column are added
// in a fixed order.  (One thread per column walking all rows took 4.4 ms at N = 16,384: 128 CTAs of serial 16 k-row loops.)
__global__ void __launch_bounds__(256) trmv_lower_t_kernel(const double* __restrict__ U, int Np, const double* __restrict__ v,
                                                           double* __restrict__ out) {
  __shared__ double part[8][33];
  const int c = threadIdx.x & 31, r = threadIdx.x >> 5, j0 = blockIdx.x * 32, j = j0 + c;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  if (j < Np) {
    int i = j0 + r;
    for (; i + 24 < Np; i += 32) {
      const double* u = U + (int64_t)i * Np + j;
      const double u0 = i >= j ? u[0] : 0.0, u1 = i + 8 >= j ? u[(int64_t)8 * Np] : 0.0;
      const double u2 = i + 16 >= j ? u[(int64_t)16 * Np] : 0.0, u3 = i + 24 >= j ? u[(int64_t)24 * Np] : 0.0;
      s0 = fma(u0, v[i], s0); s1 = fma(u1, v[i + 8], s1); s2 = fma(u2, v[i + 16], s2); s3 = fma(u3, v[i + 24], s3);
    }
    for (; i < Np; i += 8)
      if (i >= j) s0 = fma(U[(int64_t)i * Np + j], v[i], s0);
  }
  part[r][c] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (r == 0 && j < Np) {
    double t = part[0][c];
    for (int k = 1; k < 8; k++) t += part[k][c];
    out[j] = t;
  }
}
__global__ void pack_xa_kernel(const double* __restrict__ X, const double* __restrict__ alpha, int N, int Np,
                               double* __restrict__ Xa) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Np) return;
  bool ok = i < N;
  Xa[4 * i + 0] = ok ? X[3 * i] : 0.0;
  Xa[4 * i + 1] = ok ? X[3 * i + 1] : 0.0;
  Xa[4 * i + 2] = ok ? X[3 * i + 2] : 0.0;
  Xa[4 * i + 3] = ok ? alpha[i] : 0.0;
}
__global__ void sub_bias_kernel(const double* __restrict__ y, int N, int Np, double bias, double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Np) out[i] = i < N ? y[i] - bias : 0.0;
}
__global__ void extract_alpha_kernel(const double* __restrict__ Xa, int N, double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) out[i] = Xa[4 * i + 3];
}
__global__ void extract_x_kernel(const double* __restrict__ Xa, int N, double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) { out[3 * i] = Xa[4 * i]; out[3 * i + 1] = Xa[4 * i + 1]; out[3 * i + 2] = Xa[4 * i + 2]; }
}

int gpis_reserve(cdx_ctx* ctx, int64_t N) {
  cdx_gpis_state& g = ctx->g;
  int64_t Np = (N + CDX_NPAD - 1) / CDX_NPAD * CDX_NPAD;
  if (Np > g.cap) {
    cudaFree(g.Xa); cudaFree(g.y1); cudaFree(g.E11); cudaFree(g.L); cudaFree(g.U); cudaFree(g.P);
    cudaFree(g.Dinv); cudaFree(g.tmp); cudaFree(g.Pslab);
    g.Pslab = nullptr;
    g.cap = 0;
    size_t mat = sizeof(double) * (size_t)Np * (size_t)Np;
    if (cudaMalloc(&g.Xa, sizeof(double) * 4 * Np) != cudaSuccess || cudaMalloc(&g.y1, sizeof(double) * Np) != cudaSuccess ||
        cudaMalloc(&g.E11, mat) != cudaSuccess || cudaMalloc(&g.L, mat) != cudaSuccess ||
        cudaMalloc(&g.U, mat) != cudaSuccess || cudaMalloc(&g.P, mat) != cudaSuccess ||
        cudaMalloc(&g.Dinv, sizeof(double) * 64 * Np) != cudaSuccess || cudaMalloc(&g.tmp, sizeof(double) * 2 * Np) != cudaSuccess) {
      cudaGetLastError();
      cdx_set_error("out of device memory for GPIS state N=%lld", (long long)N);
      return CDX_ERR_ALLOC;
    }
    g.cap = Np;
  }
  if (Np <= 448 && !g.Pslab && cudaMalloc(&g.Pslab, sizeof(double) * 448 * 448) != cudaSuccess) {
    cudaGetLastError();
    cdx_set_error("out of device memory for the slab copy of E11^-1");
    return CDX_ERR_ALLOC;
  }
  if (!g.dR2) {
    CDX_CUDA(cudaMalloc(&g.dR2, sizeof(double)));
    CDX_CUDA(cudaMalloc(&g.flag, sizeof(int)));
  }
  g.N = N;
  g.Np = Np;
  g.ready = false;
  g.i8x_S = 0;   // the sliced copy of U belongs to the previous state
  return CDX_OK;
}

static inline void fit_mark(cdx_gpis_state& g, int i, cudaStream_t st) {
  if (g.fit_prof && g.fit_ev[i]) cudaEventRecord(g.fit_ev[i], st);
}

// factorise g.E11 (padded) -> L, U, P, alpha, Xa.  X1 is [N,3] device.
int gpis_factorise(cdx_ctx* ctx, const double* X1, cudaStream_t st, bool l_holds_e11) {
  cdx_gpis_state& g = ctx->g;
  fit_mark(g, 1, st);
  const int Np = (int)g.Np, N = (int)g.N;
  const int64_t ld = Np;
  const size_t mat = sizeof(double) * (size_t)Np * (size_t)Np;
  if (!l_holds_e11) CDX_CUDA(cudaMemcpyAsync(g.L, g.E11, mat, cudaMemcpyDeviceToDevice, st));   // the assembly kernel writes both
  CDX_CUDA(cudaMemsetAsync(g.U, 0, mat, st));
  CDX_CUDA(cudaMemsetAsync(g.flag, 0, sizeof(int), st));
  const int nb = Np / 64;
  const size_t potrf_smem_bytes = sizeof(double) * 2 * 64 * 65;
  CDX_CUDA(cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)potrf_smem_bytes));
  // Right-looking over 256- or 512-column outer panels.  Inside a panel the 64-column sub-steps (64x64 factorisation + inverse,
  // panel solve as a GEMM with the block inverse, update of the panel's own remaining columns) touch only the panel; the
  // trailing matrix is updated ONCE per panel with K = 256 / 512 - a quarter / an eighth of the read-modify-write traffic of 64-wide steps and
  // four times the work per tile launch (with K = 64 every tile's prologue / epilogue outweighed its four k-chunks: 0.29 of
  // the DMMA peak).  One panel of look-ahead: the next panel's columns are updated first, so that the next panel's
  // serial sub-steps run while the rest of the trailing update executes.  The serial chain (sub-steps + look-ahead update)
  // runs on the library's own HIGH-PRIORITY stream `hi`, the bulk update on the caller's stream: the chain's small grids
  // must not queue behind the thousands of CTAs of a bulk update, or the overlap is lost (with equal priorities the phase
  // took the SUM of the two, 89 ms at N = 16,384).
  if (!g.fit_side) {
    int least = 0, greatest = 0;
    CDX_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CDX_CUDA(cudaStreamCreateWithPriority(&g.fit_side, cudaStreamNonBlocking, greatest));
    CDX_CUDA(cudaEventCreateWithFlags(&g.fit_ev_trsm, cudaEventDisableTiming));
    CDX_CUDA(cudaEventCreateWithFlags(&g.fit_ev_rest, cudaEventDisableTiming));
  }
  cudaStream_t hi = g.fit_side;
  CDX_CUDA(cudaEventRecord(g.fit_ev_trsm, st));
  CDX_CUDA(cudaStreamWaitEvent(hi, g.fit_ev_trsm, 0));              // the chain starts after everything queued so far
  // outer panel width: 512 while more than 8,192 columns are left (K = 512 halves the trailing update's per-tile prologue /
  // epilogue share: Cholesky phase 68.8 -> 65.8 ms at N = 16,384), 256 below (N = 8,192: 13.2 ms against 13.9 - the longer
  // in-panel chain is exposed there, and the same holds for the tail of a larger matrix)
  auto panel_width = [&](int c) { const int left = Np - c, pw = left > 8192 ? 512 : 256; return left < pw ? left : pw; };
  bool rest_pending = false;
  for (int c0 = 0, w = 0; c0 < Np; c0 += w) {
    w = panel_width(c0);                                             // a multiple of 64
    for (int j0 = c0; j0 < c0 + w; j0 += 64) {
      double* Ajj = g.L + (int64_t)j0 * (ld + 1);
      potrf_diag_kernel<<<1, 256, potrf_smem_bytes, hi>>>(Ajj, ld, g.Dinv + (int64_t)(j0 / 64) * 4096, g.U + (int64_t)j0 * (ld + 1), g.flag, j0);
      CDX_CHECK_LAUNCH(ctx);
      const int m = Np - (j0 + 64);
      if (m <= 0) continue;
      double* A21 = Ajj + 64 * ld;
      // L21 = A21 * inv(Ljj)^T   (in place: each CTA reads its own 64x64 row block completely before writing it)
      int rc = cdx_dgemm(ctx, false, false, m, 64, 64, 1.0, A21, ld, g.Dinv + (int64_t)(j0 / 64) * 4096, 64, 0.0, A21, ld, false,
                         0, 1, 0, 0, 0, hi);
      if (rc) return rc;
      const int wr = c0 + w - (j0 + 64);                             // columns of this panel still to the right
      if (wr > 0) {                                                  // A[j0+64.., j0+64 .. c0+w) -= L21 L21[0:wr]^T
        rc = cdx_dgemm(ctx, false, false, m, wr, 64, -1.0, A21, ld, A21, ld, 1.0, A21 + 64, ld, false, 0, 1, 0, 0, 0, hi);
        if (rc) return rc;
      }
    }
    const int m2 = Np - (c0 + w);                                    // trailing rows / columns
    if (m2 <= 0) break;
    double* L21 = g.L + (int64_t)(c0 + w) * ld + c0;                 // [m2, w]
    double* A22 = g.L + (int64_t)(c0 + w) * (ld + 1);
    CDX_CUDA(cudaEventRecord(g.fit_ev_trsm, hi));                    // this panel's columns are final
    if (rest_pending) CDX_CUDA(cudaStreamWaitEvent(hi, g.fit_ev_rest, 0));   // the previous bulk update touched the next panel's columns
    // next panel's columns first (chain stream), K = w
    const int wn = panel_width(c0 + w);                              // the next panel's width
    int rc = cdx_dgemm(ctx, false, false, m2, wn, w, -1.0, L21, ld, L21, ld, 1.0, A22, ld, false, 0, 1, 0, 0, 0, hi);
    if (rc) return rc;
    // the rest (rows / columns >= c0 + w + 256, lower tiles only) on the caller's stream
    const int m3 = m2 - wn;
    CDX_CUDA(cudaStreamWaitEvent(st, g.fit_ev_trsm, 0));
    if (m3 > 0) {
      double* L31 = L21 + (int64_t)wn * ld;
      rc = cdx_dgemm(ctx, false, false, m3, m3, w, -1.0, L31, ld, L31, ld, 1.0, A22 + (int64_t)wn * (ld + 1), ld, true, 0, 1, 0, 0, 0, st);
      if (rc) return rc;
    }
    CDX_CUDA(cudaEventRecord(g.fit_ev_rest, st));
    rest_pending = true;
  }
  CDX_CUDA(cudaEventRecord(g.fit_ev_trsm, hi));                      // the caller's stream continues after the chain's last step
  CDX_CUDA(cudaStreamWaitEvent(st, g.fit_ev_trsm, 0));
  dim3 g2((Np + 255) / 256, Np);
  zero_upper_kernel<<<g2, 256, 0, st>>>(g.L, Np);
  CDX_CHECK_LAUNCH(ctx);
  fit_mark(g, 2, st);
  // U = L^-1 by doubling: inv([L11 0; L21 L22]) = [U11 0; -U22 L21 U11, U22]; P is scratch for T = L21 U11
  for (int s = 64; s < Np; s *= 2) {
    int npairs = (Np + 2 * s - 1) / (2 * s);
    int nfull = 0;
    for (int p = 0; p < npairs; p++) if (2 * p * s + 2 * s <= Np) nfull++;
    auto level = [&](int p0, int batch, int r) -> int {
      int64_t off11 = (int64_t)(2 * p0 * s) * (ld + 1);
      int64_t off21 = off11 + (int64_t)s * ld;
      int64_t off22 = off11 + (int64_t)s * (ld + 1);
      int64_t stride = (int64_t)(2 * s) * (ld + 1);
      int rc = cdx_dgemm(ctx, false, true, r, s, s, 1.0, g.L + off21, ld, g.U + off11, ld, 0.0, g.P + off21, ld, false, 1,
                         batch, stride, stride, stride, st);
      if (rc) return rc;
      return cdx_dgemm(ctx, false, true, r, s, r, -1.0, g.U + off22, ld, g.P + off21, ld, 0.0, g.U + off21, ld, false, 2,
                       batch, stride, stride, stride, st);
    };
    if (nfull > 0) { int rc = level(0, nfull, s); if (rc) return rc; }
    if (nfull < npairs) {
      int r = Np - (2 * nfull * s + s);
      if (r > 0) { int rc = level(nfull, 1, r); if (rc) return rc; }
    }
  }
  fit_mark(g, 3, st);
  // P = U^T U (lower tiles, k >= m0), mirrored
  int rc = cdx_dgemm(ctx, true, true, Np, Np, Np, 1.0, g.U, ld, g.U, ld, 0.0, g.P, ld, true, 3, 1, 0, 0, 0, st);
  if (rc) return rc;
  mirror_lower_kernel<<<g2, 256, 0, st>>>(g.P, Np);
  CDX_CHECK_LAUNCH(ctx);
  if (Np <= 448) {
    pack_slab_kernel<<<g2, 256, 0, st>>>(g.P, Np, g.Pslab);
    CDX_CHECK_LAUNCH(ctx);
  }
  fit_mark(g, 4, st);
  // alpha = U^T (U y1)
  trmv_lower_kernel<<<(Np + 7) / 8, 256, 0, st>>>(g.U, Np, g.y1, g.tmp);
  CDX_CHECK_LAUNCH(ctx);
  trmv_lower_t_kernel<<<(Np + 31) / 32, 256, 0, st>>>(g.U, Np, g.tmp, g.tmp + Np);
  CDX_CHECK_LAUNCH(ctx);
  pack_xa_kernel<<<(Np + 255) / 256, 256, 0, st>>>(X1, g.tmp + Np, N, Np, g.Xa);
  CDX_CHECK_LAUNCH(ctx);
  cdx_kparams kp = cdx_make_kparams(g);
  g.prior = g.kernel == CDX_KERNEL_TPS ? kp.R3 : (g.kernel == CDX_KERNEL_RBF ? 1.0 : 0.3 + 0.7 * kp.R3);
  g.ready = true;
  fit_mark(g, 5, st);
  return CDX_OK;
}

}  // namespace

extern "C" int cdx_gpis_fit_profile(cdx_ctx* ctx, int enable) {
  if (!ctx) { cdx_set_error("cdx_gpis_fit_profile: NULL context"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_gpis_state& g = ctx->g;
  if (enable)
    for (int i = 0; i < 6; i++)
      if (!g.fit_ev[i]) CDX_CUDA(cudaEventCreate(&g.fit_ev[i]));
  g.fit_prof = enable != 0;
  return CDX_OK;
}

extern "C" int cdx_gpis_fit_profile_read(cdx_ctx* ctx, double* ms_out5, void* stream) {
  if (!ctx || !ms_out5 || !ctx->g.fit_ev[0]) { cdx_set_error("cdx_gpis_fit_profile_read: enable profiling and fit first"); return CDX_ERR_STATE; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  CDX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  for (int i = 0; i < 5; i++) {
    float ms = 0.f;
    CDX_CUDA(cudaEventElapsedTime(&ms, ctx->g.fit_ev[i], ctx->g.fit_ev[i + 1]));
    ms_out5[i] = (double)ms;
  }
  return CDX_OK;
}

extern "C" int cdx_create(int device, cdx_ctx** out) {
  if (!out) { cdx_set_error("cdx_create: null out"); return CDX_ERR_ARG; }
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
    cudaGetLastError();
    cdx_set_error("cdx_create: no usable CUDA device %d (count %d); cdx has no CPU fallback", device, n);
    return CDX_ERR_CUDA;
  }
  CDX_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  CDX_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    cdx_set_error("cdx_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    return CDX_ERR_CUDA;
  }
  cdx_ctx* c = new cdx_ctx();
  c->device = device;
  c->sms = prop.multiProcessorCount;
  *out = c;
  return CDX_OK;
}

void cdx_pregrasp_free(cdx_ctx* ctx);

extern "C" int cdx_destroy(cdx_ctx* ctx) {
  if (!ctx) return CDX_OK;
  cudaSetDevice(ctx->device);
  cdx_gpis_state& g = ctx->g;
  cudaFree(g.Xa); cudaFree(g.y1); cudaFree(g.E11); cudaFree(g.L); cudaFree(g.U); cudaFree(g.P);
  cudaFree(g.Dinv); cudaFree(g.tmp); cudaFree(g.dR2); cudaFree(g.flag); cudaFree(g.Pslab);
  cudaFree(ctx->hx); cudaFree(ctx->hout);
  for (int i = 0; i < 6; i++) if (g.fit_ev[i]) cudaEventDestroy(g.fit_ev[i]);
  if (g.fit_side) { cudaStreamDestroy(g.fit_side); cudaEventDestroy(g.fit_ev_trsm); cudaEventDestroy(g.fit_ev_rest); }
  cdx_i8x_free(ctx);
  cdx_pregrasp_free(ctx);
  cdx_sdf_free(ctx);
  delete ctx;
  return CDX_OK;
}

extern "C" int64_t cdx_launch_count(cdx_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int cdx_gpis_fit(cdx_ctx* ctx, const double* X1, const double* y1, const double* noise, double noise_scalar,
                            int64_t N, double bias, int kernel, double sigma, void* stream) {
  CDX_NVTX("cdx_gpis_fit");
  if (!ctx || !X1 || !y1 || N <= 0 || kernel < 0 || kernel > 2) { cdx_set_error("cdx_gpis_fit: bad argument"); return CDX_ERR_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  CDX_CUDA(cudaSetDevice(ctx->device));
  int rc = gpis_reserve(ctx, N);
  if (rc) return rc;
  cdx_gpis_state& g = ctx->g;
  g.kernel = kernel; g.sigma = sigma; g.bias = bias; g.R = 0.0;
  const int Np = (int)g.Np;
  fit_mark(g, 0, st);
  if (kernel != CDX_KERNEL_RBF) {  // gpis.py:35-36
    CDX_CUDA(cudaMemsetAsync(g.dR2, 0, sizeof(double), st));
    const int jsplit = (int)((N + 511) / 512) < 64 ? (int)((N + 511) / 512) : 64, jlen = ((int)N + jsplit - 1) / jsplit;
    pairmax_kernel<<<dim3(((int)N + 127) / 128, jsplit), 128, 0, st>>>(X1, (int)N, jlen, (unsigned long long*)g.dR2);
    CDX_CHECK_LAUNCH(ctx);
    double r2 = 0;
    CDX_CUDA(cudaMemcpyAsync(&r2, g.dR2, sizeof(double), cudaMemcpyDeviceToHost, st));
    CDX_CUDA(cudaStreamSynchronize(st));
    g.R = sqrt(r2);
  }
  sub_bias_kernel<<<(Np + 255) / 256, 256, 0, st>>>(y1, (int)N, Np, bias, g.y1);  // gpis.py:38
  CDX_CHECK_LAUNCH(ctx);
  cdx_kparams kp = cdx_make_kparams(g);
  dim3 g2((Np + 255) / 256, Np);
  if (kernel == CDX_KERNEL_TPS) assemble_kernel<CDX_KERNEL_TPS><<<g2, 256, 0, st>>>(X1, noise, noise_scalar, (int)N, Np, kp, g.E11, g.L);
  else if (kernel == CDX_KERNEL_RBF) assemble_kernel<CDX_KERNEL_RBF><<<g2, 256, 0, st>>>(X1, noise, noise_scalar, (int)N, Np, kp, g.E11, g.L);
  else assemble_kernel<CDX_KERNEL_JOINT><<<g2, 256, 0, st>>>(X1, noise, noise_scalar, (int)N, Np, kp, g.E11, g.L);
  CDX_CHECK_LAUNCH(ctx);
  return gpis_factorise(ctx, X1, st, true);
}

extern "C" int cdx_gpis_set_state(cdx_ctx* ctx, const double* X1, const double* y1, const double* E11, double R,
                                  double bias, int64_t N, int kernel, double sigma, void* stream) {
  CDX_NVTX("cdx_gpis_set_state");
  if (!ctx || !X1 || !y1 || !E11 || N <= 0 || kernel < 0 || kernel > 2) { cdx_set_error("cdx_gpis_set_state: bad argument"); return CDX_ERR_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  CDX_CUDA(cudaSetDevice(ctx->device));
  int rc = gpis_reserve(ctx, N);
  if (rc) return rc;
  cdx_gpis_state& g = ctx->g;
  g.kernel = kernel; g.sigma = sigma; g.bias = bias; g.R = R;
  const int Np = (int)g.Np;
  fit_mark(g, 0, st);
  sub_bias_kernel<<<(Np + 255) / 256, 256, 0, st>>>(y1, (int)N, Np, 0.0, g.y1);
  CDX_CHECK_LAUNCH(ctx);
  dim3 g2((Np + 255) / 256, Np);
  pad_copy_kernel<<<g2, 256, 0, st>>>(E11, (int)N, Np, g.E11);
  CDX_CHECK_LAUNCH(ctx);
  return gpis_factorise(ctx, X1, st, false);
}

extern "C" int cdx_gpis_info(cdx_ctx* ctx, int64_t* N, double* R, double* bias, double* prior, int* not_spd_col, void* stream) {
  if (!ctx || !ctx->g.ready) { cdx_set_error("cdx_gpis_info: no fitted state"); return CDX_ERR_STATE; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  int f = 0;
  CDX_CUDA(cudaMemcpyAsync(&f, ctx->g.flag, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CDX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (N) *N = ctx->g.N;
  if (R) *R = ctx->g.R;
  if (bias) *bias = ctx->g.bias;
  if (prior) *prior = ctx->g.prior;
  if (not_spd_col) *not_spd_col = f - 1;
  if (f) { cdx_set_error("E11 is not positive definite: Cholesky pivot at column %d <= 0", f - 1); return CDX_ERR_NOT_SPD; }
  return CDX_OK;
}

extern "C" int cdx_gpis_get(cdx_ctx* ctx, int which, double* out, void* stream) {
  CDX_NVTX("cdx_gpis_get");
  if (!ctx || !ctx->g.ready || !out) { cdx_set_error("cdx_gpis_get: no fitted state / null out"); return CDX_ERR_STATE; }
  cudaStream_t st = (cudaStream_t)stream;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cdx_gpis_state& g = ctx->g;
  const int N = (int)g.N, Np = (int)g.Np;
  dim3 g2((N + 255) / 256, N);
  const double* src = nullptr;
  switch (which) {
    case CDX_MAT_E11: src = g.E11; break;
    case CDX_MAT_CHOL: src = g.L; break;
    case CDX_MAT_CHOL_INV: src = g.U; break;
    case CDX_MAT_E11_INV: src = g.P; break;
    case CDX_VEC_ALPHA: extract_alpha_kernel<<<(N + 255) / 256, 256, 0, st>>>(g.Xa, N, out); CDX_CHECK_LAUNCH(ctx); return CDX_OK;
    case CDX_VEC_Y1: CDX_CUDA(cudaMemcpyAsync(out, g.y1, sizeof(double) * N, cudaMemcpyDeviceToDevice, st)); return CDX_OK;
    case CDX_MAT_X1: extract_x_kernel<<<(N + 255) / 256, 256, 0, st>>>(g.Xa, N, out); CDX_CHECK_LAUNCH(ctx); return CDX_OK;
    default: cdx_set_error("cdx_gpis_get: unknown selector %d", which); return CDX_ERR_ARG;
  }
  unpad_copy_kernel<<<g2, 256, 0, st>>>(src, N, Np, out);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}
