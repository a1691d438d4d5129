// Generic FP64 GEMM on the DMMA tensor path (mma.sync m8n8k4 f64) used by the one-time fit:
// Cholesky trailing updates (SYRK), panel solves, the block-recursive triangular inverse and E11^-1 = U^T U.
// Two kernels: dgemm128_kernel (128x128x32 CTA tiles, 16 warps, 3-stage cp.async ring on full / empty mbarriers) for the large
// products, dgemm_kernel (64x64x16 CTA tile, 4 warps (2x2), each warp a 32x32 tile = 4x4 DMMA blocks) for the thin / in-place
// panel operations; operands staged in shared memory as [row][k] with a padded row stride (conflict-free for the DMMA fragment
// pattern row=T/4, k=T%4).
#include "cdx_internal.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, LDS = 20;

template <bool TA, bool TB>
__global__ void __launch_bounds__(128) dgemm_kernel(int M, int N, int K, double alpha, const double* __restrict__ A,
                                                    int64_t lda, const double* __restrict__ B, int64_t ldb, double beta,
                                                    double* __restrict__ C, int64_t ldc, int lower_only, int ktri,
                                                    int64_t sA, int64_t sB, int64_t sC) {
  const int tm = blockIdx.y, tn = blockIdx.x;
  if (lower_only && tn > tm) return;
  A += (int64_t)blockIdx.z * sA;
  B += (int64_t)blockIdx.z * sB;
  C += (int64_t)blockIdx.z * sC;
  __shared__ double As[BM * LDS];
  __shared__ double Bs[BN * LDS];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int wm = warp >> 1, wn = warp & 1;
  const int m0 = tm * BM, n0 = tn * BN;

  double ra[8], rb[8];
  auto load_tiles = [&](int k0) {
    if (!TA) {  // A[M,K]: thread -> row t/2, 8 consecutive k
      int r = t >> 1, kk = (t & 1) * 8;
      const double* p = A + (int64_t)(m0 + r) * lda + k0 + kk;
      bool rok = (m0 + r) < M;
#pragma unroll
      for (int i = 0; i < 8; i++) ra[i] = (rok && (k0 + kk + i) < K) ? p[i] : 0.0;
    } else {  // A[K,M]: thread -> k = t/8, 8 consecutive m
      int kk = t >> 3, mm = (t & 7) * 8;
      const double* p = A + (int64_t)(k0 + kk) * lda + m0 + mm;
      bool kok = (k0 + kk) < K;
#pragma unroll
      for (int i = 0; i < 8; i++) ra[i] = (kok && (m0 + mm + i) < M) ? p[i] : 0.0;
    }
    if (!TB) {  // B[N,K]
      int r = t >> 1, kk = (t & 1) * 8;
      const double* p = B + (int64_t)(n0 + r) * ldb + k0 + kk;
      bool rok = (n0 + r) < N;
#pragma unroll
      for (int i = 0; i < 8; i++) rb[i] = (rok && (k0 + kk + i) < K) ? p[i] : 0.0;
    } else {  // B[K,N]
      int kk = t >> 3, nn = (t & 7) * 8;
      const double* p = B + (int64_t)(k0 + kk) * ldb + n0 + nn;
      bool kok = (k0 + kk) < K;
#pragma unroll
      for (int i = 0; i < 8; i++) rb[i] = (kok && (n0 + nn + i) < N) ? p[i] : 0.0;
    }
  };
  auto store_tiles = [&]() {
    if (!TA) {
      int r = t >> 1, kk = (t & 1) * 8;
#pragma unroll
      for (int i = 0; i < 8; i++) As[r * LDS + kk + i] = ra[i];
    } else {
      int kk = t >> 3, mm = (t & 7) * 8;
#pragma unroll
      for (int i = 0; i < 8; i++) As[(mm + i) * LDS + kk] = ra[i];
    }
    if (!TB) {
      int r = t >> 1, kk = (t & 1) * 8;
#pragma unroll
      for (int i = 0; i < 8; i++) Bs[r * LDS + kk + i] = rb[i];
    } else {
      int kk = t >> 3, nn = (t & 7) * 8;
#pragma unroll
      for (int i = 0; i < 8; i++) Bs[(nn + i) * LDS + kk] = rb[i];
    }
  };

  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

  const int fr = lane >> 2, fk = lane & 3;
  // triangular operands: skip k ranges known to be zero (ktri 1: k >= n0, 2: k < m0+BM, 3: k >= m0)
  int kbeg = 0, kend = K;
  if (ktri == 1) kbeg = n0;
  if (ktri == 3) kbeg = m0;
  if (ktri == 2) kend = min(K, m0 + BM);
  load_tiles(kbeg);
  const int Kfull = K;
  K = kend;
  (void)Kfull;
  for (int k0 = kbeg; k0 < K; k0 += BK) {
    __syncthreads();
    store_tiles();
    __syncthreads();
    if (k0 + BK < K) load_tiles(k0 + BK);
#pragma unroll
    for (int ks = 0; ks < BK; ks += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = As[(wm * 32 + i * 8 + fr) * LDS + ks + fk];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = Bs[(wn * 32 + j * 8 + fr) * LDS + ks + fk];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) cdx_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int r = m0 + wm * 32 + i * 8 + fr;
      int c = n0 + wn * 32 + j * 8 + 2 * fk;
      if (r < M) {
#pragma unroll
        for (int e = 0; e < 2; e++) {
          if (c + e < N) {
            double* p = C + (int64_t)r * ldc + c + e;
            double v = alpha * acc[i][j][e];
            if (beta != 0.0) v += beta * (*p);
            *p = v;
          }
        }
      }
    }
}

// ------------------------------------------------------------------------------------------ 128 x 128 tiles, mbarrier pipeline
// The large products of the fit (trailing updates of the Cholesky with K = 256 / 512, the recursive-doubling levels of L^-1,
// E11^-1 = U^T U): 128x128x32 CTA tiles, 16 warps (4 x 4, each a 32x32 tile = 4x4 DMMA blocks) - a DMMA holds its warp for
// several cycles, so it takes four warps per scheduler to keep the pipe fed.  Operand tiles keep the layout they have in
// memory: k-contiguous sources as [row][36] (conflict-free for the fragment pattern row = T/4, k = T%4), row-contiguous
// sources (ta / tb) as [k][132].  Three ring stages and NO CTA-wide barrier in the k loop: every thread's 16-byte cp.async
// copies of a chunk arrive on the stage's `full` mbarrier (cp.async.mbarrier.arrive.noinc; rows outside the matrix are
// zero-filled by cp.async with a source size of 0, so every write of a stage is tracked), a warp releases a stage on its
// `empty` mbarrier when its DMMAs have consumed the last fragments, and a chunk is prefetched ONE iteration ahead into the
// stage that chunk c - 2 used - so a warp may run a whole chunk ahead of the slowest one instead of meeting it at every chunk.
// History (fit at N = 16,384): 64x64 single-buffered kernel 0.40-0.50 of the DMMA peak; 128x128x16, 4-stage cp.async ring with
// a __syncthreads per chunk 0.76-0.77 in the inverse phases; 32-deep chunks (2 stages, half the barriers) 0.79-0.80, ncu: 8 % of
// the stall samples on the barrier, DMMA sub-pipe 83 %; this version 0.87-0.90 (L^-1 49.8 -> 45.1 ms, E11^-1 49.2 -> 44.1 ms,
// Cholesky 61.9 -> 58.0 ms).  221 KB of shared memory, 120 registers x 512 threads: nothing else fits on the SM beside it - the
// chain kernels of the Cholesky get their SMs from the stream priority as tiles retire, which is what happened before as well
// (61 k of the 64 k registers were taken already).
constexpr int GM = 128, GN = 128, GK = 32, GST = 3, LDT = 132, LDG = GK + 4;
constexpr int G_TILE = GM * LDG > GK * LDT ? GM * LDG : GK * LDT;   // doubles per operand tile and stage


__device__ __forceinline__ void g_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void g_mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void g_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void g_cp_async_arrive(uint64_t* bar) {   // arrives when this thread's earlier cp.async copies have landed
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void g_cp_async16_zfill(void* smem, const void* gmem, bool valid) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  int bytes = valid ? 16 : 0;                                        // 0: nothing is read, the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}

template <bool TA, bool TB>
__global__ void __launch_bounds__(512, 1) dgemm128_kernel(int M, int N, int K, double alpha, const double* __restrict__ A,
                                                            int64_t lda, const double* __restrict__ B, int64_t ldb, double beta,
                                                            double* __restrict__ C, int64_t ldc, int lower_only, int ktri,
                                                            int64_t sA, int64_t sB, int64_t sC) {
  const int tm = blockIdx.y, tn = blockIdx.x;
  if (lower_only && tn > tm) return;
  A += (int64_t)blockIdx.z * sA;
  B += (int64_t)blockIdx.z * sB;
  C += (int64_t)blockIdx.z * sC;
  extern __shared__ __align__(16) double gsm[];
  double* As = gsm;                        // [GST][G_TILE]
  double* Bs = gsm + GST * G_TILE;        // [GST][G_TILE]
  uint64_t* full = reinterpret_cast<uint64_t*>(gsm + 2 * GST * G_TILE);
  uint64_t* empty = full + GST;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int wm = warp >> 2, wn = warp & 3;
  const int fr = lane >> 2, fk = lane & 3;
  const int m0 = tm * GM, n0 = tn * GN;
  int kbeg = 0, kend = K;
  if (ktri == 1) kbeg = n0;
  if (ktri == 3) kbeg = m0;
  if (ktri == 2) kend = min(K, m0 + GM);
  const int nch = (kend - kbeg + GK - 1) / GK;
  if (t == 0) {
    for (int s = 0; s < GST; s++) { g_mbar_init(full + s, 512); g_mbar_init(empty + s, 16); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  auto load_tile = [&](const double* __restrict__ P, int64_t ld, bool trans, int r0, int R, double* dst, int k0) {
    if (!trans) {
      for (int p = t; p < GM * (GK / 2); p += 512) {
        const int row = p / (GK / 2), seg = p % (GK / 2);
        const bool ok = r0 + row < R && k0 + seg * 2 < kend;
        g_cp_async16_zfill(dst + row * LDG + seg * 2, ok ? P + (int64_t)(r0 + row) * ld + k0 + seg * 2 : P, ok);
      }
    } else {
      for (int p = t; p < GK * 64; p += 512) {
        const int kk = p >> 6, seg = p & 63;
        const bool ok = k0 + kk < kend && r0 + seg * 2 < R;
        g_cp_async16_zfill(dst + kk * LDT + seg * 2, ok ? P + (int64_t)(k0 + kk) * ld + r0 + seg * 2 : P, ok);
      }
    }
  };
  auto issue = [&](int c) {                // c < nch
    const int buf = c % GST, k0 = kbeg + c * GK;
    load_tile(A, lda, TA, m0, M, As + buf * G_TILE, k0);
    load_tile(B, ldb, TB, n0, N, Bs + buf * G_TILE, k0);
    g_cp_async_arrive(full + buf);
  };

  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

  if (nch > 0) issue(0);
  for (int c = 0; c < nch; c++) {
    if (c + 1 < nch) {
      const int nb = (c + 1) % GST, use = (c + 1) / GST;             // the stage's use number `use`; chunk c - 2 was use - 1
      if (use > 0) g_mbar_wait(empty + nb, (unsigned)((use - 1) & 1));
      issue(c + 1);
    }
    g_mbar_wait(full + c % GST, (unsigned)((c / GST) & 1));
    const double* a_s = As + (c % GST) * G_TILE;
    const double* b_s = Bs + (c % GST) * G_TILE;
#pragma unroll
    for (int ks = 0; ks < GK; ks += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++)
        a[i] = TA ? a_s[(ks + fk) * LDT + wm * 32 + i * 8 + fr] : a_s[(wm * 32 + i * 8 + fr) * LDG + ks + fk];
#pragma unroll
      for (int j = 0; j < 4; j++)
        b[j] = TB ? b_s[(ks + fk) * LDT + wn * 32 + j * 8 + fr] : b_s[(wn * 32 + j * 8 + fr) * LDG + ks + fk];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) cdx_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    // the DMMAs above have consumed the stage's last fragments (the loads have completed): release it
    __syncwarp();
    if (lane == 0) g_mbar_arrive(empty + c % GST);
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int r = m0 + wm * 32 + i * 8 + fr;
      const int c = n0 + wn * 32 + j * 8 + 2 * fk;
      if (r < M && c < N) {
        double2* p = reinterpret_cast<double2*>(C + (int64_t)r * ldc + c);
        double2 v = make_double2(alpha * acc[i][j][0], alpha * acc[i][j][1]);
        if (beta != 0.0) { const double2 o = *p; v.x += beta * o.x; v.y += beta * o.y; }
        *p = v;
      }
    }
}

}  // namespace

int cdx_dgemm(cdx_ctx* ctx, bool ta, bool tb, int M, int N, int K, double alpha, const double* A, int64_t lda,
              const double* B, int64_t ldb, double beta, double* C, int64_t ldc, bool lower_only, int ktri, int batch,
              int64_t sA, int64_t sB, int64_t sC, cudaStream_t st) {
  if (M <= 0 || N <= 0 || batch <= 0) return CDX_OK;
  // large, aligned products go to the 128x128 mbarrier-pipelined kernel; thin / in-place panel operations keep the 64x64 one
  if (M >= 128 && N >= 128 && (M % 2) == 0 && (N % 2) == 0 && (K % 2) == 0 && (lda % 2) == 0 && (ldb % 2) == 0 && (ldc % 2) == 0 &&
      A != C && B != C) {
    const size_t smem = sizeof(double) * 2 * GST * G_TILE + 2 * GST * sizeof(uint64_t);
    dim3 grid128((N + GN - 1) / GN, (M + GM - 1) / GM, batch);
    const int lo128 = lower_only ? 1 : 0;
#define CDX_G128(TA_, TB_)                                                                                                 \
    do {                                                                                                                   \
      if (int rc = cdx_smem_attr(ctx, dgemm128_kernel<TA_, TB_>, smem)) return rc;                                       \
      dgemm128_kernel<TA_, TB_><<<grid128, 512, smem, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lo128, ktri, sA, sB, sC); \
    } while (0)
    if (!ta && !tb) CDX_G128(false, false); else if (!ta && tb) CDX_G128(false, true);
    else if (ta && !tb) CDX_G128(true, false); else CDX_G128(true, true);
#undef CDX_G128
    CDX_CHECK_LAUNCH(ctx);
    return CDX_OK;
  }
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, batch);
  int lo = lower_only ? 1 : 0;
  if (!ta && !tb)
    dgemm_kernel<false, false><<<grid, 128, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lo, ktri, sA, sB, sC);
  else if (!ta && tb)
    dgemm_kernel<false, true><<<grid, 128, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lo, ktri, sA, sB, sC);
  else if (ta && !tb)
    dgemm_kernel<true, false><<<grid, 128, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lo, ktri, sA, sB, sC);
  else
    dgemm_kernel<true, true><<<grid, 128, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, lo, ktri, sA, sB, sC);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}
