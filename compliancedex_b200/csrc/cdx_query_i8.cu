// GPIS posterior std on the 5th-generation tensor cores (tcgen05, INT8 kind, INT32 accumulators in TMEM).
//
// sigma(x)^2 = k(x,x) - |v|^2, v = L^-1 k*(x)  (gpis.py:51-59; the reference solves an LU per call).  v = U k* with the
// precomputed lower-triangular U = L^-1 is a dense triangular contraction, but tcgen05 has no FP64 kind.  This kernel
// evaluates it as an error-free INTEGER product (Ozaki-style slicing):
//   k*(x)_k  = S_q 256^-S  sum_i a_i 256^(S-1-i)      a_i  in [0,255]      (S unsigned byte slices of a fixed-point value)
//   U[n,k]   = 2^e_n 256^-S sum_j b_j 256^(S-1-j)      b_j  in [-128,127]   (balanced signed digits, per-row exponent)
//   v_n      = S_q 2^e_n 256^-(S+1) sum_{d<S} 256^(S-1-d) sum_{i+j=d} sum_k a_i[k] b_j[n,k]
// every inner sum is an exact INT32 tcgen05.mma accumulation (no rounding at all); the only error is the truncation of
// the two operands to 8S-2 bits and of the slice pairs to i+j < S, 2^-8S-class relative to |row|.|column| - measured on
// B200 against the FP64 DMMA path (tests/test_gpis_i8_gpu.py): S = 5 gives |d var| < 1e-8 prior, the gate is 1e-6 prior.
//
// Roles per CTA (128 queries x all training rows, one CTA per SM; TMEM: S x 64 accumulator columns + 2 stages of k*):
//   warps 0-15  generate k* for (query row, 16 training points) in FP64 (9 DFMA-class ops + 1 MUFU.RSQ per pair: expanded
//               squared distance, float rsqrt seed built by integer bit moves instead of F2F conversions, one Newton step,
//               cubic, fixed-point conversion by the 1.5*2^52 trick), split the value into bytes with PRMT and write the S
//               slices straight into tensor memory (tcgen05.st; lane = query row, 4 int8 per column) - the A operand never
//               touches shared memory;
//   warp 17     streams the pre-tiled U slices and the stage's training points with two cp.async.bulk (1-D TMA) per stage;
//   warp 16     stays converged, one elected lane issues S(S+1)/2 tcgen05.mma (A from TMEM, B from smem, M=128, N=64,
//               K=32) per 32 training points; stages are recycled by tcgen05.commit -> mbarrier;
//   warps 18-21 drain the S diagonal accumulators from TMEM (tcgen05.ld), recombine them (integer Horner + FP64) and
//               accumulate |v|^2.
// Measured on B200 (N=4096): the kernel is bound by the FP64 generation of k* (about 3.6k cycles per 128x64 stage against
// 1.0k cycles of MMA), because k* is regenerated for each of the N/64 row tiles; see DESIGN.md for the numbers.
#include "cdx_internal.cuh"

namespace {

constexpr int I8_TQ = 128;   // queries per CTA  (UMMA M)
constexpr int I8_TN = 64;    // training rows per tile (UMMA N)
constexpr int I8_GEN_WARPS = 16, I8_WARP_MMA = 16, I8_WARP_LOAD = 17, I8_WARP_EPI = 18;
constexpr int I8_THREADS = 22 * 32;
constexpr double I8_MAGIC = 6755399441055744.0;  // 1.5 * 2^52: adding it leaves rint(x) in the low mantissa bits

#ifndef CDX_I8_BK5
#define CDX_I8_BK5 64      // training points per stage for S <= 5
#endif
#ifndef CDX_I8_NSA
#define CDX_I8_NSA 2       // A stages in tensor memory
#endif
#ifndef CDX_I8_LATE_ARRIVE
#define CDX_I8_LATE_ARRIVE 0   // 1: a stage's tcgen05.st are completed (wait::st + arrive) after the NEXT stage has been generated
#endif
#ifndef CDX_I8_EXP
#define CDX_I8_EXP 0   // timing experiments only (results are wrong): 1 = no tcgen05.mma issued, 2 = no k* arithmetic in the generators,
                       // 3 = no MMA and no tcgen05.st, 6 = MMAs issued but no tcgen05.st (A operand = stale tensor memory), 5 = generator loop only (no synchronisation at all, other warps idle), 4 = no MMA, generators do not wait for / read the TMA-staged points
#endif
#ifndef CDX_I8_NSB
#define CDX_I8_NSB 6       // shared-memory stages of (U slices, points)
#endif
template <int S> struct i8cfg {
  static constexpr int BK = S >= 6 ? 32 : CDX_I8_BK5;  // training points per stage (TMEM budget: S*64 + NSA*S*BK/4 <= 512 columns)
  static constexpr int KG = BK / 16;                   // 16-byte k groups per stage
  static constexpr int GSPLIT = 4 / KG;                // generator warps are split over this many consecutive stages
  static constexpr int ACOLS = S * BK / 4;             // TMEM columns of one A stage (4 int8 per 32-bit column)
  static constexpr int NSA = S >= 6 ? 2 : CDX_I8_NSA;  // A stages in TMEM
  static constexpr int ACC_COLS = S * I8_TN, A_BASE = ACC_COLS;
  static constexpr int TMEM_COLS = 512;
  static constexpr int B_SLICE = I8_TN * BK;           // bytes of one slice of one U tile
  static constexpr int B_BYTES = S * B_SLICE;
  static constexpr int P_BYTES = BK * 32;              // the stage's training points (-2x,-2y,-2z,|x|^2)
  static constexpr int SLOT = B_BYTES + P_BYTES;
  static constexpr int NSB = CDX_I8_NSB;               // smem stages of (U slices, points)
  static constexpr size_t SMEM = (size_t)NSB * SLOT + 256;
  static_assert(ACC_COLS + NSA * ACOLS <= 512, "TMEM budget");
};

// ------------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ unsigned s32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(s32(bar)), "r"(count));
}
__device__ __forceinline__ void mb_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mb_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mb_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
// for waits that are long by design (epilogue / loader): poll with back-off so the spin does not take issue slots
__device__ __forceinline__ void mb_wait_relaxed(unsigned bar, unsigned parity, unsigned ns = 128) {
  unsigned done = 0;
  for (;;) {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(ns);
  }
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, INT8 kind, INT32 accumulate
__device__ __forceinline__ void tc_mma_i8(unsigned tmem_d, uint64_t da, uint64_t db, unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
// same with the A operand in tensor memory (lane = row, 4 int8 per 32-bit column)
__device__ __forceinline__ void tc_mma_i8_ts(unsigned tmem_d, unsigned tmem_a, uint64_t db, unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tc_st4(unsigned taddr, unsigned a, unsigned b, unsigned c, unsigned d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float rsqrt_ftz(float x) {   // one MUFU.RSQ, no denormal fix-up (inputs are in [1e-16, 1e2])
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ bool elect_one() {   // one lane of the (converged) warp
  unsigned pred;
  asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, 0xffffffff;\nselp.u32 %0, 1, 0, px;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
// K-major, no swizzle: 8x16-byte core matrices; LBO = byte stride between the two 16-byte K columns of one MMA,
// SBO = byte stride between 8-row groups (cute/arch/mma_sm100_desc.hpp SmemDescriptor; version 1 = sm_100)
__device__ __forceinline__ uint64_t tc_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void tc_ld8(unsigned taddr, int (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// ------------------------------------------------------------------------------------------- state preparation
// per training row n: exponent e_n with |U[n,:]| 2^-e_n < 1/4 (so S balanced base-256 digits never overflow)
__global__ void __launch_bounds__(256) i8_rowscale_kernel(const double* __restrict__ U, int N, int Np, double* __restrict__ sb,
                                                          int* __restrict__ eb) {
  const int n = blockIdx.x;
  __shared__ double red[256];
  double m = 0.0;
  for (int k = threadIdx.x; k <= n; k += 256) m = fmax(m, fabs(U[(int64_t)n * Np + k]));
  red[threadIdx.x] = m;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    int e = 0;
    if (red[0] > 0.0) frexp(red[0], &e);   // max = f 2^e, f in [.5,1)
    e += 2;
    eb[n] = e;
    sb[n] = n < N ? ldexp(1.0, e) : 0.0;
  }
}

// digits of U in the smem image order of the B operand: [n tile][k chunk][slice][k group of 16][row 64][16 bytes]
template <int S>
__global__ void __launch_bounds__(256) i8_slice_kernel(const double* __restrict__ U, int N, int Np, const int* __restrict__ eb,
                                                       int8_t* __restrict__ Bsl) {
  const int k = blockIdx.x * 256 + threadIdx.x, n = blockIdx.y;
  if (k >= Np) return;
  constexpr int BK = i8cfg<S>::BK;
  const int NC = Np / BK;
  long long V = 0;
  if (n < N && k <= n) V = llrint(ldexp(U[(int64_t)n * Np + k], 8 * S - eb[n]));
  const int nt = n / I8_TN, nl = n % I8_TN, c = k / BK, kk = k % BK;
  int8_t* base = Bsl + ((size_t)(nt * NC + c) * S) * i8cfg<S>::B_SLICE + (kk / 16) * (I8_TN * 16) + nl * 16 + (kk % 16);
#pragma unroll
  for (int j = S - 1; j >= 0; j--) {
    const long long d = ((V + 128) & 255) - 128;
    V = (V - d) >> 8;
    base[(size_t)j * i8cfg<S>::B_SLICE] = (int8_t)d;
  }
}

// (-2x, -2y, -2z, |x|^2) per training point and the bounding box of the training set
__global__ void __launch_bounds__(256) i8_points_kernel(const double* __restrict__ Xa, int N, int Np, double* __restrict__ pts,
                                                        double* __restrict__ bb) {
  __shared__ double red[6][256];
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = threadIdx.x; i < Np; i += 256) {
    const double x = Xa[4 * i], y = Xa[4 * i + 1], z = Xa[4 * i + 2];
    pts[4 * i] = -2.0 * x; pts[4 * i + 1] = -2.0 * y; pts[4 * i + 2] = -2.0 * z;
    pts[4 * i + 3] = fma(x, x, fma(y, y, z * z));
    if (i < N) {
      lo[0] = fmin(lo[0], x); lo[1] = fmin(lo[1], y); lo[2] = fmin(lo[2], z);
      hi[0] = fmax(hi[0], x); hi[1] = fmax(hi[1], y); hi[2] = fmax(hi[2], z);
    }
  }
  for (int r = 0; r < 3; r++) { red[r][threadIdx.x] = lo[r]; red[3 + r][threadIdx.x] = hi[r]; }
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if ((int)threadIdx.x < o)
      for (int r = 0; r < 3; r++) {
        red[r][threadIdx.x] = fmin(red[r][threadIdx.x], red[r][threadIdx.x + o]);
        red[3 + r][threadIdx.x] = fmax(red[3 + r][threadIdx.x], red[3 + r][threadIdx.x + o]);
      }
    __syncthreads();
  }
  if (threadIdx.x < 6) bb[threadIdx.x] = red[threadIdx.x][0];
}

// per-query fixed-point scale: k* <= kmax = max(R^3, k(dmax)), dmax = distance to the farthest corner of the training
// box;  S_q = 2^(e+1) > 2 kmax.  Returns sc = 256^S / S_q (a power of two).
template <int S>
__device__ __forceinline__ double i8_query_scale(double x, double y, double z, const double* __restrict__ bb, double R,
                                                 double R3) {
  const double ax = fmax(fabs(x - bb[0]), fabs(x - bb[3])), ay = fmax(fabs(y - bb[1]), fabs(y - bb[4])),
               az = fmax(fabs(z - bb[2]), fabs(z - bb[5]));
  const double d2 = fma(ax, ax, fma(ay, ay, az * az)), dm = sqrt(d2);
  const double kmax = fmax(R3, fma(d2, fma(2.0, dm, -3.0 * R), R3));
  int e;
  frexp(kmax, &e);
  return ldexp(1.0, 8 * S - e - 1);
}

// ------------------------------------------------------------------------------------------- the query kernel
template <int S>
__global__ void __launch_bounds__(I8_THREADS, 1)
gpis_std_i8_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ pts, const int8_t* __restrict__ Bsl,
                   const double* __restrict__ sb, const double* __restrict__ bb, int Np, double R, double R3, double prior,
                   double* __restrict__ stdv) {
  using C = i8cfg<S>;
  constexpr int NSA = C::NSA, NSB = C::NSB, BK = C::BK, KG = C::KG;
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)NSB * C::SLOT);
  // mbarriers, addressed by their 32-bit shared-window offsets (8 bytes each)
  const unsigned smem0 = s32(smem), bar0 = s32(bars);
  const unsigned a_full = bar0;                         // [NSA] generator warps arrive (k* slices are in tensor memory)
  const unsigned a_empty = bar0 + 8 * NSA;              // [NSA] tcgen05.commit
  const unsigned b_full = bar0 + 8 * 2 * NSA;           // [NSB] bulk copies (tx bytes)
  const unsigned b_empty = b_full + 8 * NSB;            // [NSB] tcgen05.commit
  const unsigned acc_full = b_empty + 8 * NSB;
  const unsigned acc_empty = acc_full + 8;
  unsigned* tmem_slot = reinterpret_cast<unsigned*>(bars + 2 * NSA + 2 * NSB + 2);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int64_t q0 = (int64_t)blockIdx.x * I8_TQ;
  const int NT = Np / I8_TN, NC = Np / BK;
  constexpr int CPT = I8_TN / BK;           // k chunks per n tile (triangular bound: chunks c < (nt + 1) CPT)

  if (t == 0) {
    for (int i = 0; i < NSA; i++) { mb_init(bars + i, 4 * KG); mb_init(bars + NSA + i, 1); }
    for (int i = 0; i < 2 * NSB; i++) mb_init(bars + 2 * NSA + i, 1);
    mb_init(bars + 2 * NSA + 2 * NSB, 1);
    mb_init(bars + 2 * NSA + 2 * NSB + 1, 4);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == I8_WARP_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(s32(tmem_slot)), "r"(C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = *tmem_slot;

  if (warp < I8_GEN_WARPS) {
    // ===================================================== generators: row = t % 128, 16 training points per owned stage
    const int row = t & 127, kq = t >> 7, kg = kq % KG, sub = kq / KG;
    const int64_t q = q0 + row;
    const bool ok = q < M;
    const double xq = ok ? Xq[q * 3] : 0.0, yq = ok ? Xq[q * 3 + 1] : 0.0, zq = ok ? Xq[q * 3 + 2] : 0.0;
    const double qq = fma(xq, xq, fma(yq, yq, zq * zq)) + 1e-16;   // keeps the expanded squared distance positive
    const double sc = i8_query_scale<S>(xq, yq, zq, bb, R, R3);
    const double c2 = 2.0 * sc, c3 = -3.0 * R * sc, c0m = R3 * sc + I8_MAGIC;
    const unsigned trow = tmem + ((unsigned)((warp & 3) * 32) << 16) + C::A_BASE + kg * 4;
    int g = 0;
#if CDX_I8_LATE_ARRIVE
    int pending = -1;
#endif
    for (int nt = 0; nt < NT; nt++) {
      const int nch = (nt + 1) * CPT;
      for (int c = 0; c < nch; c++, g++) {
        if (C::GSPLIT > 1 && (g % C::GSPLIT) != sub) continue;
        const int sa = g % NSA, sbk = g % NSB;
        if (CDX_I8_EXP != 4 && CDX_I8_EXP != 5) mb_wait(b_full + 8 * sbk, (unsigned)(g / NSB) & 1u);     // the stage's training points arrive with the U slices
        const double2* P2 = reinterpret_cast<const double2*>(smem + (size_t)(CDX_I8_EXP == 4 ? 0 : sbk) * C::SLOT + C::B_BYTES) + kg * 32;
        unsigned ow[S][4];
#pragma unroll
        for (int gq = 0; gq < 4; gq++) {
          unsigned L[4], H[4];
#pragma unroll
          for (int p = 0; p < 4; p++) {
            const int j = gq * 4 + p;
            const double2 a = P2[2 * j], b = P2[2 * j + 1];
            if (CDX_I8_EXP == 2) { L[p] = (unsigned)__double2loint(a.x); H[p] = (unsigned)__double2hiint(b.y); continue; }
            const double s = fma(xq, a.x, fma(yq, a.y, fma(zq, b.x, b.y + qq)));
            const unsigned shi = (unsigned)__double2hiint(s), slo = (unsigned)__double2loint(s);
            const unsigned fb = __funnelshift_l(slo, shi - 0x38000000u, 3);     // float(s) by bit moves (no F2F)
            const unsigned yb = __float_as_uint(rsqrt_ftz(__uint_as_float(fb)));
            const double yd = __hiloint2double((int)((yb >> 3) + 0x38000000u), (int)(yb << 29));
            const double hy = __hiloint2double((int)((yb >> 3) + 0x37F00000u), (int)(yb << 29));   // y / 2
            double d = s * yd;
            d = fma(fma(-d, d, s), hy, d);
            if (S >= 6) d = fma(fma(-d, d, s), hy, d);
            const double res = fma(s, fma(c2, d, c3), c0m);     // rint(k sc) in the low mantissa bits
            L[p] = (unsigned)__double2loint(res);
            H[p] = (unsigned)__double2hiint(res);
          }
          // 4x4 byte transposes: slice i holds byte S-1-i of the four values
          const unsigned t0 = __byte_perm(L[0], L[1], 0x5140), t1 = __byte_perm(L[2], L[3], 0x5140);
          const unsigned t2 = __byte_perm(L[0], L[1], 0x7362), t3 = __byte_perm(L[2], L[3], 0x7362);
          const unsigned t4 = __byte_perm(H[0], H[1], 0x5140), t5 = __byte_perm(H[2], H[3], 0x5140);
          unsigned by[6];
          by[0] = __byte_perm(t0, t1, 0x5410); by[1] = __byte_perm(t0, t1, 0x7632);
          by[2] = __byte_perm(t2, t3, 0x5410); by[3] = __byte_perm(t2, t3, 0x7632);
          by[4] = __byte_perm(t4, t5, 0x5410); by[5] = __byte_perm(t4, t5, 0x7632);
#pragma unroll
          for (int i = 0; i < S; i++) ow[i][gq] = by[S - 1 - i];
        }
#if CDX_I8_LATE_ARRIVE
        // the stores of the previous owned stage have had this stage's whole generation to land: complete them now
        if (pending >= 0) {
          tc_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mb_arrive(a_full + 8 * pending);
        }
#endif
        // the TMEM stage is free once the MMAs of stage g - NSA have completed (the values above did not need it)
        if (CDX_I8_EXP != 5) {
          if (lane == 0) mb_wait(a_empty + 8 * sa, ((unsigned)(g / NSA) & 1u) ^ 1u);
          __syncwarp();
          tc_fence_after();
        }
        if (CDX_I8_EXP == 3 || CDX_I8_EXP == 5 || CDX_I8_EXP == 6) {   // timing experiment: no tensor-memory stores (keep the values alive)
          unsigned x = 0;
#pragma unroll
          for (int i = 0; i < S; i++) x ^= ow[i][0] ^ ow[i][1] ^ ow[i][2] ^ ow[i][3];
          if (x == 0x12345677u) stdv[0] = 0.0;
        } else {
#pragma unroll
          for (int i = 0; i < S; i++) tc_st4(trow + sa * C::ACOLS + i * (BK / 4), ow[i][0], ow[i][1], ow[i][2], ow[i][3]);
        }
#if CDX_I8_LATE_ARRIVE
        pending = sa;
#else
        if (CDX_I8_EXP != 5) {
          tc_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mb_arrive(a_full + 8 * sa);
        }
#endif
      }
    }
#if CDX_I8_LATE_ARRIVE
    if (pending >= 0) {
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mb_arrive(a_full + 8 * pending);
    }
#endif
  } else if (warp == I8_WARP_LOAD) {
    // ===================================================== U slices + points: two bulk copies per stage
    if (lane == 0 && CDX_I8_EXP != 5) {
      int g = 0;
      for (int nt = 0; nt < NT; nt++) {
        const int nch = (nt + 1) * CPT;
        for (int c = 0; c < nch; c++, g++) {
          const int sbk = g % NSB;
          mb_wait_relaxed(b_empty + 8 * sbk, ((unsigned)(g / NSB) & 1u) ^ 1u);
          mb_expect_tx(b_full + 8 * sbk, (unsigned)(C::B_BYTES + C::P_BYTES));
          const unsigned dst = smem0 + sbk * C::SLOT;
          bulk_g2s(dst, Bsl + (size_t)(nt * NC + c) * C::B_BYTES, (unsigned)C::B_BYTES, b_full + 8 * sbk);
          bulk_g2s(dst + C::B_BYTES, pts + (size_t)c * BK * 4, (unsigned)C::P_BYTES, b_full + 8 * sbk);
        }
      }
    }
  } else if (warp == I8_WARP_MMA) {
    // ===================================================== MMA issue: the warp stays converged, one elected lane issues
    // INT32 accumulate (c_format 2 @4), A unsigned 8-bit (0 @7), B signed 8-bit (1 @10), K-major both, N>>3 @17, M>>4 @24
    constexpr unsigned idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((unsigned)(I8_TN >> 3) << 17) | ((unsigned)(I8_TQ >> 4) << 24);
    int g = 0;
    for (int nt = 0; nt < (CDX_I8_EXP == 5 ? 0 : NT); nt++) {
      mb_wait(acc_empty, ((unsigned)nt & 1u) ^ 1u);
      tc_fence_after();
      const int nch = (nt + 1) * CPT;
      for (int c = 0; c < nch; c++, g++) {
        const int sa = g % NSA, sbk = g % NSB;
        mb_wait(b_full + 8 * sbk, (unsigned)(g / NSB) & 1u);
        mb_wait_relaxed(a_full + 8 * sa, (unsigned)(g / NSA) & 1u, 32);
        tc_fence_after();
        const unsigned bbase = smem0 + sbk * C::SLOT;
        const unsigned abase = tmem + C::A_BASE + sa * C::ACOLS;
        if (elect_one()) {
#pragma unroll
          for (int d = 0; d < S; d++)
#pragma unroll
            for (int i = 0; i <= d; i++)
#pragma unroll
              for (int ks = 0; ks < BK / 32; ks++) {
                const uint64_t db = tc_desc(bbase + (d - i) * C::B_SLICE + ks * (2 * I8_TN * 16), I8_TN * 16, 128);
                if (CDX_I8_EXP == 0 || CDX_I8_EXP == 2 || CDX_I8_EXP == 6)
                  tc_mma_i8_ts(tmem + d * I8_TN, abase + i * (BK / 4) + ks * 8, db, idesc, (c > 0 || ks > 0 || i > 0) ? 1u : 0u);
              }
          tc_commit(a_empty + 8 * sa);
          tc_commit(b_empty + 8 * sbk);
          if (c == nch - 1) tc_commit(acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    // ===================================================== epilogue: lanes 32 w .. 32 w + 31 of TMEM, w = warp % 4
    const int lq = warp & 3, row = lq * 32 + lane;
    const int64_t q = q0 + row;
    const bool ok = q < M;
    const double xq = ok ? Xq[q * 3] : 0.0, yq = ok ? Xq[q * 3 + 1] : 0.0, zq = ok ? Xq[q * 3 + 2] : 0.0;
    const double sc = i8_query_scale<S>(xq, yq, zq, bb, R, R3);
    double ss = 0.0;
    for (int nt = 0; nt < (CDX_I8_EXP == 5 ? 0 : NT); nt++) {
      mb_wait_relaxed(acc_full, (unsigned)nt & 1u);
      tc_fence_after();
      const unsigned trow = tmem + ((unsigned)(lq * 32) << 16);
#pragma unroll 1
      for (int cg = 0; cg < I8_TN / 8; cg++) {
        int r[S][8];
#pragma unroll
        for (int d = 0; d < S; d++) tc_ld8(trow + d * I8_TN + cg * 8, r[d]);
        tc_wait_ld();
#pragma unroll
        for (int e = 0; e < 8; e++) {
          // sum_d 256^(S-1-d) acc_d: integer Horner for the leading S-2 terms and the last two (two conversions, not S)
          long long hi = r[0][e];
#pragma unroll
          for (int d = 1; d < S - 2; d++) hi = (hi << 8) + r[d][e];
          const long long lo = ((long long)r[S - 2][e] << 8) + r[S - 1][e];
          double h = fma((double)hi, 65536.0, (double)lo);
          h *= __ldg(sb + nt * I8_TN + cg * 8 + e);
          ss = fma(h, h, ss);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mb_arrive(acc_empty);
    }
    if (ok) {
      const double f = 1.0 / (sc * 256.0);   // S_q 256^-(S+1)
      stdv[q] = sqrt(fabs(prior - f * f * ss));   // gpis.py:59
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == I8_WARP_MMA) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(C::TMEM_COLS));
  }
}

template <int S>
int i8_prepare(cdx_ctx* ctx, cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  if (g.i8_S == S) return CDX_OK;
  const int Np = (int)g.Np, N = (int)g.N;
  const size_t need = (size_t)(Np / I8_TN) * (Np / i8cfg<S>::BK) * i8cfg<S>::B_BYTES;
  if (need > g.i8_cap) {
    cudaFree(g.i8_B);
    g.i8_B = nullptr;
    g.i8_cap = 0;
    if (cudaMalloc((void**)&g.i8_B, need) != cudaSuccess) {
      cudaGetLastError();
      cdx_set_error("out of device memory for the INT8 slices of L^-1 (N=%d, %d slices)", N, S);
      return CDX_ERR_ALLOC;
    }
    g.i8_cap = need;
  }
  if (g.Np > g.i8_np) {
    cudaFree(g.i8_sb); cudaFree(g.i8_eb); cudaFree(g.i8_pts);
    g.i8_np = 0;
    CDX_CUDA(cudaMalloc((void**)&g.i8_sb, sizeof(double) * Np));
    CDX_CUDA(cudaMalloc((void**)&g.i8_eb, sizeof(int) * Np));
    CDX_CUDA(cudaMalloc((void**)&g.i8_pts, sizeof(double) * 4 * Np));
    g.i8_np = g.Np;
  }
  if (!g.i8_bb) CDX_CUDA(cudaMalloc((void**)&g.i8_bb, sizeof(double) * 6));
  i8_rowscale_kernel<<<Np, 256, 0, st>>>(g.U, N, Np, g.i8_sb, g.i8_eb);
  CDX_CHECK_LAUNCH(ctx);
  i8_slice_kernel<S><<<dim3((Np + 255) / 256, Np), 256, 0, st>>>(g.U, N, Np, g.i8_eb, g.i8_B);
  CDX_CHECK_LAUNCH(ctx);
  i8_points_kernel<<<1, 256, 0, st>>>(g.Xa, N, Np, g.i8_pts, g.i8_bb);
  CDX_CHECK_LAUNCH(ctx);
  g.i8_S = S;
  if (!g.i8_ready) CDX_CUDA(cudaEventCreateWithFlags(&g.i8_ready, cudaEventDisableTiming));
  CDX_CUDA(cudaEventRecord(g.i8_ready, st));
  return CDX_OK;
}

template <int S>
int i8_launch(cdx_ctx* ctx, const double* X, int64_t M, double* stdv, cudaStream_t st) {
  int rc = i8_prepare<S>(ctx, st);
  if (rc) return rc;
  cdx_gpis_state& g = ctx->g;
  if (int rca = cdx_smem_attr(ctx, gpis_std_i8_kernel<S>, i8cfg<S>::SMEM)) return rca;
  CDX_CUDA(cudaStreamWaitEvent(st, g.i8_ready, 0));   // the slices may have been built on another stream
  const int64_t grid = (M + I8_TQ - 1) / I8_TQ;
  gpis_std_i8_kernel<S><<<(unsigned)grid, I8_THREADS, i8cfg<S>::SMEM, st>>>(X, M, g.i8_pts, g.i8_B, g.i8_sb, g.i8_bb, (int)g.Np,
                                                                             g.R, g.R * g.R * g.R, g.prior, stdv);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

}  // namespace

// TPS only (the fixed-point scale uses the cubic's bound); the INT32 accumulators hold 255*128*S*Np < 2^31 up to Np = 8192
bool cdx_i8_eligible(const cdx_ctx* ctx) {
  const cdx_gpis_state& g = ctx->g;
  return g.kernel == CDX_KERNEL_TPS && g.Np >= 128 && g.Np <= 8192;
}

int cdx_i8_std(cdx_ctx* ctx, int S, const double* X, int64_t M, double* stdv, cudaStream_t st) {
  switch (S) {
    case 4: return i8_launch<4>(ctx, X, M, stdv, st);
    case 5: return i8_launch<5>(ctx, X, M, stdv, st);
    case 6: return i8_launch<6>(ctx, X, M, stdv, st);
    default: cdx_set_error("cdx_i8_std: slices must be 4, 5 or 6"); return CDX_ERR_ARG;
  }
}

void cdx_i8_free(cdx_ctx* ctx) {
  cdx_gpis_state& g = ctx->g;
  cudaFree(g.i8_B); cudaFree(g.i8_sb); cudaFree(g.i8_eb); cudaFree(g.i8_pts); cudaFree(g.i8_bb);
  g.i8_B = nullptr; g.i8_sb = nullptr; g.i8_eb = nullptr; g.i8_pts = nullptr; g.i8_bb = nullptr;
  g.i8_S = 0; g.i8_cap = 0; g.i8_np = 0;
  if (g.i8_ready) { cudaEventDestroy(g.i8_ready); g.i8_ready = nullptr; }
}

extern "C" int cdx_gpis_set_var_mode(cdx_ctx* ctx, int mode) {
  if (!ctx || !(mode == 0 || mode == -1 || (mode >= 4 && mode <= 6))) {
    cdx_set_error("cdx_gpis_set_var_mode: mode must be 0 (auto), -1 (FP64 DMMA) or 4..6 (INT8 slices)");
    return CDX_ERR_ARG;
  }
  ctx->g.var_mode = mode;
  return CDX_OK;
}
