// GPIS posterior std on tcgen05 (INT8 kind, INT32 accumulators in TMEM) - "rows as M" formulation with the generated
// operand shared by a thread-block cluster.
//
// sigma(x)^2 = k(x,x) - |v|^2, v = L^-1 k*(x)  (gpis.py:51-59; the reference solves an LU per call).  v = U k* with the
// precomputed lower-triangular U = L^-1 is a dense triangular contraction, but tcgen05 has no FP64 kind.  It is evaluated as an
// error-free INTEGER product (Ozaki-style slicing): k*(x) as S unsigned byte slices a_i of a per-query fixed-point value, U as
// S balanced signed base-256 digits b_j with a per-row exponent,
//   v_n = S_q 2^e_n 256^-(S+1) sum_{d<S} 256^(S-1-d) sum_{i+j=d} sum_k b_j[n,k] a_i[k],   every inner sum an exact INT32 MMA;
// the only error is the truncation of the operands to 8S-2 bits and of the slice pairs to i+j < S (measured on B200 against
// the FP64 DMMA path: S = 5 gives |d var| < 2e-8 prior, the gate is 1e-6 prior).
// Round 1's kernel made the 128 QUERIES of a CTA the M dimension (k* slices generated into tensor memory) and 64 training rows
// N; tensor memory holds only S x 64 accumulator columns, so k* was re-generated in FP64 for every one of the N/64 row tiles -
// 64 times at N = 4096 - and that generation, not the tensor pipe, bounded the kernel.  Here 128 ROWS of U are M (A operand, pre-sliced in HBM in the exact
// shared-memory image, one bulk copy per stage), NQ = 96 QUERIES are N (S x 96 = 480 accumulator columns): the k* slices
// of a (query block, 32 training points) stage are the B operand in shared memory, and the C CTAs of a cluster take
// DIFFERENT row tiles of the same query block, each generating 1/C of the stage's k* and writing it into every CTA's
// shared memory (st.shared::cluster).  A generated value now serves 128 C rows instead of 64: 4x (C = 2) fewer FP64
// generations per query, and each CTA generates only its share.
//
// Roles per CTA (22 warps): warps 0-11 generate (4 stage groups of 3 warps: one (query, 16 training points) item per lane
// and stage; 9 FP64 ops + 1 MUFU.RSQ per value, PRMT byte transposes, one 16-byte st.shared per slice locally and one st.async
// per slice into each peer - which completes tx bytes on the peer's mbarrier, so no cluster-scope fence or remote arrive is
// needed: both cost a MEMBAR.ALL.GPU per warp and stage and made the first version generator-bound); warp 12 issues the S(S+1)/2 tcgen05.mma per stage
// (both operands from shared memory, M = 128, N = 96, K = 32) and releases the stage in every CTA with a multicast
// tcgen05.commit; warp 13 streams the U slices + the stage's training points (cp.async.bulk); warps 14-21 drain the S
// diagonal accumulators after each row tile (tcgen05.ld), recombine them, square, and reduce over the 128 rows.
#include <algorithm>
#include "cdx_internal.cuh"

namespace {

constexpr int X_TM = 128;        // U rows per tile (UMMA M)
constexpr int X_BK = 32;         // training points per stage (one K = 32 MMA step per slice pair)
constexpr int X_GEN_WARPS = 12, X_WARP_MMA = 12, X_WARP_LOAD = 13, X_WARP_EPI = 14, X_EPI_WARPS = 8;
constexpr int X_THREADS = 22 * 32;
constexpr double X_MAGIC = 6755399441055744.0;  // 1.5 * 2^52

// PAIR: the two CTAs of the cluster issue ONE tcgen05.mma.cta_group::2 (M = 256: 128 rows of U from each CTA) whose B operand
// is split between them - each CTA holds (and generates) the k* slices of HALF the block's queries, so nothing crosses the
// cluster through shared memory and every MMA reads half as many B bytes per CTA.
template <int S, bool PAIR = false> struct xcfg {
  static constexpr int NQ = S <= 4 ? 128 : (S == 5 ? 96 : (PAIR ? 64 : 80));   // queries per cluster block: S * NQ <= 512 TMEM columns (cta_group::2: N % 32 == 0)
  static constexpr int NB = PAIR ? NQ / 2 : NQ;         // rows of the B tile held by one CTA
  static constexpr int A_SLICE = X_TM * X_BK;           // bytes of one slice of one U tile stage
  static constexpr int A_BYTES = S * A_SLICE;
  static constexpr int B_SLICE = NB * X_BK;
  static constexpr int B_BYTES = S * B_SLICE;
  static constexpr int P_BYTES = X_BK * 32;             // (-2x,-2y,-2z,|x|^2) of one stage's training points
  static constexpr int NP = 16;                         // stages of the points ring: the generators run this far ahead of the MMAs
  static constexpr int SLOT = A_BYTES + B_BYTES;        // one stage of the operand ring: U digit slices | k* byte slices
  static constexpr int NST = (220 * 1024 - NP * P_BYTES) / SLOT;   // S = 5: 5 x 35 KB; pair: 7 x 27.5 KB
  static constexpr int RING = NST * SLOT + NP * P_BYTES;
  static constexpr size_t SMEM = (size_t)RING + 1024 + 8 * 2 * NQ;   // + barriers + cross-warp / cross-CTA partial sums
  static_assert(S * NQ <= 512 && NQ % 16 == 0 && (!PAIR || NQ % 32 == 0), "TMEM budget / UMMA N");
  static_assert(NST >= 3 && NP > NST, "pipeline depth");
};

__device__ __forceinline__ unsigned xs32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void xmb_init(unsigned bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
// CTA-scope release: a plain SYNCS.ARRIVE (a cluster-scope release costs a MEMBAR.ALL.GPU per arrive)
__device__ __forceinline__ void xmb_arrive_local(unsigned bar) {
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void xmb_arrive_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.release.cta.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void xmb_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
// cluster-scope acquire: the data guarded by the barrier may have been written by another CTA of the cluster
__device__ __forceinline__ void xmb_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nXW_%=:\n"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
      "@p bra XD_%=;\nbra XW_%=;\nXD_%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void xmb_wait_relaxed(unsigned bar, unsigned parity, unsigned ns = 64) {
  unsigned done = 0;
  for (;;) {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(ns);
  }
}
__device__ __forceinline__ unsigned xmapa(unsigned local_addr, unsigned rank) {
  unsigned r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void xbulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void xst_local16(unsigned addr, unsigned a, unsigned b, unsigned c, unsigned d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// 16 bytes into a peer CTA's shared memory; the store itself completes 16 tx bytes on the PEER's mbarrier (release at
// cluster scope is part of complete_tx), so the sender needs neither a fence nor a remote arrive
__device__ __forceinline__ void xst_async16(unsigned addr, unsigned a, unsigned b, unsigned c, unsigned d, unsigned rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];\n" ::"r"(addr),
               "r"(a), "r"(b), "r"(c), "r"(d), "r"(rbar) : "memory");
}
__device__ __forceinline__ void xfence_proxy_async_cta() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void xtc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void xtc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void xtc_commit(unsigned bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// the same barrier (CTA-relative offset) in every CTA of `mask`
__device__ __forceinline__ void xtc_commit_mc(unsigned bar, unsigned short mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(bar),
               "h"(mask) : "memory");
}
__device__ __forceinline__ void xtc_mma_i8(unsigned tmem_d, uint64_t da, uint64_t db, unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
// cta_group::2: issued by the leader CTA for both; the barrier of every CTA in `mask` is signalled when the MMAs complete
__device__ __forceinline__ void xtc_commit_pair(unsigned bar, unsigned short mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(bar),
               "h"(mask) : "memory");
}
__device__ __forceinline__ void xtc_mma_i8_pair(unsigned tmem_d, uint64_t da, uint64_t db, unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
// release at cluster scope (a MEMBAR.ALL.GPU): once per round by the peer's epilogue
__device__ __forceinline__ void xmb_arrive_remote(unsigned bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(bar_cluster_addr) : "memory");
}
// relaxed: the peer's relay signals "my half of the stage is in MY shared memory" - data that only the peer SM's own tensor
// core reads (cta_group::2), that the relay has acquired through the peer's local barriers, and that its writers have already
// fenced into the async proxy; a release here would cost a MEMBAR.ALL.GPU per stage on the relay's serial path
__device__ __forceinline__ void xmb_arrive_remote_relaxed(unsigned bar_cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ uint64_t xtc_desc(unsigned saddr, unsigned lbo, unsigned sbo) {   // K-major, no swizzle (8x16-byte core matrices)
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void xtc_ld8(unsigned taddr, int (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ void xtc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ bool xelect_one() {
  unsigned pred;
  asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, 0xffffffff;\nselp.u32 %0, 1, 0, px;\n}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ float xrsqrt_ftz(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void xcluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ unsigned xcluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}

// ------------------------------------------------------------------------------------------- state preparation
// per training row n: exponent e_n with |U[n,:]| 2^-e_n < 1/4 (so S balanced base-256 digits never overflow); rows >= N
// (padding, up to the tile-rounded row count) get scale 0
__global__ void __launch_bounds__(256) x_rowscale_kernel(const double* __restrict__ U, int N, int Np, int rows,
                                                         double* __restrict__ sb, int* __restrict__ eb) {
  const int n = blockIdx.x;
  __shared__ double red[256];
  double m = 0.0;
  if (n < N)
    for (int k = threadIdx.x; k <= n; k += 256) m = fmax(m, fabs(U[(int64_t)n * Np + k]));
  red[threadIdx.x] = m;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0 && n < rows) {
    int e = 0;
    if (red[0] > 0.0) frexp(red[0], &e);
    e += 2;
    eb[n] = e;
    sb[n] = n < N ? ldexp(1.0, e) : 0.0;
  }
}

// digits of U in the shared-memory image of the A operand: [row tile 128][k chunk 32][slice][k group of 16][row 128][16 bytes]
template <int S>
__global__ void __launch_bounds__(256) x_slice_kernel(const double* __restrict__ U, int N, int Np, int rows, const int* __restrict__ eb,
                                                      int8_t* __restrict__ Asl) {
  const int k = blockIdx.x * 256 + threadIdx.x, n = blockIdx.y;
  if (k >= Np || n >= rows) return;
  const int NC = Np / X_BK;
  long long V = 0;
  if (n < N && k <= n) V = llrint(ldexp(U[(int64_t)n * Np + k], 8 * S - eb[n]));
  const int nt = n / X_TM, nl = n % X_TM, c = k / X_BK, kk = k % X_BK;
  int8_t* base = Asl + ((size_t)(nt * NC + c) * S) * xcfg<S>::A_SLICE + (kk / 16) * (X_TM * 16) + nl * 16 + (kk % 16);
#pragma unroll
  for (int j = S - 1; j >= 0; j--) {
    const long long d = ((V + 128) & 255) - 128;
    V = (V - d) >> 8;
    base[(size_t)j * xcfg<S>::A_SLICE] = (int8_t)d;
  }
}

__global__ void __launch_bounds__(256) x_points_kernel(const double* __restrict__ Xa, int N, int Np, double* __restrict__ pts,
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
__device__ __forceinline__ double x_query_scale(double x, double y, double z, const double* __restrict__ bb, double R, double R3) {
  const double ax = fmax(fabs(x - bb[0]), fabs(x - bb[3])), ay = fmax(fabs(y - bb[1]), fabs(y - bb[4])),
               az = fmax(fabs(z - bb[2]), fabs(z - bb[5]));
  const double d2 = fma(ax, ax, fma(ay, ay, az * az)), dm = sqrt(d2);
  const double kmax = fmax(R3, fma(d2, fma(2.0, dm, -3.0 * R), R3));
  int e;
  frexp(kmax, &e);
  return ldexp(1.0, 8 * S - e - 1);
}

// ------------------------------------------------------------------------------------------- k* slices, generated once
// per-query fixed-point scale (a power of two), once per query
template <int S>
__global__ void __launch_bounds__(256) i8x_scale_kernel(const double* __restrict__ Xq, int64_t M, int64_t Mc, const double* __restrict__ bb,
                                                        double R, double R3, double* __restrict__ scq) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= Mc) return;
  const bool ok = i < M;
  scq[i] = x_query_scale<S>(ok ? Xq[i * 3] : 0.0, ok ? Xq[i * 3 + 1] : 0.0, ok ? Xq[i * 3 + 2] : 0.0, bb, R, R3);
}

// One thread per (query, 512 training points): per value 9 FP64 operations + 1 MUFU.RSQ (expanded squared distance, float
// rsqrt seed built by integer bit moves, one Newton step, cubic, fixed-point conversion by the 1.5 * 2^52 trick), PRMT byte
// transposes, and per 16 points one 16-byte store per slice straight into the image the MMA kernel's bulk copies expect:
//   kq[(query block, CTA of the pair)][stage of 32 points][slice][16-point group][query row of that CTA][16 bytes].
// A warp = 32 consecutive queries at the same points (uniform point loads, 512 contiguous bytes per store instruction).
// MEAN: the posterior mean and its gradient (gpis.py:55, :80-82) need the same distances: mu = sum_k alpha_k k,
// grad mu = sum_k alpha_k c_k (x - x_k) are accumulated on the way (10 more FP64 operations per pair instead of a second
// kernel's 20) into one partial per (512-point part, query); i8x_mean_reduce_kernel adds the parts in order - the split
// depends on the state only, so a query's mean is the same for every batch.
constexpr int X_PART = 512;
template <int S, int C, bool PAIR, bool MEAN>
__global__ void __launch_bounds__(128) i8x_kstar_kernel(const double* __restrict__ Xq, int64_t M, int64_t Mc, const double* __restrict__ pts,
                                                        const double* __restrict__ Xa, const double* __restrict__ scq, int Np, double R,
                                                        double R3, int8_t* __restrict__ kq, double* __restrict__ part) {
  using G = xcfg<S, PAIR>;
  // the part's training records (32 B each) and weights, staged once per CTA: every warp reads them at warp-uniform
  // addresses, which from shared memory is a broadcast with a short, fixed latency - from global memory (through L1) the
  // same loads left the kernel waiting on the scoreboard (ncu: long_scoreboard 7.5 stall cycles per issue, FP64 pipe 32 %)
  __shared__ double2 sP[X_PART * 2];
  __shared__ double sAl[X_PART];
  {
    const int p0 = blockIdx.y * X_PART, np = min(X_PART, Np - p0);
    const double2* src = reinterpret_cast<const double2*>(pts) + (size_t)p0 * 2;
    for (int i = threadIdx.x; i < np * 2; i += 128) sP[i] = __ldg(src + i);
    if (MEAN)
      for (int i = threadIdx.x; i < np; i += 128) sAl[i] = __ldg(Xa + 4 * (size_t)(p0 + i) + 3);
  }
  __syncthreads();
  const int64_t ql = (int64_t)blockIdx.x * 128 + threadIdx.x;
  if (ql >= Mc) return;
  const bool ok = ql < M;
  const double xq = ok ? Xq[ql * 3] : 0.0, yq = ok ? Xq[ql * 3 + 1] : 0.0, zq = ok ? Xq[ql * 3 + 2] : 0.0;
  const double qq = fma(xq, xq, fma(yq, yq, zq * zq)) + 1e-16;   // keeps the expanded squared distance positive
  const double sc = scq[ql];
  const double m3R = -3.0 * R, m6R = -6.0 * R;
  const int64_t block = ql / G::NQ;
  const int qcol = (int)(ql % G::NQ), cta = PAIR ? qcol / G::NB : 0, row = PAIR ? qcol % G::NB : qcol;
  const int NC = Np / X_BK;
  int8_t* dst0 = kq + (size_t)(block * C + cta) * NC * G::B_BYTES + row * 16;
  const int g0 = blockIdx.y * (X_PART / 16), g1 = min(g0 + X_PART / 16, Np / 16);
  double am[4] = {0, 0, 0, 0}, ax[4] = {0, 0, 0, 0}, ay[4] = {0, 0, 0, 0}, az[4] = {0, 0, 0, 0};
  for (int c16 = g0; c16 < g1; c16++) {
    const double2* P2 = sP + (c16 - g0) * 32;
    unsigned ow[S][4];
#pragma unroll
    for (int gq = 0; gq < 4; gq++) {
      unsigned L[4], H[4];
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const int j = gq * 4 + p;
        const double2 a = P2[2 * j], b = P2[2 * j + 1];
        const double sd = fma(xq, a.x, fma(yq, a.y, fma(zq, b.x, b.y + qq)));
        const unsigned shi = (unsigned)__double2hiint(sd), slo = (unsigned)__double2loint(sd);
        const unsigned fb = __funnelshift_l(slo, shi - 0x38000000u, 3);     // float(s) by bit moves (no F2F)
        const unsigned yb = __float_as_uint(xrsqrt_ftz(__uint_as_float(fb)));
        const double yd = __hiloint2double((int)((yb >> 3) + 0x38000000u), (int)(yb << 29));
        const double hy = __hiloint2double((int)((yb >> 3) + 0x37F00000u), (int)(yb << 29));   // y / 2
        double d = sd * yd;
        d = fma(fma(-d, d, sd), hy, d);
        d = fma(fma(-d, d, sd), hy, d);                      // second Newton step: generated once per query, it is cheap here
        const double k = fma(sd, fma(2.0, d, m3R), R3);      // 2 d^3 - 3 R d^2 + R^3 (gpis.py:26)
        const double res = fma(k, sc, X_MAGIC);              // rint(k sc) in the low mantissa bits (sc is a power of two)
        L[p] = (unsigned)__double2loint(res);
        H[p] = (unsigned)__double2hiint(res);
        if (MEAN) {
          const double al = sAl[(c16 - g0) * 16 + j];      // alpha (0 on padded rows)
          am[p] = fma(al, k, am[p]);
          const double ac = al * fma(6.0, d, m6R);           // alpha c,  c = 6 (d - R)
          ax[p] = fma(ac, fma(0.5, a.x, xq), ax[p]);        // x - x_k  (the record holds -2 x_k)
          ay[p] = fma(ac, fma(0.5, a.y, yq), ay[p]);
          az[p] = fma(ac, fma(0.5, b.x, zq), az[p]);
        }
      }
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
    int8_t* dst = dst0 + (size_t)(c16 >> 1) * G::B_BYTES + (c16 & 1) * (G::NB * 16);
#pragma unroll
    for (int i = 0; i < S; i++) *reinterpret_cast<uint4*>(dst + (size_t)i * G::B_SLICE) = make_uint4(ow[i][0], ow[i][1], ow[i][2], ow[i][3]);
  }
  if (MEAN) {
    double* o = part + ((size_t)blockIdx.y * Mc + ql) * 4;
    o[0] = (am[0] + am[1]) + (am[2] + am[3]); o[1] = (ax[0] + ax[1]) + (ax[2] + ax[3]);
    o[2] = (ay[0] + ay[1]) + (ay[2] + ay[3]); o[3] = (az[0] + az[1]) + (az[2] + az[3]);
  }
}

__global__ void __launch_bounds__(256) i8x_mean_reduce_kernel(const double* __restrict__ part, int parts, int64_t M, int64_t Mc, double bias,
                                                              double* __restrict__ mean, double* __restrict__ gmean) {
  const int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (q >= M) return;
  double v[4] = {0, 0, 0, 0};
  for (int p = 0; p < parts; p++)
#pragma unroll
    for (int e = 0; e < 4; e++) v[e] += part[((size_t)p * Mc + q) * 4 + e];
  if (mean) mean[q] = v[0] + bias;
  if (gmean) { gmean[q * 3] = v[1]; gmean[q * 3 + 1] = v[2]; gmean[q * 3 + 2] = v[3]; }
}

// ------------------------------------------------------------------------------------------- the query kernel
// One cluster of C CTAs per block of NQ queries; CTA `rank` takes the row tiles rank, rank + C, ... (a "round" = one tile per
// CTA); every CTA of the cluster walks the same stage sequence (round r: k chunks 0 .. (r C + C) 128/32 - 1, the bound of
// the round's LAST tile; the few chunks above a tile's own diagonal hold zero digits).
// PRE: the k* byte slices were generated beforehand (i8x_kstar_kernel) into `kq`, laid out per (query block, CTA, stage) in
// the shared-memory image of the B operand: the loader brings them in with one more bulk copy per stage and the generator
// warps have nothing to do.  On B200 FP64 instructions and tcgen05 MMAs exclude each other (csrc/cdx_mixbench.cu: while
// MMAs are queued the FP64 pipe issues ~2 % of its rate), so generating inside this kernel only ADDS the generation time to
// the MMA time - and repeats it for every round; generated once, on the whole chip, it costs a few per cent of the MMAs.
template <int S, int C, bool PAIR, bool PRE>
__global__ void __launch_bounds__(X_THREADS, 1)
gpis_std_i8x_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ pts, const int8_t* __restrict__ Asl,
                    const double* __restrict__ sb, const double* __restrict__ bb, int Np, int NT, double R, double R3, double prior,
                    double* __restrict__ stdv, const int8_t* __restrict__ kq) {
  static_assert(!PAIR || C == 2, "a CTA pair is a cluster of two");
  using G = xcfg<S, PAIR>;
  constexpr int NQ = G::NQ, NB = G::NB, NST = G::NST, QPC = NQ / C;  // QPC: queries whose k* this CTA generates
  constexpr int ITEMS = QPC * (X_BK / 16);                            // (query, 16-point group) items per stage and CTA
  constexpr int WPS = (ITEMS + 31) / 32;                              // generator warps per stage
  constexpr int NGRP = X_GEN_WARPS / WPS;                             // stage groups the generator warps are split into
  static_assert(NGRP >= 1 && NQ % C == 0, "generator split");
  extern __shared__ __align__(128) unsigned char smem[];
  const unsigned smem0 = xs32(smem);
  constexpr int NP = G::NP;
  const unsigned pts0 = smem0 + NST * G::SLOT;         // points ring [NP][32 points][4 doubles]
  const unsigned bar0 = smem0 + G::RING;
  const unsigned a_full = bar0;                        // [NST] U slices landed (tx bytes)
  const unsigned b_full = bar0 + 8 * NST;              // [NST] every generator warp of every CTA has written its k* share
  const unsigned empty = bar0 + 16 * NST;              // [NST] the MMAs of EVERY CTA have finished with the stage (multicast commit)
  const unsigned peer_ready = bar0 + 24 * NST;         // [NST] (pair, leader only) the peer CTA's half of the stage is complete
  const unsigned p_full = bar0 + 32 * NST;             // [NP] a stage's training points landed (tx bytes)
  const unsigned p_empty = p_full + 8 * NP;            // [NP] the stage's generator warps have read them
  const unsigned acc_full = p_empty + 8 * NP;
  const unsigned acc_empty = acc_full + 8;
  unsigned* tmem_slot = reinterpret_cast<unsigned*>(smem + G::RING + 32 * NST + 16 * NP + 16);
  double* part = reinterpret_cast<double*>(smem + G::RING + 1024);   // [2][NQ]: [0] this CTA's sums, [1] the peer's

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const unsigned rank = C > 1 ? xcluster_rank() : 0u;
  const int64_t q0 = (int64_t)(blockIdx.x / C) * NQ;
  const int NC = Np / X_BK;
  const int rounds = (NT + C - 1) / C;

  if (t == 0) {
    for (int i = 0; i < NST; i++) {
      xmb_init(a_full + 8 * i, 1);
      xmb_init(b_full + 8 * i, WPS);
      xmb_init(empty + 8 * i, PAIR ? 1 : C);             // pair: one multicast commit by the leader; else one per CTA
      xmb_init(peer_ready + 8 * i, 1);
    }
    for (int i = 0; i < NP; i++) {
      xmb_init(p_full + 8 * i, 1);
      xmb_init(p_empty + 8 * i, WPS);
    }
    xmb_init(acc_full, 1);
    xmb_init(acc_empty, PAIR ? 2 * X_EPI_WARPS : X_EPI_WARPS);   // pair: the leader waits for both CTAs' epilogues
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (PAIR) {                            // both CTAs of the pair take part in a cta_group::2 allocation (same warp, same slot offset)
    __syncthreads();
    xcluster_sync();
  }
  if (warp == X_WARP_MMA) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(xs32(tmem_slot)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(xs32(tmem_slot)));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
    }
  }
  xtc_fence_before();
  __syncthreads();
  if (C > 1) xcluster_sync();          // every CTA's barriers are initialised before any remote arrive / multicast commit
  xtc_fence_after();
  const unsigned tmem = *tmem_slot;
  auto stages_of = [&](int r) { return min((r * C + C) * (X_TM / X_BK), NC); };

  if (warp < X_GEN_WARPS) {
    // ===================================================== generators
    const int grp = PRE ? NGRP : warp / WPS, wg = warp % WPS;     // PRE: no generation in this kernel
    const int item = wg * 32 + lane;
    const bool active = grp < NGRP && item < ITEMS;
    const int nact = min(32, ITEMS - wg * 32);                       // active lanes of this warp (the same in every CTA)
    const int ql = active ? item % QPC : 0, kg = active ? item / QPC : 0;
    const int qcol = (int)rank * QPC + ql;                            // column of the query in the cluster's block
    const int64_t q = q0 + qcol;
    const bool ok = q < M;
    const double xq = ok ? Xq[q * 3] : 0.0, yq = ok ? Xq[q * 3 + 1] : 0.0, zq = ok ? Xq[q * 3 + 2] : 0.0;
    const double qq = fma(xq, xq, fma(yq, yq, zq * zq)) + 1e-16;
    const double sc = x_query_scale<S>(xq, yq, zq, bb, R, R3);
    const double m3R = -3.0 * R;
    const unsigned boff = (unsigned)(G::A_BYTES + kg * (NB * 16) + (PAIR ? ql : qcol) * 16);     // within a stage slot, slice 0
    if (grp < NGRP) {
      int g = 0;
      for (int r = 0; r < rounds; r++) {
        const int ns = stages_of(r);
        for (int c = 0; c < ns; c++, g++) {
          if ((g % NGRP) != grp) continue;
          const int s = g % NST;
          const unsigned ph = (unsigned)(g / NST) & 1u;
          const int ps = g % NP;
          xmb_wait(p_full + 8 * ps, (unsigned)(g / NP) & 1u);         // the stage's training points have landed
          const double2* P2 = reinterpret_cast<const double2*>(smem + (size_t)NST * G::SLOT + (size_t)ps * G::P_BYTES) + kg * 32;
          unsigned ow[S][4];
#pragma unroll
          for (int gq = 0; gq < 4; gq++) {
            unsigned L[4], H[4];
#pragma unroll
            for (int p = 0; p < 4; p++) {
              const int j = gq * 4 + p;
              const double2 a = P2[2 * j], b = P2[2 * j + 1];
              const double sd = fma(xq, a.x, fma(yq, a.y, fma(zq, b.x, b.y + qq)));
              const unsigned shi = (unsigned)__double2hiint(sd), slo = (unsigned)__double2loint(sd);
              const unsigned fb = __funnelshift_l(slo, shi - 0x38000000u, 3);     // float(s) by bit moves (no F2F)
              const unsigned yb = __float_as_uint(xrsqrt_ftz(__uint_as_float(fb)));
              const double yd = __hiloint2double((int)((yb >> 3) + 0x38000000u), (int)(yb << 29));
              const double hy = __hiloint2double((int)((yb >> 3) + 0x37F00000u), (int)(yb << 29));   // y / 2
              double d = sd * yd;
              d = fma(fma(-d, d, sd), hy, d);
              d = fma(fma(-d, d, sd), hy, d);
              const double res = fma(fma(sd, fma(2.0, d, m3R), R3), sc, X_MAGIC);   // the arithmetic of i8x_kstar_kernel, bit for bit
              L[p] = (unsigned)__double2loint(res);
              H[p] = (unsigned)__double2hiint(res);
            }
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
          __syncwarp();
          if (lane == 0) xmb_arrive_local(p_empty + 8 * ps);          // the points slot can be refilled
          // the slot's previous k* has been consumed by the MMAs of EVERY CTA of the cluster
          xmb_wait(empty + 8 * s, ph ^ 1u);
          const unsigned dst = smem0 + (unsigned)s * G::SLOT + boff;
          if (active) {
#pragma unroll
            for (int i = 0; i < S; i++) xst_local16(dst + i * G::B_SLICE, ow[i][0], ow[i][1], ow[i][2], ow[i][3]);
            if (C > 1 && !PAIR) {
#pragma unroll
              for (int pr = 1; pr < C; pr++) {
                const unsigned peer = (rank + pr) % C;
                const unsigned rdst = xmapa(dst, peer), rbar = xmapa(b_full + 8 * s, peer);
#pragma unroll
                for (int i = 0; i < S; i++) xst_async16(rdst + i * G::B_SLICE, ow[i][0], ow[i][1], ow[i][2], ow[i][3], rbar);
              }
            }
          }
          xfence_proxy_async_cta();       // the local generic-proxy stores -> visible to the tensor core's async proxy
          __syncwarp();
          // one arrival per warp; it also announces the bytes the SAME warp of every peer CTA sends into this CTA's stage
          if (lane == 0) {
            if (C > 1 && !PAIR) xmb_arrive_expect_tx(b_full + 8 * s, (unsigned)(nact * S * 16 * (C - 1)));
            else xmb_arrive_local(b_full + 8 * s);
          }
        }
      }
    }
  } else if (warp == X_WARP_LOAD) {
    // ===================================================== U slices (operand ring) + training points (their own, deeper ring)
    // one thread walks both schedules: the points of stage k are requested NP - NST stages before the U tile of stage k - so
    // the generators' arithmetic is never inside the operand ring's re-use loop (TMA latency + generation + MMA per NST stages)
    if (lane == 0 && PRE) {
      // U digit slices + the pre-generated k* slices of this CTA's rows of the B operand: two bulk copies per stage
      const int8_t* kqc = kq + ((size_t)(blockIdx.x / C) * C + rank) * (size_t)NC * G::B_BYTES;
      int g = 0;
      for (int r = 0; r < rounds; r++) {
        const int ns = stages_of(r);
        const int nt = r * C + (int)rank;
        for (int c = 0; c < ns; c++, g++) {
          const int sa = g % NST;
          xmb_wait_relaxed(empty + 8 * sa, ((unsigned)(g / NST) & 1u) ^ 1u);
          xmb_expect_tx(a_full + 8 * sa, (unsigned)(G::A_BYTES + G::B_BYTES));
          const unsigned dst = smem0 + (unsigned)sa * G::SLOT;
          xbulk_g2s(dst, Asl + (size_t)(nt * NC + c) * G::A_BYTES, (unsigned)G::A_BYTES, a_full + 8 * sa);
          xbulk_g2s(dst + G::A_BYTES, kqc + (size_t)c * G::B_BYTES, (unsigned)G::B_BYTES, a_full + 8 * sa);
        }
      }
    } else if (lane == 0) {
      int total = 0;
      for (int r = 0; r < rounds; r++) total += stages_of(r);
      constexpr int D = NP - NST;
      int ra = 0, ca = 0;                                              // (round, chunk) of the next U tile
      int rp = 0, cp = 0;                                              // ... of the next points stage
      for (int k = 0; k < total + D; k++) {
        if (k < total) {
          const int ps = k % NP;
          xmb_wait_relaxed(p_empty + 8 * ps, ((unsigned)(k / NP) & 1u) ^ 1u);
          xmb_expect_tx(p_full + 8 * ps, (unsigned)G::P_BYTES);
          xbulk_g2s(pts0 + (unsigned)ps * G::P_BYTES, pts + (size_t)cp * X_BK * 4, (unsigned)G::P_BYTES, p_full + 8 * ps);
          if (++cp == stages_of(rp)) { cp = 0; rp++; }
        }
        const int ga = k - D;
        if (ga >= 0) {
          const int sa = ga % NST;
          const int nt = ra * C + (int)rank;                           // rows >= Np: the slice image holds zero tiles up to rounds * C
          xmb_wait_relaxed(empty + 8 * sa, ((unsigned)(ga / NST) & 1u) ^ 1u);
          xmb_expect_tx(a_full + 8 * sa, (unsigned)G::A_BYTES);
          xbulk_g2s(smem0 + (unsigned)sa * G::SLOT, Asl + (size_t)(nt * NC + ca) * G::A_BYTES, (unsigned)G::A_BYTES, a_full + 8 * sa);
          if (++ca == stages_of(ra)) { ca = 0; ra++; }
        }
      }
    }
  } else if (warp == X_WARP_MMA) {
    // ===================================================== MMA issue
    // INT32 accumulate (c_format 2 @4), A signed 8-bit (1 @7), B unsigned 8-bit (0 @10), K-major both, N>>3 @17, M>>4 @24
    constexpr unsigned idesc = (2u << 4) | (1u << 7) | (0u << 10) | ((unsigned)(NQ >> 3) << 17) |
                               ((unsigned)((PAIR ? 2 * X_TM : X_TM) >> 4) << 24);
    constexpr unsigned short mask = (unsigned short)((1u << C) - 1u);
    int g = 0;
    if (PAIR && rank != 0) {
      // the peer of a pair issues nothing: this warp tells the leader when the peer's half of each stage is complete
      const unsigned lead_ready = xmapa(peer_ready, 0);
      for (int r = 0; r < rounds; r++) {
        const int ns = stages_of(r);
        for (int c = 0; c < ns; c++, g++) {
          const int s = g % NST;
          const unsigned ph = (unsigned)(g / NST) & 1u;
          xmb_wait(a_full + 8 * s, ph);
          if (!PRE) xmb_wait(b_full + 8 * s, ph);
          if (lane == 0) xmb_arrive_remote_relaxed(lead_ready + 8 * s);
          __syncwarp();
        }
      }
    } else {
      for (int r = 0; r < rounds; r++) {
        xmb_wait(acc_empty, ((unsigned)r & 1u) ^ 1u);
        xtc_fence_after();
        const int ns = stages_of(r);
        for (int c = 0; c < ns; c++, g++) {
          const int s = g % NST;
          const unsigned ph = (unsigned)(g / NST) & 1u;
          xmb_wait(a_full + 8 * s, ph);
          if (!PRE) xmb_wait(b_full + 8 * s, ph);
          if (PAIR) xmb_wait(peer_ready + 8 * s, ph);
          if (!PRE) xfence_proxy_async_cta();   // k* bytes written through the generic proxy (locally, or by st.async from the peers)
          xtc_fence_after();
          const unsigned abase = smem0 + (unsigned)s * G::SLOT, bbase = abase + G::A_BYTES;
          if (xelect_one()) {
            // k* slice i outermost: consecutive MMAs then accumulate into DIFFERENT diagonals d (no back-to-back dependency on
            // one accumulator) and share the B operand
#pragma unroll
            for (int i = 0; i < S; i++)
#pragma unroll
              for (int d = i; d < S; d++) {
                const uint64_t da = xtc_desc(abase + (d - i) * G::A_SLICE, X_TM * 16, 128);   // U digit j = d - i
                const uint64_t db = xtc_desc(bbase + i * G::B_SLICE, NB * 16, 128);           // k* byte slice i
                if (PAIR) xtc_mma_i8_pair(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);
                else xtc_mma_i8(tmem + d * NQ, da, db, idesc, (c > 0 || i > 0) ? 1u : 0u);
              }
            if (PAIR) xtc_commit_pair(empty + 8 * s, mask);
            else if (C > 1) xtc_commit_mc(empty + 8 * s, mask);
            else xtc_commit(empty + 8 * s);
            if (c == ns - 1) { if (PAIR) xtc_commit_pair(acc_full, mask); else xtc_commit(acc_full); }
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ===================================================== epilogue: 8 warps; lane quarter = warp % 4, column half = (warp - 14) / 4
    const int ew = warp - X_WARP_EPI, lq = warp & 3, half = ew >> 2;
    constexpr int HC = NQ / 2, NGC = HC / 8;                         // columns per half, 8-column groups per half
    double acc[NGC];
#pragma unroll
    for (int i = 0; i < NGC; i++) acc[i] = 0.0;
    const unsigned trow = tmem + ((unsigned)(lq * 32) << 16) + (unsigned)(half * HC);
    for (int r = 0; r < rounds; r++) {
      const int n = (r * C + (int)rank) * X_TM + lq * 32 + lane;
      const double sbn = __ldg(sb + n);
      xmb_wait_relaxed(acc_full, (unsigned)r & 1u);
      xtc_fence_after();
#pragma unroll
      for (int cg = 0; cg < NGC; cg++) {
        int rr[S][8];
#pragma unroll
        for (int d = 0; d < S; d++) xtc_ld8(trow + d * NQ + cg * 8, rr[d]);
        xtc_wait_ld();
        double v[8];
#pragma unroll
        for (int e = 0; e < 8; e++) {
          long long hi = rr[0][e];
#pragma unroll
          for (int d = 1; d < S - 2; d++) hi = (hi << 8) + rr[d][e];
          const long long lo = ((long long)rr[S - 2][e] << 8) + rr[S - 1][e];
          double h = fma((double)hi, 65536.0, (double)lo);
          h *= sbn;
          v[e] = h * h;
        }
        // sum over the warp's 32 rows, 8 columns at a time: butterfly "transpose-reduce" (the same tree for every column)
        const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4;
        double w4[4], w2[2], w1;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const double keep = b16 ? v[4 + i] : v[i], send = b16 ? v[i] : v[4 + i];
          w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const double keep = b8 ? w4[2 + i] : w4[i], send = b8 ? w4[i] : w4[2 + i];
          w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
        {
          const double keep = b4 ? w2[1] : w2[0], send = b4 ? w2[0] : w2[1];
          w1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
        w1 += __shfl_xor_sync(0xffffffffu, w1, 2);
        w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
        acc[cg] += w1;     // lane L holds column cg*8 + 4*bit4 + 2*bit3 + bit2 of its half (replicated over L & 3)
      }
      xtc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR && rank != 0) xmb_arrive_remote(xmapa(acc_empty, 0));
        else xmb_arrive_local(acc_empty);
      }
    }
    // cross-warp: the 4 lane quarters (rows) of each column, in quarter order
    double* red = reinterpret_cast<double*>(smem);                   // [8 warps][HC]: the stage buffers are idle now
    if ((lane & 3) == 0) {
      const int col = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
#pragma unroll
      for (int cg = 0; cg < NGC; cg++) red[ew * HC + cg * 8 + col] = acc[cg];
    }
    asm volatile("bar.sync 1, %0;\n" ::"n"(X_EPI_WARPS * 32) : "memory");
    const int et = ew * 32 + lane;                                   // 0..255
    if (et < NQ) {
      const int hh = et / HC, cc = et % HC;
      const double* rp = red + (hh * 4) * HC + cc;
      const double sum = ((rp[0] + rp[HC]) + rp[2 * HC]) + rp[3 * HC];
      part[et] = sum;
      if (C > 1 && rank != 0) {
        const unsigned raddr = xmapa(xs32(part + NQ + et), 0);
        asm volatile("st.shared::cluster.f64 [%0], %1;\n" ::"r"(raddr), "d"(sum) : "memory");
      }
    }
  }
  xtc_fence_before();
  __syncthreads();
  if (C > 1) xcluster_sync();          // the peer's partial sums have arrived; nobody leaves while a peer may still signal it
  if (rank == 0 && t < NQ) {
    const int64_t q = q0 + t;
    if (q < M) {
      const double xq = Xq[q * 3], yq = Xq[q * 3 + 1], zq = Xq[q * 3 + 2];
      const double sc = x_query_scale<S>(xq, yq, zq, bb, R, R3);
      double ss = part[t];
      if (C > 1) ss += part[NQ + t];
      const double f = 1.0 / (sc * 256.0);   // S_q 256^-(S+1)
      stdv[q] = sqrt(fabs(prior - f * f * ss));   // gpis.py:59
    }
  }
  if (warp == X_WARP_MMA) {
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;\n" ::"r"(tmem));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem));
  }
}

template <int S>
int x_prepare(cdx_ctx* ctx, cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  if (g.i8x_S == S) return CDX_OK;
  const int Np = (int)g.Np, N = (int)g.N;
  const int NT = (Np + X_TM - 1) / X_TM, NTp = (NT + 1) / 2 * 2, rows = NTp * X_TM, NC = Np / X_BK;
  const size_t need = (size_t)NTp * NC * xcfg<S>::A_BYTES;
  if (need > g.i8x_cap) {
    cudaFree(g.i8x_A);
    g.i8x_A = nullptr;
    g.i8x_cap = 0;
    if (cudaMalloc((void**)&g.i8x_A, need) != cudaSuccess) {
      cudaGetLastError();
      cdx_set_error("out of device memory for the INT8 slices of L^-1 (N=%d, %d slices)", N, S);
      return CDX_ERR_ALLOC;
    }
    g.i8x_cap = need;
  }
  if (rows > g.i8x_rows) {
    cudaFree(g.i8x_sb); cudaFree(g.i8x_eb); cudaFree(g.i8x_pts);
    g.i8x_rows = 0;
    CDX_CUDA(cudaMalloc((void**)&g.i8x_sb, sizeof(double) * rows));
    CDX_CUDA(cudaMalloc((void**)&g.i8x_eb, sizeof(int) * rows));
    CDX_CUDA(cudaMalloc((void**)&g.i8x_pts, sizeof(double) * 4 * rows));
    g.i8x_rows = rows;
  }
  if (!g.i8x_bb) CDX_CUDA(cudaMalloc((void**)&g.i8x_bb, sizeof(double) * 6));
  x_rowscale_kernel<<<rows, 256, 0, st>>>(g.U, N, Np, rows, g.i8x_sb, g.i8x_eb);
  CDX_CHECK_LAUNCH(ctx);
  x_slice_kernel<S><<<dim3((Np + 255) / 256, rows), 256, 0, st>>>(g.U, N, Np, rows, g.i8x_eb, g.i8x_A);
  CDX_CHECK_LAUNCH(ctx);
  x_points_kernel<<<1, 256, 0, st>>>(g.Xa, N, Np, g.i8x_pts, g.i8x_bb);
  CDX_CHECK_LAUNCH(ctx);
  g.i8x_S = S;
  if (!g.i8_ready) CDX_CUDA(cudaEventCreateWithFlags(&g.i8_ready, cudaEventDisableTiming));
  CDX_CUDA(cudaEventRecord(g.i8_ready, st));
  return CDX_OK;
}

template <int S, int C, bool PAIR>
int x_launch(cdx_ctx* ctx, const double* X, int64_t M, double* stdv, double* mean, double* gmean, bool* mean_done, cudaStream_t st) {
  using G = xcfg<S, PAIR>;
  constexpr bool PRE = PAIR;           // the CTA-pair path consumes k* slices generated beforehand (see gpis_std_i8x_kernel)
  int rc = x_prepare<S>(ctx, st);
  if (rc) return rc;
  cdx_gpis_state& g = ctx->g;
  if (int rca = cdx_smem_attr(ctx, gpis_std_i8x_kernel<S, C, PAIR, PRE>, G::SMEM)) return rca;
  CDX_CUDA(cudaStreamWaitEvent(st, g.i8_ready, 0));   // the slices may have been built on another stream
  const int NT = (int)((g.Np + X_TM - 1) / X_TM), NC = (int)(g.Np / X_BK);
  // queries are processed in chunks whose k* image fits the scratch buffer (20 KB per query at N = 4096, S = 5)
  const int64_t chunk_blocks = PRE ? 8 * (int64_t)(ctx->sms / C) : (M + G::NQ - 1) / G::NQ;
  const int64_t chunk_q = chunk_blocks * G::NQ;
  const int parts = (int)((g.Np + X_PART - 1) / X_PART);
  const bool with_mean = PRE && (mean || gmean);
  *mean_done = with_mean;
  if (PRE) {
    const int64_t blocks = std::min<int64_t>(chunk_blocks, (M + G::NQ - 1) / G::NQ);
    const size_t need = (size_t)blocks * C * NC * G::B_BYTES;
    if (need > g.i8x_kq_cap) {
      CDX_CUDA(cudaStreamSynchronize(st));
      cudaFree(g.i8x_kq); cudaFree(g.i8x_scq); cudaFree(g.i8x_part);
      g.i8x_kq = nullptr; g.i8x_scq = nullptr; g.i8x_part = nullptr; g.i8x_kq_cap = 0;
      if (cudaMalloc((void**)&g.i8x_kq, need) != cudaSuccess || cudaMalloc((void**)&g.i8x_scq, sizeof(double) * blocks * G::NQ) != cudaSuccess ||
          cudaMalloc((void**)&g.i8x_part, sizeof(double) * 4 * parts * blocks * G::NQ) != cudaSuccess) {
        cudaGetLastError();
        cdx_set_error("out of device memory for the k* slices of %lld queries (%zu bytes)", (long long)(blocks * G::NQ), need);
        return CDX_ERR_ALLOC;
      }
      g.i8x_kq_cap = need;
    }
  }
  for (int64_t q0 = 0; q0 < M; q0 += chunk_q) {
    const int64_t mc = std::min<int64_t>(chunk_q, M - q0);
    const int64_t blocks = (mc + G::NQ - 1) / G::NQ, Mc = blocks * G::NQ;
    if (PRE) {
      i8x_scale_kernel<S><<<(unsigned)((Mc + 255) / 256), 256, 0, st>>>(X + q0 * 3, mc, Mc, g.i8x_bb, g.R, g.R * g.R * g.R, g.i8x_scq);
      CDX_CHECK_LAUNCH(ctx);
      const dim3 kgrid((unsigned)((Mc + 127) / 128), (unsigned)parts);
      if (with_mean)
        i8x_kstar_kernel<S, C, PAIR, true><<<kgrid, 128, 0, st>>>(X + q0 * 3, mc, Mc, g.i8x_pts, g.Xa, g.i8x_scq, (int)g.Np, g.R,
                                                                g.R * g.R * g.R, g.i8x_kq, g.i8x_part);
      else
        i8x_kstar_kernel<S, C, PAIR, false><<<kgrid, 128, 0, st>>>(X + q0 * 3, mc, Mc, g.i8x_pts, g.Xa, g.i8x_scq, (int)g.Np, g.R,
                                                                 g.R * g.R * g.R, g.i8x_kq, g.i8x_part);
      CDX_CHECK_LAUNCH(ctx);
      if (with_mean) {
        i8x_mean_reduce_kernel<<<(unsigned)((mc + 255) / 256), 256, 0, st>>>(g.i8x_part, parts, mc, Mc, g.bias, mean ? mean + q0 : nullptr,
                                                                             gmean ? gmean + q0 * 3 : nullptr);
        CDX_CHECK_LAUNCH(ctx);
      }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(blocks * C), 1, 1);
    cfg.blockDim = dim3(X_THREADS, 1, 1);
    cfg.dynamicSmemBytes = G::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = C > 1 ? 1 : 0;
    CDX_CUDA(cudaLaunchKernelEx(&cfg, gpis_std_i8x_kernel<S, C, PAIR, PRE>, X + q0 * 3, mc, (const double*)g.i8x_pts, (const int8_t*)g.i8x_A,
                                (const double*)g.i8x_sb, (const double*)g.i8x_bb, (int)g.Np, NT, g.R, g.R * g.R * g.R, g.prior,
                                stdv + q0, (const int8_t*)g.i8x_kq));
    CDX_CHECK_LAUNCH(ctx);
  }
  return CDX_OK;
}

}  // namespace

// cluster: 1 = no cluster, 2 = two CTAs sharing the generated k* through distributed shared memory, 3 = a CTA PAIR
// (tcgen05 cta_group::2, M = 256)
// mean / gmean (nullable): computed here only on the CTA-pair path (*mean_done tells the caller)
int cdx_i8x_std(cdx_ctx* ctx, int S, int cluster, const double* X, int64_t M, double* stdv, double* mean, double* gmean,
                bool* mean_done, cudaStream_t st) {
  if (cluster == 1) {
    switch (S) {
      case 4: return x_launch<4, 1, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 5: return x_launch<5, 1, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 6: return x_launch<6, 1, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
    }
  } else if (cluster == 2) {
    switch (S) {
      case 4: return x_launch<4, 2, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 5: return x_launch<5, 2, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 6: return x_launch<6, 2, false>(ctx, X, M, stdv, mean, gmean, mean_done, st);
    }
  } else {
    switch (S) {
      case 4: return x_launch<4, 2, true>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 5: return x_launch<5, 2, true>(ctx, X, M, stdv, mean, gmean, mean_done, st);
      case 6: return x_launch<6, 2, true>(ctx, X, M, stdv, mean, gmean, mean_done, st);
    }
  }
  cdx_set_error("cdx_i8x_std: slices must be 4, 5 or 6");
  return CDX_ERR_ARG;
}

// TPS only (the fixed-point scale uses the cubic's bound); the INT32 accumulators hold 255*128*S*Np < 2^31 up to Np = 8192
bool cdx_i8_eligible(const cdx_ctx* ctx) {
  const cdx_gpis_state& g = ctx->g;
  return g.kernel == CDX_KERNEL_TPS && g.Np >= 128 && g.Np <= 8192;
}

extern "C" int cdx_gpis_set_i8_variant(cdx_ctx* ctx, int variant) {
  if (!ctx || variant < 1 || variant > 3) { cdx_set_error("cdx_gpis_set_i8_variant: 1, 2 or 3"); return CDX_ERR_ARG; }
  ctx->g.i8_variant = variant;
  return CDX_OK;
}

extern "C" int cdx_gpis_set_var_mode(cdx_ctx* ctx, int mode) {
  if (!ctx || !(mode == 0 || mode == -1 || (mode >= 4 && mode <= 6))) {
    cdx_set_error("cdx_gpis_set_var_mode: mode must be 0 (auto), -1 (FP64 DMMA) or 4..6 (INT8 slices)");
    return CDX_ERR_ARG;
  }
  ctx->g.var_mode = mode;
  return CDX_OK;
}

void cdx_i8x_free(cdx_ctx* ctx) {
  cdx_gpis_state& g = ctx->g;
  cudaFree(g.i8x_A); cudaFree(g.i8x_sb); cudaFree(g.i8x_eb); cudaFree(g.i8x_pts); cudaFree(g.i8x_bb);
  cudaFree(g.i8x_kq); cudaFree(g.i8x_scq); cudaFree(g.i8x_part);
  g.i8x_kq = nullptr; g.i8x_scq = nullptr; g.i8x_part = nullptr; g.i8x_kq_cap = 0;
  g.i8x_A = nullptr; g.i8x_sb = nullptr; g.i8x_eb = nullptr; g.i8x_pts = nullptr; g.i8x_bb = nullptr;
  g.i8x_S = 0; g.i8x_cap = 0; g.i8x_rows = 0;
  if (g.i8_ready) { cudaEventDestroy(g.i8_ready); g.i8_ready = nullptr; }
}
