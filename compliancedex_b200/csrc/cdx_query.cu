// GPIS posterior queries (gpis.py:43-59 pred, :63-87 compute_normal, and the autograd gradients the optimizers
// take through pred).  Three kernels:
//  * gpis_mean_kernel      mean + grad_mean only: training set (x,y,z,alpha) staged in shared memory by bulk TMA
//                          (cp.async.bulk + mbarrier), 4 lanes per query, shuffle reduction.   24 N flop + N sqrt / query
//  * gpis_var_kernel<EXT=0> unit query {mean, grad_mean, std}:  v = L^-1 k* as a triangular product with the
//                          precomputed inverse factor U on the FP64 tensor path (DMMA), k* generated on the fly,
//                          |v|^2 reduced in the epilogue; nothing of size N per query ever reaches HBM.   N^2 + 26 N flop
//  * gpis_var_kernel<EXT=1> extended query (+ grad_std): w = E11^-1 k* with the full inverse, epilogue forms
//                          var = k(x,x) - k*.w and grad var = -2 sum_i w_i c_i (x - x_i).                 2 N^2 + 34 N flop
#include <cooperative_groups.h>
#include "cdx_internal.cuh"

namespace cg = cooperative_groups;

namespace {

// queries per CTA (TQ) is a template parameter: 128 for throughput (one 512-thread CTA per SM), 32 when the batch is too
// small to fill the chip with 128-query tiles (the pregrasp loop: a few thousand queries per launch)
constexpr int TK = 16;    // training points per k-chunk
constexpr int LDS = 20;   // shared-memory row stride (doubles): conflict-free for lane -> (row = T/4, k = T%4)
constexpr int NA = 3, NB = 4;  // ring depths: generated k* tiles / cp.async tiles of U or P

template <int KERNEL, int TN, bool EXT, int TQ>
__global__ void __launch_bounds__(TQ * 4, TQ == 128 ? 1 : (TN == 64 ? 4 : 2))
gpis_var_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ Xa, const double* __restrict__ Bmat,
                int N, int Np, cdx_kparams kp, double bias, double prior, double* __restrict__ mean,
                double* __restrict__ gmean, double* __restrict__ stdv, double* __restrict__ gstd, int csplit) {
  constexpr int WN = TN / 4, NBW = WN / 8;
  // 32-query CTAs use a 3-deep ring for the matrix tiles (4 CTAs of 52 KB per SM: the 4 x 128 CTAs of the pregrasp batch are
  // all resident at once); the tile for chunk c+2 is then requested AFTER the barrier of chunk c, when its slot is free
  constexpr int NBq = TQ == 32 ? 3 : NB;
  extern __shared__ __align__(16) double smem[];
  double* As = smem;                          // [NA][TQ*LDS]
  double* Bs = As + NA * TQ * LDS;            // [NBq][TN*LDS]
  double* sxq = Bs + NBq * TN * LDS;          // [TQ][3]
  double* red = sxq + TQ * 3;                 // [4][TQ][4]
  double* tot = red + 4 * TQ * 4;             // [TQ][4] running sum of the per-n-tile partials (fixed order)

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int fr = lane >> 2, fk = lane & 3;
  // row group wm (32 query rows) / column group wn (TN/4 columns) of this warp's accumulators.  The row group is the fast index:
  // a scheduler (warp % 4) then holds all four column groups, whose work differs on a diagonal tile of the unit query (below)
  constexpr int RG = TQ / 32;
  const int wm = warp % RG, wn = warp / RG;
  const int64_t q0 = (int64_t)blockIdx.x * TQ;

  constexpr int NTHR = TQ * 4;
  for (int e = t; e < TQ * 3; e += NTHR) {
    int64_t q = q0 + e / 3;
    sxq[e] = q < M ? Xq[q * 3 + e % 3] : 0.0;
  }
  __syncthreads();
  // generator role: query row 8*warp + lane/4, training points (lane%4) + 4 i of each chunk
  const int qg = warp * 8 + fr;
  const double xq = sxq[qg * 3], yq = sxq[qg * 3 + 1], zq = sxq[qg * 3 + 2];
  double m_acc = 0.0, gx = 0.0, gy = 0.0, gz = 0.0;


  const int NT = (Np + TN - 1) / TN;
  const double2* Xa2 = reinterpret_cast<const double2*>(Xa);

  auto issueB = [&](int n0, int c, int buf) {
    double* dst = Bs + buf * TN * LDS;
    for (int p = t; p < TN * 8; p += NTHR) {
      int row = p >> 3, seg = p & 7;
      double* d = dst + row * LDS + seg * 2;
      if (n0 + row < Np) cdx_cp_async16(d, Bmat + (int64_t)(n0 + row) * Np + c * TK + seg * 2);
      else { d[0] = 0.0; d[1] = 0.0; }
    }
  };
  auto genA = [&](int c, int buf, bool meanpass) {
    double* dst = As + buf * TQ * LDS + qg * LDS + fk;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      int j = c * TK + fk + 4 * i;
      double2 p0 = __ldg(Xa2 + 2 * j), p1 = __ldg(Xa2 + 2 * j + 1);
      double dx = xq - p0.x, dy = yq - p0.y, dz = zq - p1.x;
      double k, cc;
      cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);
      if (j >= N) k = 0.0;
      dst[4 * i] = k;
      if (meanpass) {
        double a = p1.y;  // alpha (0 on padded rows)
        m_acc = fma(a, k, m_acc);
        double ac = a * cc;
        gx = fma(ac, dx, gx); gy = fma(ac, dy, gy); gz = fma(ac, dz, gz);
      }
    }
  };

  // n-tiles are visited in DESCENDING order; with csplit > 1 (a thread-block cluster of csplit == NT CTAs along y) CTA
  // rank r owns the single tile NT-1-r and rank 0 adds the partials in rank order over distributed shared memory, which is
  // the same association as the sequential loop: a query's result does not depend on the variant that computed it.
  const int crank = csplit > 1 ? (int)cg::this_cluster().block_rank() : 0;
  bool first_tile = true;
  for (int nt = NT - 1 - crank; nt >= 0; nt -= csplit) {
    const int n0 = nt * TN;
    double ss[4];          // !EXT: sum v^2 per owned row (this tile)
    double ev[4][4];       // EXT: (k.w, w c dx, w c dy, w c dz) per owned row (this tile)
#pragma unroll
    for (int i = 0; i < 4; i++) { ss[i] = 0.0; ev[i][0] = ev[i][1] = ev[i][2] = ev[i][3] = 0.0; }
    // training rows >= N are padding (k* = 0, identity rows of the factor): whole chunks of them contribute exact zeros
    const int kpad = (N + TK - 1) / TK * TK;
    const int kmax = EXT ? kpad : min(kpad, n0 + TN);
    const int nch = kmax / TK;
    const bool meanpass = (nt == NT - 1);
    double acc[4][NBW][2];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < NBW; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    issueB(n0, 0, 0);
    cdx_cp_async_commit();
    if (nch > 1) issueB(n0, 1, 1);
    cdx_cp_async_commit();
    genA(0, 0, meanpass);
    for (int c = 0; c < nch; c++) {
      if (NBq == NB) {
        if (c + 2 < nch) issueB(n0, c + 2, (c + 2) % NBq);
        cdx_cp_async_commit();
      }
      if (c + 1 < nch) genA(c + 1, (c + 1) % NA, meanpass);
      if (NBq == NB) cdx_cp_async_wait<2>(); else cdx_cp_async_wait<1>();
      __syncthreads();
      if (NBq != NB) {
        if (c + 2 < nch) issueB(n0, c + 2, (c + 2) % NBq);
        cdx_cp_async_commit();
      }
      const double* a_s = As + (c % NA) * TQ * LDS + (wm * 32 + fr) * LDS + fk;
      const double* b_s = Bs + (c % NBq) * TN * LDS + (wn * WN + fr) * LDS + fk;
      // unit query, diagonal tile: U is lower triangular, so a chunk of training points past this warp's last column meets
      // only exact zeros of U - the warp skips its DMMAs (the accumulators would not change)
      if (!EXT && c * TK >= n0 + (wn + 1) * WN) continue;
      if (n0 + wn * WN >= N) continue;         // this warp's columns are all padding (identity rows of the factor meet k* = 0: exact zeros)
#pragma unroll
      for (int ks = 0; ks < TK; ks += 4) {
        double a[4], b[NBW];
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = a_s[i * 8 * LDS + ks];
#pragma unroll
        for (int j = 0; j < NBW; j++) b[j] = b_s[j * 8 * LDS + ks];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < NBW; j++) cdx_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
      }
    }
    // epilogue of this n-tile
    if (!EXT) {
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < NBW; j++) ss[i] = fma(acc[i][j][0], acc[i][j][0], fma(acc[i][j][1], acc[i][j][1], ss[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int r = wm * 32 + i * 8 + fr;
        const double rx = sxq[r * 3], ry = sxq[r * 3 + 1], rz = sxq[r * 3 + 2];
#pragma unroll
        for (int j = 0; j < NBW; j++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            int col = n0 + wn * WN + j * 8 + 2 * fk + e;
            if (col < N) {
              double2 p0 = __ldg(Xa2 + 2 * col), p1 = __ldg(Xa2 + 2 * col + 1);
              double dx = rx - p0.x, dy = ry - p0.y, dz = rz - p1.x;
              double k, cc;
              cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);
              double w = acc[i][j][e];
              ev[i][0] = fma(k, w, ev[i][0]);
              double wc = w * cc;
              ev[i][1] = fma(wc, dx, ev[i][1]); ev[i][2] = fma(wc, dy, ev[i][2]); ev[i][3] = fma(wc, dz, ev[i][3]);
            }
          }
      }
    }
    // this tile's partial: quad reduce, then across the 4 column-warps through shared memory in a fixed order
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int r = wm * 32 + i * 8 + fr;
      if (!EXT) {
        double s = ss[i];
        s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2);
        if (fk == 0) red[(wn * TQ + r) * 4] = s;
      } else {
#pragma unroll
        for (int e = 0; e < 4; e++) {
          double s = ev[i][e];
          s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2);
          if (fk == 0) red[(wn * TQ + r) * 4 + e] = s;
        }
      }
    }
    __syncthreads();  // partials visible; all warps done with the rings before the next n-tile refills them
    if (t < TQ) {
#pragma unroll
      for (int e = 0; e < (EXT ? 4 : 1); e++) {
        const double pv = (red[(0 * TQ + t) * 4 + e] + red[(1 * TQ + t) * 4 + e]) + (red[(2 * TQ + t) * 4 + e] + red[(3 * TQ + t) * 4 + e]);
        tot[t * 4 + e] = first_tile ? pv : tot[t * 4 + e] + pv;
      }
    }
    first_tile = false;
    __syncthreads();  // red is rewritten by the next tile
  }

  // mean / grad_mean (accumulated by the CTA that owns the last n-tile): reduce over the 4 lanes that share a query row
  if (crank == 0) {
    m_acc += __shfl_xor_sync(0xffffffffu, m_acc, 1); m_acc += __shfl_xor_sync(0xffffffffu, m_acc, 2);
    gx += __shfl_xor_sync(0xffffffffu, gx, 1); gx += __shfl_xor_sync(0xffffffffu, gx, 2);
    gy += __shfl_xor_sync(0xffffffffu, gy, 1); gy += __shfl_xor_sync(0xffffffffu, gy, 2);
    gz += __shfl_xor_sync(0xffffffffu, gz, 1); gz += __shfl_xor_sync(0xffffffffu, gz, 2);
    if (fk == 0 && q0 + qg < M) {
      if (mean) mean[q0 + qg] = m_acc + bias;
      if (gmean) { gmean[(q0 + qg) * 3] = gx; gmean[(q0 + qg) * 3 + 1] = gy; gmean[(q0 + qg) * 3 + 2] = gz; }
    }
  }
  if (csplit > 1) cg::this_cluster().sync();   // every rank's `tot` is complete and visible cluster-wide
  if (crank == 0 && t < TQ && q0 + t < M) {
    const int64_t q = q0 + t;
    double v[4];
#pragma unroll
    for (int e = 0; e < (EXT ? 4 : 1); e++) v[e] = tot[t * 4 + e];
    for (int r = 1; r < csplit; r++) {           // rank r holds tile NT-1-r: descending-tile order, as the sequential loop
      const double* rt = cg::this_cluster().map_shared_rank(tot, r);
#pragma unroll
      for (int e = 0; e < (EXT ? 4 : 1); e++) v[e] += rt[t * 4 + e];
    }
    double var = prior - v[0];
    double sd = sqrt(fabs(var));  // gpis.py:59
    if (stdv) stdv[q] = sd;
    if (EXT && gstd) {
      // d sqrt|var| = sign(var) dvar / (2 sd),  dvar = -2 sum_i w_i c_i (x - x_i)
      double f = (var < 0.0 ? 1.0 : -1.0) / sd;
      gstd[q * 3] = f * v[1]; gstd[q * 3 + 1] = f * v[2]; gstd[q * 3 + 2] = f * v[3];
    }
  }
  if (csplit > 1) cg::this_cluster().sync();   // remote shared memory stays alive until rank 0 has read it
}

// ------------------------------------------------------------------------------------------ extended query, small states
// gpis_ext_kernel: {mean, grad mean, std, grad std} for states of up to 448 padded training points - the shape of every
// pregrasp iteration (optimize_pregrasp.py:714-716,733-736: 4 contact points per candidate, N = 196..401).  w = E11^-1 k* is a
// [TQ, Np] x [Np, Np] product.  gpis_var_kernel walks it column tile by column tile and regenerates k* (FP64: distance, sqrt,
// cubic) for every one of the Np/64 column tiles on the pipe the DMMA shares.  Here a CTA of 16 warps keeps the accumulators
// of ALL Np columns of its TQ query rows in registers (8 column groups x 2 row groups: a warp owns Np/8 columns = NBW = Np/64
// DMMA tiles of TQ/2 rows), so the K loop runs once: each 16-point chunk of k* is generated exactly once and
// E11^-1 streams through shared memory exactly once per query tile, in 16-row slabs of Np columns (cp.async ring, 2-4 slabs deep;
// E11^-1 is symmetric, so slab rows are contiguous).  The epilogue evaluates k and c once per (query, column) for
// var = k(x,x) - k*.w and grad var = -2 sum_i w_i c_i (x - x_i): 2 kernel evaluations per (query, training point) instead
// of Np/64 + 1.  CTAs are persistent over query tiles.  A query's arithmetic - the 16 interleaved partial sums of the mean,
// the chunk order of the DMMA accumulation, the fixed column-warp order of the epilogue - does not depend on TQ or on the
// tile it lands in, so results are bit-identical for every batch size and shard.
constexpr int EW = 8;      // column groups (warps along the columns)
constexpr int ERG = 2;     // row groups: 16 warps per CTA - a DMMA holds its warp for several cycles, four warps per scheduler keep the pipe fed
__host__ __device__ constexpr int ext_slab_stages(int nbw) { return nbw <= 4 ? 4 : (nbw <= 5 ? 3 : 2); }   // 227 KB of shared memory per CTA
template <int KERNEL, int NBW, int TQ>
__global__ void __launch_bounds__(EW * ERG * 32, 1)
gpis_ext_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ Xa, const double* __restrict__ Pm,
                int N, int Np, cdx_kparams kp, double bias, double prior, double* __restrict__ mean,
                double* __restrict__ gmean, double* __restrict__ stdv, double* __restrict__ gstd, int64_t ntiles) {
  constexpr int RPG = TQ / ERG;         // query rows per row group
  constexpr int MT = RPG / 8;           // 8-row DMMA blocks per warp
  constexpr int NTHR = EW * ERG * 32;
  constexpr int NBS = ext_slab_stages(NBW);   // slabs of E11^-1 in flight
  extern __shared__ __align__(16) double smem[];
  double* Bs = smem;                          // [NBS][Np * LDS] slab of E11^-1: rows = output columns n, 16 k each
  double* As = Bs + NBS * (size_t)Np * LDS;   // [NA][TQ * LDS]  generated k* chunks
  double* sxq = As + NA * TQ * LDS;           // [TQ][3]
  double* red = sxq + TQ * 3;                 // [EW][TQ][4]
  double* sXa = red + EW * TQ * 4;            // [Np][4] the training records (x, y, z, alpha): resident for the CTA's lifetime
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int fr = lane >> 2, fk = lane & 3;
  const int cg = warp % EW, rg = warp / EW;   // this warp's column group and row group
  const int WN = NBW * 8;                     // columns per warp
  const int kpad = (N + TK - 1) / TK * TK, nch = kpad / TK;
  for (int e = t; e < Np * 2; e += NTHR)
    reinterpret_cast<double2*>(sXa)[e] = __ldg(reinterpret_cast<const double2*>(Xa) + e);
  const double2* Xa2 = reinterpret_cast<const double2*>(sXa);
  // generator role: 16 threads per query row, thread s of them owns training point s of every 16-point chunk and one of the
  // query's 16 interleaved partial sums of the mean - for every TQ (with TQ = 16 the upper half of the CTA has no row)
  const int qg = t >> 4, kslot = t & 15;
  const bool gen = qg < TQ;

  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t q0 = tile * TQ;
    __syncthreads();                          // the previous tile's epilogue has finished with sxq / red / the rings
    for (int e = t; e < TQ * 3; e += NTHR) {
      const int64_t q = q0 + e / 3;
      sxq[e] = q < M ? Xq[q * 3 + e % 3] : 0.0;
    }
    __syncthreads();
    const double xq = gen ? sxq[qg * 3] : 0.0, yq = gen ? sxq[qg * 3 + 1] : 0.0, zq = gen ? sxq[qg * 3 + 2] : 0.0;
    double ma = 0.0, gxa = 0.0, gya = 0.0, gza = 0.0;
    double acc[MT][NBW][2];
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int j = 0; j < NBW; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto issueB = [&](int c, int buf) {
      double* dst = Bs + (size_t)buf * Np * LDS;
      for (int p = t; p < Np * 8; p += NTHR) {
        const int row = p >> 3, seg = p & 7;
        cdx_cp_async16(dst + row * LDS + seg * 2, Pm + (int64_t)row * Np + c * TK + seg * 2);
      }
    };
    auto genA = [&](int c, int buf) {
      if (!gen) return;
      const int j = c * TK + kslot;
      const double2 p0 = Xa2[2 * j], p1 = Xa2[2 * j + 1];
      const double dx = xq - p0.x, dy = yq - p0.y, dz = zq - p1.x;
      double k, cc;
      cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);
      if (j >= N) k = 0.0;
      As[buf * TQ * LDS + qg * LDS + kslot] = k;
      const double a = p1.y;  // alpha (0 on padded rows)
      ma = fma(a, k, ma);
      const double ac = a * cc;
      gxa = fma(ac, dx, gxa); gya = fma(ac, dy, gya); gza = fma(ac, dz, gza);
    };

#pragma unroll
    for (int p = 0; p < NBS - 1; p++) {       // NBS - 1 slabs in flight before the first product
      if (p < nch) issueB(p, p);
      cdx_cp_async_commit();
    }
    genA(0, 0);
    for (int c = 0; c < nch; c++) {
      if (c + 1 < nch) genA(c + 1, (c + 1) % NA);
      cdx_cp_async_wait<NBS - 2>();           // slab c has landed (the NBS - 2 younger groups may still be in flight)
      __syncthreads();                        // chunk c (slab + k*) visible; every warp is done with chunk c - 1
      if (c + NBS - 1 < nch) issueB(c + NBS - 1, (c + NBS - 1) % NBS);   // into the slot chunk c - 1 has just left
      cdx_cp_async_commit();
      const double* a_s = As + (c % NA) * TQ * LDS + (rg * RPG + fr) * LDS + fk;
      const double* b_s = Bs + (size_t)(c % NBS) * Np * LDS + (cg * WN + fr) * LDS + fk;
#pragma unroll
      for (int ks = 0; ks < TK; ks += 4) {
        double a[MT], b[NBW];
#pragma unroll
        for (int i = 0; i < MT; i++) a[i] = a_s[i * 8 * LDS + ks];
#pragma unroll
        for (int j = 0; j < NBW; j++) b[j] = b_s[j * 8 * LDS + ks];
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
          for (int j = 0; j < NBW; j++) cdx_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
      }
    }
    // ---- mean / grad mean: the 16 slot sums of a query, butterfly over its 16 lanes (every lane takes part)
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) {
      ma += __shfl_xor_sync(0xffffffffu, ma, o); gxa += __shfl_xor_sync(0xffffffffu, gxa, o);
      gya += __shfl_xor_sync(0xffffffffu, gya, o); gza += __shfl_xor_sync(0xffffffffu, gza, o);
    }
    if (gen && kslot == 0 && q0 + qg < M) {
      if (mean) mean[q0 + qg] = ma + bias;
      if (gmean) { gmean[(q0 + qg) * 3] = gxa; gmean[(q0 + qg) * 3 + 1] = gya; gmean[(q0 + qg) * 3 + 2] = gza; }
    }
    // ---- epilogue: (k.w, w c dx, w c dy, w c dz) over this warp's columns, per owned row
#pragma unroll
    for (int i = 0; i < MT; i++) {
      const int r = rg * RPG + i * 8 + fr;
      const double rx = sxq[r * 3], ry = sxq[r * 3 + 1], rz = sxq[r * 3 + 2];
      double ev[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int j = 0; j < NBW; j++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int col = cg * WN + j * 8 + 2 * fk + e;
          if (col < N) {
            const double2 p0 = Xa2[2 * col], p1 = Xa2[2 * col + 1];
            const double dx = rx - p0.x, dy = ry - p0.y, dz = rz - p1.x;
            double k, cc;
            cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);
            const double w = acc[i][j][e];
            ev[0] = fma(k, w, ev[0]);
            const double wc = w * cc;
            ev[1] = fma(wc, dx, ev[1]); ev[2] = fma(wc, dy, ev[2]); ev[3] = fma(wc, dz, ev[3]);
          }
        }
#pragma unroll
      for (int e = 0; e < 4; e++) {
        double sres = ev[e];
        sres += __shfl_xor_sync(0xffffffffu, sres, 1); sres += __shfl_xor_sync(0xffffffffu, sres, 2);
        if (fk == 0) red[(cg * TQ + r) * 4 + e] = sres;
      }
    }
    __syncthreads();
    if (t < TQ && q0 + t < M) {
      const int64_t q = q0 + t;
      double v[4];
#pragma unroll
      for (int e = 0; e < 4; e++) {
        double sres = red[(0 * TQ + t) * 4 + e];
#pragma unroll
        for (int w = 1; w < EW; w++) sres += red[(w * TQ + t) * 4 + e];      // column groups in order
        v[e] = sres;
      }
      const double var = prior - v[0];
      const double sd = sqrt(fabs(var));  // gpis.py:59
      if (stdv) stdv[q] = sd;
      if (gstd) {
        // d sqrt|var| = sign(var) dvar / (2 sd),  dvar = -2 sum_i w_i c_i (x - x_i)
        const double f = (var < 0.0 ? 1.0 : -1.0) / sd;
        gstd[q * 3] = f * v[1]; gstd[q * 3 + 1] = f * v[2]; gstd[q * 3 + 2] = f * v[3];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ mean only
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned phase) {
  unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(a), "r"(phase) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx_only(uint64_t* bar, unsigned bytes) {   // adds to the tx-count, does not arrive
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------ extended query, warp-specialised
// gpis_extw_kernel: the all-columns extended query (see gpis_ext_kernel above for the arithmetic) with the roles separated.
// Ablations of gpis_ext_kernel at 131,072 queries on hammer (N = 317): 1,245 us as is, 835 us with the DMMAs removed, 492 us
// with the fragment loads removed as well, and the DMMAs alone are 730 us - the kernel took the SUM of its tensor time and
// of everything else, because all sixteen warps did "everything else" (generate k*, issue 2,560 cp.async per slab, wait,
// __syncthreads) together, chunk after chunk, and the tensor pipe idled meanwhile.  Here
//   * warps 0-15 (consumers) only load fragments and issue DMMAs; they wait on / release mbarriers per ring slot, never
//     the whole CTA; the epilogue uses a named barrier of their 512 threads;
//   * warps 16-19 (generators) produce the k* chunks (and the mean / grad-mean sums) up to D chunks ahead;
//   * the slabs of E11^-1 arrive by ONE bulk-TMA stream (a lane of the first generator warp issues <= 4 cp.async.bulk per
//     slab) from a copy packed at fit time in slab order [chunk][column][16] (cdx_gpis_state::Pslab), whose 16-byte
//     pieces are XOR-swizzled by the column index (piece ^ 2 (column & 3)): rows are 128 B with no padding, and the DMMA
//     fragment pattern - lane (fr, fk) of step s reads k = 4s + fk of column fr, i.e. pieces 2s and 2s + 1 - touches 8
//     distinct pieces per half-warp: conflict-free.  The k order of the accumulation is the one of the other two kernels,
//     so w = E11^-1 k* is the same bit for bit.
// A query's arithmetic (chunk order of the accumulation; column order inside a thread, then the four lanes of a row, then
// the column groups in order in the epilogue) does not depend on TQ or the tile it lands in: results are bit-identical for
// every batch size and shard.
constexpr int XW_GEN = 4;                       // generator warps
constexpr int XW_THREADS = (EW * ERG + XW_GEN) * 32;
constexpr int XW_LA = LDS;                      // k* tile row stride (doubles)
__host__ __device__ constexpr int extw_depth(int nbw) { return nbw <= 5 ? 4 : 3; }
__host__ __device__ constexpr size_t extw_smem(int nbw, int tq) {
  return (size_t)extw_depth(nbw) * (nbw * 64) * 128 + sizeof(double) * ((size_t)extw_depth(nbw) * tq * XW_LA + 4 * (size_t)(nbw * 64) + 2 * tq * 3 +
                                                                       EW * tq * 8) + 8 * (2 * extw_depth(nbw) + 2);
}
template <int KERNEL, int NBW, int TQ>
__global__ void __launch_bounds__(XW_THREADS, 1)
gpis_extw_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ Xa, const double* __restrict__ Pslab,
                 int N, cdx_kparams kp, double bias, double prior, double* __restrict__ mean,
                 double* __restrict__ gmean, double* __restrict__ stdv, double* __restrict__ gstd, int64_t ntiles) {
  constexpr int Np = NBW * 64;
  constexpr int RPG = TQ / ERG, MT = RPG / 8, WN = NBW * 8, D = extw_depth(NBW);
  constexpr unsigned SLAB = Np * 128u;          // bytes
  constexpr int QPG = TQ / 8;                   // queries per generator thread
  extern __shared__ __align__(128) unsigned char xw_smem[];
  double* Bs = reinterpret_cast<double*>(xw_smem);          // [D][Np][16] swizzled
  double* As = Bs + (size_t)D * Np * 16;                    // [D][TQ][XW_LA]
  double* sXa = As + D * TQ * XW_LA;                        // [Np][4]
  double* sxq = sXa + 4 * Np;                               // [2][TQ][3]
  double* red = sxq + 2 * TQ * 3;                           // [EW][TQ][8]
  uint64_t* full = reinterpret_cast<uint64_t*>(red + EW * TQ * 8);   // [D] slab landed (tx bytes) + k* written (XW_GEN arrivals)
  uint64_t* empty = full + D;                                        // [D] the 16 consumer warps have finished with the slot
  uint64_t* epi = empty + D;                                         // [2] the consumers have finished the epilogue of a tile of this parity
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int kpad = (N + TK - 1) / TK * TK, nch = kpad / TK;
  if (t == 0) {
    for (int i = 0; i < D; i++) { mbar_init(full + i, XW_GEN); mbar_init(empty + i, EW * ERG); }
    mbar_init(epi, EW * ERG); mbar_init(epi + 1, EW * ERG);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  for (int e = t; e < Np * 2; e += XW_THREADS)
    reinterpret_cast<double2*>(sXa)[e] = __ldg(reinterpret_cast<const double2*>(Xa) + e);
  __syncthreads();
  const double2* Xa2 = reinterpret_cast<const double2*>(sXa);

  if (warp >= EW * ERG) {
    // ===================================================== generators: k* chunks, slab stream
    // thread (kslot, qb) owns training point kslot of every chunk for the queries qb + 8 i.  Only k* is produced here: the
    // generators share the FP64 pipe with the DMMAs, so their chain is kept short (the mean and its gradient, which need
    // the same k and c, are formed in the consumers' epilogue, where k and c are evaluated per (query, column) anyway)
    // and the QPG evaluations of a chunk are written stage by stage so that their dependent chains overlap.
    const int gt = t - EW * ERG * 32, kslot = gt & 15, qb = gt >> 4;   // queries qb + 8 i
    int64_t gcs = 0;
    int ti = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ti++) {
      const int64_t q0 = tile * TQ;
      const int par = ti & 1;
      mbar_wait(epi + par, ((ti >> 1) & 1) ^ 1);     // the epilogue of tile ti - 2 has finished with sxq[par]
      double xq[QPG], yq[QPG], zq[QPG];
#pragma unroll
      for (int i = 0; i < QPG; i++) {
        const int64_t q = q0 + qb + 8 * i;
        const bool ok = q < M;
        xq[i] = ok ? Xq[q * 3] : 0.0; yq[i] = ok ? Xq[q * 3 + 1] : 0.0; zq[i] = ok ? Xq[q * 3 + 2] : 0.0;
        if (kslot < 3) sxq[par * TQ * 3 + (qb + 8 * i) * 3 + kslot] = kslot == 0 ? xq[i] : (kslot == 1 ? yq[i] : zq[i]);
      }
      // two chunks per pass: 2 QPG independent evaluation chains in flight per thread - the chain (distance, conversion,
      // rsqrt seed, two Newton steps, cubic) is latency-bound, and the generators pace the kernel whenever they fall behind
      for (int c = 0; c < nch; c += 2, gcs += 2) {
        const bool two = c + 1 < nch;
        const int slot0 = (int)(gcs % D), slot1 = (int)((gcs + 1) % D);
        mbar_wait(empty + slot0, (unsigned)(((gcs / D) & 1) ^ 1));
        if (two) mbar_wait(empty + slot1, (unsigned)((((gcs + 1) / D) & 1) ^ 1));
        if (gt == 0) {
          for (int h = 0; h < (two ? 2 : 1); h++) {
            const int slot = h ? slot1 : slot0;
            mbar_expect_tx_only(full + slot, SLAB);
            const char* src = reinterpret_cast<const char*>(Pslab) + (size_t)(c + h) * SLAB;
            char* dst = reinterpret_cast<char*>(Bs) + (size_t)slot * SLAB;
            for (unsigned off = 0; off < SLAB; off += 16384u) tma_bulk_g2s(dst + off, src + off, min(16384u, SLAB - off), full + slot);
          }
        }
        const int j0 = c * TK + kslot, j1 = two ? j0 + TK : j0;
        const double2 p0 = Xa2[2 * j0], p1 = Xa2[2 * j0 + 1], r0 = Xa2[2 * j1], r1 = Xa2[2 * j1 + 1];
        double sq[2 * QPG], kv[2 * QPG];
#pragma unroll
        for (int i = 0; i < QPG; i++) {
          const double dx = xq[i] - p0.x, dy = yq[i] - p0.y, dz = zq[i] - p1.x;
          sq[i] = fma(dx, dx, fma(dy, dy, dz * dz));
          const double ex = xq[i] - r0.x, ey = yq[i] - r0.y, ez = zq[i] - r1.x;
          sq[QPG + i] = fma(ex, ex, fma(ey, ey, ez * ez));
        }
        cdx_kernel_k_batch<KERNEL, 2 * QPG>(sq, kp, kv);
        double* arow0 = As + (size_t)slot0 * TQ * XW_LA + kslot;
#pragma unroll
        for (int i = 0; i < QPG; i++) arow0[(qb + 8 * i) * XW_LA] = j0 >= N ? 0.0 : kv[i];
        __syncwarp();
        if (lane == 0) mbar_arrive(full + slot0);
        if (two) {
          double* arow1 = As + (size_t)slot1 * TQ * XW_LA + kslot;
#pragma unroll
          for (int i = 0; i < QPG; i++) arow1[(qb + 8 * i) * XW_LA] = j1 >= N ? 0.0 : kv[QPG + i];
          __syncwarp();
          if (lane == 0) mbar_arrive(full + slot1);
        }
      }
      if (nch & 1) gcs--;        // the last pass of an odd tile produced one chunk
    }
    return;
  }

  // ======================================================= consumers: fragments + DMMA, epilogue
  const int fr = lane >> 2, fk = lane & 3;
  const int cg = warp % EW, rg = warp / EW;
  const int a_off = (rg * RPG + fr) * XW_LA + fk;           // k = 4 s + fk
  const int b_off = (cg * WN + fr) * 16 + (fk & 1);
  const int xs = fk >> 1, xr = (fr & 3) << 1;               // 16-byte piece of step s: (2 s + xs) ^ xr
  int64_t gcs = 0;
  int ti = 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ti++) {
    const int64_t q0 = tile * TQ;
    const int par = ti & 1;
    double acc[MT][NBW][2];
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int j = 0; j < NBW; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
    for (int c = 0; c < nch; c++, gcs++) {
      const int slot = (int)(gcs % D);
      mbar_wait(full + slot, (unsigned)((gcs / D) & 1));
      const double* a_s = As + (size_t)slot * TQ * XW_LA + a_off;
      const double* b_s = Bs + (size_t)slot * Np * 16 + b_off;
#pragma unroll
      for (int s4 = 0; s4 < 4; s4++) {
        const int xo = ((2 * s4 + xs) ^ xr) * 2;
        double a[MT], b[NBW];
#pragma unroll
        for (int i = 0; i < MT; i++) a[i] = a_s[i * 8 * XW_LA + 4 * s4];
#pragma unroll
        for (int j = 0; j < NBW; j++) b[j] = b_s[j * 128 + xo];
#pragma unroll
        for (int i = 0; i < MT; i++)
#pragma unroll
          for (int j = 0; j < NBW; j++) cdx_dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + slot);
    }
    // ---- epilogue: (k.w, w c dx, w c dy, w c dz) over this warp's columns, per owned row
    const double* sq = sxq + par * TQ * 3;
    asm volatile("bar.sync 1, %0;\n" ::"n"(EW * ERG * 32) : "memory");     // the previous tile's final sums have been read from `red`
#pragma unroll
    for (int i = 0; i < MT; i++) {
      const int r = rg * RPG + i * 8 + fr;
      const double rx = sq[r * 3], ry = sq[r * 3 + 1], rz = sq[r * 3 + 2];
      double ev[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};   // k.w, w c (x - x_j) | alpha.k, alpha c (x - x_j)
#pragma unroll
      for (int j = 0; j < NBW; j++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int col = cg * WN + j * 8 + 2 * fk + e;
          if (col < N) {
            const double2 p0 = Xa2[2 * col], p1 = Xa2[2 * col + 1];
            const double dx = rx - p0.x, dy = ry - p0.y, dz = rz - p1.x;
            double k, cc;
            cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, cc);
            const double w = acc[i][j][e];
            ev[0] = fma(k, w, ev[0]);
            const double wc = w * cc;
            ev[1] = fma(wc, dx, ev[1]); ev[2] = fma(wc, dy, ev[2]); ev[3] = fma(wc, dz, ev[3]);
            ev[4] = fma(p1.y, k, ev[4]);
            const double ac = p1.y * cc;
            ev[5] = fma(ac, dx, ev[5]); ev[6] = fma(ac, dy, ev[6]); ev[7] = fma(ac, dz, ev[7]);
          }
        }
#pragma unroll
      for (int e = 0; e < 8; e++) {
        double sres = ev[e];
        sres += __shfl_xor_sync(0xffffffffu, sres, 1); sres += __shfl_xor_sync(0xffffffffu, sres, 2);
        if (fk == 0) red[(cg * TQ + r) * 8 + e] = sres;
      }
    }
    asm volatile("bar.sync 1, %0;\n" ::"n"(EW * ERG * 32) : "memory");
    if (t < TQ && q0 + t < M) {
      const int64_t q = q0 + t;
      double v[8];
#pragma unroll
      for (int e = 0; e < 8; e++) {
        double sres = red[(0 * TQ + t) * 8 + e];
#pragma unroll
        for (int w = 1; w < EW; w++) sres += red[(w * TQ + t) * 8 + e];      // column groups in order
        v[e] = sres;
      }
      if (mean) mean[q] = v[4] + bias;
      if (gmean) { gmean[q * 3] = v[5]; gmean[q * 3 + 1] = v[6]; gmean[q * 3 + 2] = v[7]; }
      const double var = prior - v[0];
      const double sd = sqrt(fabs(var));  // gpis.py:59
      if (stdv) stdv[q] = sd;
      if (gstd) {
        const double f = (var < 0.0 ? 1.0 : -1.0) / sd;
        gstd[q * 3] = f * v[1]; gstd[q * 3 + 1] = f * v[2]; gstd[q * 3 + 2] = f * v[3];
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(epi + par);      // sxq[par] may be overwritten (tile ti + 2)
  }
}

constexpr int MQ = 64;  // queries per CTA pass (4 lanes each)

template <int KERNEL>
__global__ void __launch_bounds__(256)
gpis_mean_kernel(const double* __restrict__ Xq, int64_t M, const double* __restrict__ Xa, int N, int Np, int tile_pts,
                 cdx_kparams kp, double bias, double* __restrict__ mean, double* __restrict__ gmean) {
  extern __shared__ __align__(16) double smem[];
  double* sX = smem;                                              // [tile_pts][4]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + (size_t)tile_pts * 4);
  const int t = threadIdx.x, jsub = t & 3;
  if (t == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  const int ntt = (Np + tile_pts - 1) / tile_pts;
  const int64_t nqt = (M + MQ - 1) / MQ;
  int staged = -1;
  unsigned phase = 0;
  const double2* s2 = reinterpret_cast<const double2*>(sX);
  for (int64_t qt = blockIdx.x; qt < nqt; qt += gridDim.x) {
    const int64_t q = qt * MQ + (t >> 2);
    const bool ok = q < M;
    const double xq = ok ? Xq[q * 3] : 0.0, yq = ok ? Xq[q * 3 + 1] : 0.0, zq = ok ? Xq[q * 3 + 2] : 0.0;
    double m0 = 0, m1 = 0, gx0 = 0, gy0 = 0, gz0 = 0, gx1 = 0, gy1 = 0, gz1 = 0;
    for (int tt = 0; tt < ntt; tt++) {
      const int p0 = tt * tile_pts, cnt = min(tile_pts, Np - p0);
      if (staged != tt) {
        __syncthreads();  // previous tile no longer being read
        if (t == 0) {
          unsigned bytes = (unsigned)cnt * 32u;
          mbar_expect_tx(bar, bytes);
          for (unsigned off = 0; off < bytes; off += 16384u)
            tma_bulk_g2s(reinterpret_cast<char*>(sX) + off, reinterpret_cast<const char*>(Xa + (size_t)p0 * 4) + off,
                         min(16384u, bytes - off), bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        staged = tt;
      }
      // padded rows carry alpha = 0, so no masking is needed; cnt is a multiple of 64
      for (int j = jsub; j < cnt; j += 8) {
        double2 a0 = s2[2 * j], a1 = s2[2 * j + 1];
        double2 b0 = s2[2 * (j + 4)], b1 = s2[2 * (j + 4) + 1];
        double dx = xq - a0.x, dy = yq - a0.y, dz = zq - a1.x, k, c;
        cdx_kernel_eval<KERNEL>(fma(dx, dx, fma(dy, dy, dz * dz)), kp, k, c);
        m0 = fma(a1.y, k, m0);
        double ac = a1.y * c;
        gx0 = fma(ac, dx, gx0); gy0 = fma(ac, dy, gy0); gz0 = fma(ac, dz, gz0);
        double ex = xq - b0.x, ey = yq - b0.y, ez = zq - b1.x, k2, c2;
        cdx_kernel_eval<KERNEL>(fma(ex, ex, fma(ey, ey, ez * ez)), kp, k2, c2);
        m1 = fma(b1.y, k2, m1);
        double bc = b1.y * c2;
        gx1 = fma(bc, ex, gx1); gy1 = fma(bc, ey, gy1); gz1 = fma(bc, ez, gz1);
      }
    }
    double m = m0 + m1, gx = gx0 + gx1, gy = gy0 + gy1, gz = gz0 + gz1;
    m += __shfl_xor_sync(0xffffffffu, m, 1); m += __shfl_xor_sync(0xffffffffu, m, 2);
    gx += __shfl_xor_sync(0xffffffffu, gx, 1); gx += __shfl_xor_sync(0xffffffffu, gx, 2);
    gy += __shfl_xor_sync(0xffffffffu, gy, 1); gy += __shfl_xor_sync(0xffffffffu, gy, 2);
    gz += __shfl_xor_sync(0xffffffffu, gz, 1); gz += __shfl_xor_sync(0xffffffffu, gz, 2);
    if (ok && jsub == 0) {
      if (mean) mean[q] = m + bias;
      if (gmean) { gmean[q * 3] = gx; gmean[q * 3 + 1] = gy; gmean[q * 3 + 2] = gz; }
    }
  }
}

__global__ void pack_weights_kernel(const double* __restrict__ Xa, const double* __restrict__ w, int N, int Np,
                                    double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Np) return;
  out[4 * i] = Xa[4 * i]; out[4 * i + 1] = Xa[4 * i + 1]; out[4 * i + 2] = Xa[4 * i + 2];
  out[4 * i + 3] = i < N ? w[i] : 0.0;
}

template <int KERNEL>
int launch_mean(cdx_ctx* ctx, const double* Xa, const double* X, int64_t M, double bias, double* mean, double* gmean,
                cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  int tile = (int)min((int64_t)4096, g.Np);
  size_t smem = (size_t)tile * 32 + 16;
  if (int rc = cdx_smem_attr(ctx, gpis_mean_kernel<KERNEL>, 4096 * 32 + 16)) return rc;
  int per_sm = (int)max((size_t)1, min((size_t)8, (size_t)(200 * 1024) / smem));
  int64_t nqt = (M + MQ - 1) / MQ;
  int grid = (int)min(nqt, (int64_t)ctx->sms * per_sm);
  gpis_mean_kernel<KERNEL><<<grid, 256, smem, st>>>(X, M, Xa, (int)g.N, (int)g.Np, tile, cdx_make_kparams(g), bias, mean, gmean);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

template <int KERNEL, int TN, bool EXT, int TQ>
int launch_var(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv, double* gstd,
               cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  size_t smem = sizeof(double) * (size_t)(NA * TQ * LDS + (TQ == 32 ? 3 : NB) * TN * LDS + TQ * 3 + 4 * TQ * 4 + TQ * 4);
  if (int rc = cdx_smem_attr(ctx, gpis_var_kernel<KERNEL, TN, EXT, TQ>, smem)) return rc;
  const int64_t grid = (M + TQ - 1) / TQ;
  const int NT = (int)((g.Np + TN - 1) / TN);
  // small batch and few column tiles: one CTA per (query tile, column tile), the column tiles of a query tile forming a
  // thread-block cluster that reduces over distributed shared memory
  const int csplit = (TQ == 32 && NT >= 2 && NT <= 8 && grid * NT <= 8 * (int64_t)ctx->sms) ? NT : 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, (unsigned)csplit, 1);
  cfg.blockDim = dim3(TQ * 4, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = (unsigned)csplit;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = csplit > 1 ? 1 : 0;
  const double* Bm = EXT ? g.P : g.U;
  CDX_CUDA(cudaLaunchKernelEx(&cfg, gpis_var_kernel<KERNEL, TN, EXT, TQ>, X, M, (const double*)g.Xa, Bm, (int)g.N, (int)g.Np,
                              cdx_make_kparams(g), g.bias, g.prior, mean, gmean, stdv, gstd, csplit));
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}


template <int KERNEL, int NBW, int TQ>
int launch_ext(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv, double* gstd,
               cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  const size_t smem = sizeof(double) * (ext_slab_stages(NBW) * (size_t)g.Np * LDS + NA * TQ * LDS + TQ * 3 + EW * TQ * 4 + 4 * (size_t)g.Np);
  if (int rc = cdx_smem_attr(ctx, gpis_ext_kernel<KERNEL, NBW, TQ>, smem)) return rc;
  const int64_t ntiles = (M + TQ - 1) / TQ;
  const int grid = (int)min(ntiles, (int64_t)ctx->sms);
  gpis_ext_kernel<KERNEL, NBW, TQ><<<grid, EW * ERG * 32, smem, st>>>(X, M, g.Xa, g.P, (int)g.N, (int)g.Np, cdx_make_kparams(g), g.bias,
                                                               g.prior, mean, gmean, stdv, gstd, ntiles);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

template <int KERNEL, int NBW, int TQ>
int launch_extw(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv, double* gstd,
                cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  const size_t smem = extw_smem(NBW, TQ);
  if (int rc = cdx_smem_attr(ctx, gpis_extw_kernel<KERNEL, NBW, TQ>, smem)) return rc;
  const int64_t ntiles = (M + TQ - 1) / TQ;
  const int grid = (int)min(ntiles, (int64_t)ctx->sms);
  gpis_extw_kernel<KERNEL, NBW, TQ><<<grid, XW_THREADS, smem, st>>>(X, M, g.Xa, g.Pslab, (int)g.N, cdx_make_kparams(g), g.bias, g.prior,
                                                                    mean, gmean, stdv, gstd, ntiles);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

// the all-columns extended kernel covers states of up to 448 padded points (7 DMMA column tiles per warp); 16-query tiles
// while 32-query tiles would leave SMs idle
template <int KERNEL>
int dispatch_ext(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv, double* gstd,
                 cudaStream_t st) {
  const int nbw = (int)(ctx->g.Np / 64);
  // tile size by rounds x cost of a round: a persistent CTA walks ceil(tiles / SMs) tiles, and a 16-query tile costs ~0.6 of a
  // 32-query one inside the pregrasp iteration (it streams E11^-1 twice per 32 queries; alone it is cheaper - hammer, 16,384
  // queries: 140 us in 7 rounds of 16 against 151 us in 4 rounds of 32 - but beside the other candidate group's kernels the
  // 8,192-candidate iteration measured 0.397 ms with 16-query tiles and 0.384 ms with 32).  So 16-query tiles only where they
  // save rounds outright: batches below one round of 32-query tiles.  Results do not depend on the choice (bit-identical).
  const int64_t r32 = ((M + 31) / 32 + ctx->sms - 1) / ctx->sms, r16 = ((M + 15) / 16 + ctx->sms - 1) / ctx->sms;
  const bool small = (double)r16 * 0.6 < (double)r32;
  if (ctx->g.ext_mode == 0 && ctx->g.Pslab) {      // warp-specialised kernel (bulk-TMA slab stream)
#define CDX_EXTW(NB) case NB: return small ? launch_extw<KERNEL, NB, 16>(ctx, X, M, mean, gmean, stdv, gstd, st) \
                                           : launch_extw<KERNEL, NB, 32>(ctx, X, M, mean, gmean, stdv, gstd, st)
    switch (nbw) {
      CDX_EXTW(1); CDX_EXTW(2); CDX_EXTW(3); CDX_EXTW(4); CDX_EXTW(5); CDX_EXTW(6); CDX_EXTW(7);
    }
#undef CDX_EXTW
  }
  const bool small_single = (M + 31) / 32 < (int64_t)ctx->sms;     // the single-role kernel's 16-query tiles only pay below one round
#define CDX_EXT(NB) case NB: return small_single ? launch_ext<KERNEL, NB, 16>(ctx, X, M, mean, gmean, stdv, gstd, st) \
                                          : launch_ext<KERNEL, NB, 32>(ctx, X, M, mean, gmean, stdv, gstd, st)
  switch (nbw) {
    CDX_EXT(1); CDX_EXT(2); CDX_EXT(3); CDX_EXT(4); CDX_EXT(5); CDX_EXT(6); CDX_EXT(7);
  }
#undef CDX_EXT
  return -1;
}

template <int KERNEL>
int query_dispatch(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv, double* gstd,
                   cudaStream_t st) {
  cdx_gpis_state& g = ctx->g;
  const bool big = (g.Np % 128 == 0) && g.Np >= 512;
  // fewer 128-query tiles than SMs: spread the batch over the chip with 32-query CTAs (3 resident per SM)
  const bool small_batch = (M + 127) / 128 < (int64_t)ctx->sms;
  // the column tile TN depends on the state only, never on the batch size: a query's result is then bit-identical
  // whichever TQ variant (or shard) evaluates it
  if (gstd) {
    if (g.Np <= 448 && g.ext_mode != 1) return dispatch_ext<KERNEL>(ctx, X, M, mean, gmean, stdv, gstd, st);
    if (small_batch) return big ? launch_var<KERNEL, 128, true, 32>(ctx, X, M, mean, gmean, stdv, gstd, st)
                                : launch_var<KERNEL, 64, true, 32>(ctx, X, M, mean, gmean, stdv, gstd, st);
    return big ? launch_var<KERNEL, 128, true, 128>(ctx, X, M, mean, gmean, stdv, gstd, st)
               : launch_var<KERNEL, 64, true, 128>(ctx, X, M, mean, gmean, stdv, gstd, st);
  }
  if (stdv) {
    // large TPS states: std on the INT8 tensor cores (tcgen05, cdx_query_i8x.cu), mean / grad_mean by the
    // TMA-staged mean kernel.  auto = 5 slices (|d var| < 1e-8 prior measured; the gate is 1e-6 prior).
    const int vm = g.var_mode;
    // the choice depends on the STATE only, never on the batch: a query's result is bit-identical for every chunking / shard
    const bool i8 = vm != -1 && cdx_i8_eligible(ctx) && (vm >= 4 || g.Np >= 1024);
    if (i8) {
      const int S = vm >= 4 ? vm : 5;
      bool mean_done = false;
      int rc = cdx_i8x_std(ctx, S, g.i8_variant == 1 ? 2 : (g.i8_variant == 3 ? 3 : 1), X, M, stdv, mean, gmean, &mean_done, st);
      if (rc) return rc;
      if ((mean || gmean) && !mean_done) return launch_mean<KERNEL>(ctx, g.Xa, X, M, g.bias, mean, gmean, st);
      return CDX_OK;
    }
    // unit query: 128-column tiles whenever they divide the padded size.  Tile nt regenerates the k* chunks 0 .. n0 + TN, so
    // halving the tile count takes ~40 % off the FP64 generation that shares the pipe with the DMMAs, and the per-warp skips
    // of a diagonal tile keep the tensor work at that of 64-column tiles (262,144 queries: banana N = 361 2.04 -> 1.80 ms,
    // mug N = 196 0.92 -> 0.81 ms; 4,608 queries on mug 42 -> 29 us).  With a partial last tile (Np = 320, 448) large batches
    // gain 4-6 % as well but the clustered small-batch launch loses 10-18 % (fewer, longer CTAs: 1.5 waves), and the tile must
    // depend on the state alone - so those states keep 64.
    const bool ubig = g.Np % 128 == 0;
    if (small_batch) return ubig ? launch_var<KERNEL, 128, false, 32>(ctx, X, M, mean, gmean, stdv, nullptr, st)
                                 : launch_var<KERNEL, 64, false, 32>(ctx, X, M, mean, gmean, stdv, nullptr, st);
    return ubig ? launch_var<KERNEL, 128, false, 128>(ctx, X, M, mean, gmean, stdv, nullptr, st)
                : launch_var<KERNEL, 64, false, 128>(ctx, X, M, mean, gmean, stdv, nullptr, st);
  }
  return launch_mean<KERNEL>(ctx, g.Xa, X, M, g.bias, mean, gmean, st);
}

}  // namespace

extern "C" int cdx_gpis_query(cdx_ctx* ctx, const double* X, int64_t M, double* mean, double* gmean, double* stdv,
                              double* gstd, void* stream) {
  CDX_NVTX("cdx_gpis_query");
  if (!ctx || !ctx->g.ready) { cdx_set_error("cdx_gpis_query: no fitted state"); return CDX_ERR_STATE; }
  if (M < 0 || (M > 0 && !X)) { cdx_set_error("cdx_gpis_query: bad argument"); return CDX_ERR_ARG; }
  if (M == 0 || (!mean && !gmean && !stdv && !gstd)) return CDX_OK;
  if (M > (int64_t)2147483647 * 64) { cdx_set_error("cdx_gpis_query: M too large"); return CDX_ERR_ARG; }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  switch (ctx->g.kernel) {
    case CDX_KERNEL_TPS: return query_dispatch<CDX_KERNEL_TPS>(ctx, X, M, mean, gmean, stdv, gstd, st);
    case CDX_KERNEL_RBF: return query_dispatch<CDX_KERNEL_RBF>(ctx, X, M, mean, gmean, stdv, gstd, st);
    default: return query_dispatch<CDX_KERNEL_JOINT>(ctx, X, M, mean, gmean, stdv, gstd, st);
  }
}

extern "C" int cdx_gpis_set_ext_mode(cdx_ctx* ctx, int mode) {
  if (!ctx || mode < 0 || mode > 2) { cdx_set_error("cdx_gpis_set_ext_mode: 0, 1 or 2"); return CDX_ERR_ARG; }
  ctx->g.ext_mode = mode;
  return CDX_OK;
}

extern "C" int cdx_gpis_query_weights(cdx_ctx* ctx, const double* alpha, const double* X, int64_t M, double* mean,
                                      double* gmean, void* stream) {
  CDX_NVTX("cdx_gpis_query_weights");
  if (!ctx || !ctx->g.ready) { cdx_set_error("cdx_gpis_query_weights: no fitted state"); return CDX_ERR_STATE; }
  if (!alpha || M < 0 || (M > 0 && !X)) { cdx_set_error("cdx_gpis_query_weights: bad argument"); return CDX_ERR_ARG; }
  if (M == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  cdx_gpis_state& g = ctx->g;
  double* Xw = nullptr;  // (x,y,z,w) copy of the training set with the caller's weights
  CDX_CUDA(cudaMallocAsync((void**)&Xw, sizeof(double) * 4 * g.Np, st));
  pack_weights_kernel<<<((int)g.Np + 255) / 256, 256, 0, st>>>(g.Xa, alpha, (int)g.N, (int)g.Np, Xw);
  CDX_CHECK_LAUNCH(ctx);
  int rc;
  switch (g.kernel) {
    case CDX_KERNEL_TPS: rc = launch_mean<CDX_KERNEL_TPS>(ctx, Xw, X, M, 0.0, mean, gmean, st); break;
    case CDX_KERNEL_RBF: rc = launch_mean<CDX_KERNEL_RBF>(ctx, Xw, X, M, 0.0, mean, gmean, st); break;
    default: rc = launch_mean<CDX_KERNEL_JOINT>(ctx, Xw, X, M, 0.0, mean, gmean, st); break;
  }
  cudaFreeAsync(Xw, st);
  return rc;
}

extern "C" int cdx_gpis_query_host(cdx_ctx* ctx, const double* Xh, int64_t M, double* mean_h, double* gmean_h,
                                   double* std_h, double* gstd_h, void* stream) {
  CDX_NVTX("cdx_gpis_query_host");
  if (!ctx || !ctx->g.ready) { cdx_set_error("cdx_gpis_query_host: no fitted state"); return CDX_ERR_STATE; }
  if (M < 0 || (M > 0 && !Xh)) { cdx_set_error("cdx_gpis_query_host: bad argument"); return CDX_ERR_ARG; }
  if (M == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  if (M > ctx->hcap) {
    cudaFree(ctx->hx); cudaFree(ctx->hout);
    ctx->hcap = 0;
    CDX_CUDA(cudaMalloc(&ctx->hx, sizeof(double) * 3 * M));
    CDX_CUDA(cudaMalloc(&ctx->hout, sizeof(double) * 8 * M));
    ctx->hcap = M;
  }
  double* dm = mean_h ? ctx->hout : nullptr;
  double* ds = std_h ? ctx->hout + M : nullptr;
  double* dgm = gmean_h ? ctx->hout + 2 * M : nullptr;
  double* dgs = gstd_h ? ctx->hout + 5 * M : nullptr;
  CDX_CUDA(cudaMemcpyAsync(ctx->hx, Xh, sizeof(double) * 3 * M, cudaMemcpyHostToDevice, st));
  int rc = cdx_gpis_query(ctx, ctx->hx, M, dm, dgm, ds, dgs, stream);
  if (rc) return rc;
  if (dm) CDX_CUDA(cudaMemcpyAsync(mean_h, dm, sizeof(double) * M, cudaMemcpyDeviceToHost, st));
  if (ds) CDX_CUDA(cudaMemcpyAsync(std_h, ds, sizeof(double) * M, cudaMemcpyDeviceToHost, st));
  if (dgm) CDX_CUDA(cudaMemcpyAsync(gmean_h, dgm, sizeof(double) * 3 * M, cudaMemcpyDeviceToHost, st));
  if (dgs) CDX_CUDA(cudaMemcpyAsync(gstd_h, dgs, sizeof(double) * 3 * M, cudaMemcpyDeviceToHost, st));
  CDX_CUDA(cudaStreamSynchronize(st));
  return CDX_OK;
}
