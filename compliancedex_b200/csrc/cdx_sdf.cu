// Point-to-triangle-mesh distance for sm_100a: the drop-in for TorchSDF's compute_sdf
// (thirdparty/TorchSDF/torchsdf/sdf.py:34-64, kernels csrc/unbatched_triangle_distance_cuda.cu:176-270).
// Per point: squared distance to the closest triangle, sign (+1 on the side the face normal points to), unit vector from the
// closest point to the point, the closest point; backward: grad_points = 2 grad_dist (p - closest).
//
// The reference runs one thread per point over all F faces and rebuilds every face's edges, normal, edge normals and squared
// edge lengths for every (point, face) pair.  Here:
//   sdf_face_kernel   one thread per face builds a 160-byte record once per call (vertices, edges, squared edge lengths,
//                     in-plane edge normals, unit normal) - the quantities the reference recomputes P times;
//   sdf_pair_kernel   grid = point tiles x face chunks, so that a small batch (4 B fingertips) still fills 148 SMs; records
//                     stream through a 2-stage shared-memory ring with bulk TMA (cp.async.bulk + mbarrier), every thread
//                     keeps PPT points in registers and tracks (float distance, face index) only; chunks combine with one
//                     64-bit atomicMin on (distance bits << 32 | face index) = the reference's "first face with the
//                     smallest distance" rule (strict < inside a tile and between tiles, :236,:243);
//   sdf_final_kernel  one thread per point re-evaluates the winning face with the same device function and writes the four
//                     outputs.
//   sdf_tile_kernel + sdf_cull_kernel  (large batches) bounding-sphere culled search instead of sdf_pair_kernel: same face,
//                     bit for bit, visiting only the tiles that can still hold it (see the comment above sdf_tile_kernel).
// The arithmetic follows the reference expression by expression, including its quirks: the edge parameter is truncated to
// float (point_at takes `float t`, :170) and so is the squared distance (`float dist`, :235), also for float64 inputs.
// Bound: FP32 (FP64 for double inputs) FMA pipe - ~1e2 instructions per pair against 12 bytes per point.
#include <math.h>
#include "cdx_internal.cuh"

namespace {

constexpr int SDF_REC = 40;     // scalars per face record (10 x vec4)
constexpr int SDF_FT = 64;      // faces per shared-memory stage
constexpr int SDF_THREADS = 256;

__device__ __forceinline__ void sdf_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void sdf_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sdf_mbar_wait(uint64_t* bar, unsigned phase) {
  unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(a), "r"(phase) : "memory");
}
__device__ __forceinline__ void sdf_bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

template <typename T> struct v3 { T x, y, z; };
template <typename T> __device__ __forceinline__ v3<T> vsub(v3<T> a, v3<T> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename T> __device__ __forceinline__ T vdot(v3<T> a, v3<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __device__ __forceinline__ v3<T> vcross(v3<T> a, v3<T> b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// |v|^2 with the contraction nvcc gives the reference's dot(dist_vec, dist_vec) (:233,:235): fma(z, z, fma(y, y, x * x)) - measured
// against the reference kernel on B200 (the golden squared distances are reproduced bit for bit by this pattern only)
__device__ __forceinline__ float vnorm2(v3<float> a) { return __fmaf_rn(a.z, a.z, __fmaf_rn(a.y, a.y, __fmul_rn(a.x, a.x))); }
__device__ __forceinline__ double vnorm2(v3<double> a) { return __fma_rn(a.z, a.z, __fma_rn(a.y, a.y, __dmul_rn(a.x, a.x))); }
__device__ __forceinline__ float sdf_rsqrt(float x) { return rsqrtf(x); }
__device__ __forceinline__ double sdf_rsqrt(double x) { return rsqrt(x); }

// record layout (scalars): 0 v1 | 3 len12 | 4 v2 | 7 len23 | 8 v3 | 11 len31 | 12 e12 | 16 e23 | 20 e31 | 24 en12 | 28 en23 |
// 32 en31 | 36 unit normal | 15,19,23 raw normal x,y,z (the sign test uses the unnormalised normal, :234)
template <typename T>
__global__ void __launch_bounds__(128) sdf_face_kernel(const T* __restrict__ fv, int64_t F, int64_t Fpad, T* __restrict__ rec) {
  const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= Fpad) return;
  T* r = rec + f * SDF_REC;
  if (f >= F) {   // padding faces of the last tile: never selected (the pair kernel bounds its loop by F)
    for (int i = 0; i < SDF_REC; i++) r[i] = (T)0;
    return;
  }
  const T* s = fv + f * 9;
  const v3<T> a = {s[0], s[1], s[2]}, b = {s[3], s[4], s[5]}, c = {s[6], s[7], s[8]};
  const v3<T> e12 = vsub(b, a), e23 = vsub(c, b), e31 = vsub(a, c);          // :196-198
  const v3<T> n = vcross(vsub(a, b), e31);                                   // :199
  const v3<T> n12 = vcross(n, e12), n23 = vcross(n, e23), n31 = vcross(n, e31);   // is_not_above, :163-166
  const T inv = sdf_rsqrt(vdot(n, n));                                       // project_plane, :140
  r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = vdot(e12, e12);
  r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = vdot(e23, e23);
  r[8] = c.x; r[9] = c.y; r[10] = c.z; r[11] = vdot(e31, e31);
  r[12] = e12.x; r[13] = e12.y; r[14] = e12.z; r[15] = n.x;
  r[16] = e23.x; r[17] = e23.y; r[18] = e23.z; r[19] = n.y;
  r[20] = e31.x; r[21] = e31.y; r[22] = e31.z; r[23] = n.z;
  r[24] = n12.x; r[25] = n12.y; r[26] = n12.z; r[27] = (T)0;
  r[28] = n23.x; r[29] = n23.y; r[30] = n23.z; r[31] = (T)0;
  r[32] = n31.x; r[33] = n31.y; r[34] = n31.z; r[35] = (T)0;
  r[36] = n.x * inv; r[37] = n.y * inv; r[38] = n.z * inv; r[39] = (T)0;
}

template <typename T> struct q4 { T x, y, z, w; };
__device__ __forceinline__ q4<float> ld4(const float* p) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  return {v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ q4<double> ld4(const double* p) {
  const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
  return {a.x, a.y, b.x, b.y};
}

// closest point of one face to p: the region tests of unbatched_triangle_distance_cuda.cu:192-231 in the reference's order.
// r = the face's record (shared or global memory).  Returns |p - closest|^2 rounded to float (:235).
//   The reference divides first (u = dot(p - v, e) / |e|^2, project_edge :128-134) and compares u with 0 and 1.  For a positive
//   finite |e|^2 the rounded quotient satisfies u > 1 <=> dot > |e|^2, u < 0 <=> dot < 0, u <= 1 <=> dot <= |e|^2 and
//   u >= 0 <=> dot >= 0 (dot > len implies dot/len >= 1 + ulp(len)/len > 1 + 2^-24, above the rounding midpoint), so the
//   comparisons are made on the dot products and only the ONE quotient that becomes an edge parameter is computed - the same
//   IEEE division, hence the same bits - instead of three per pair.  (A branch-free select of the region was measured 10 %
//   slower at 1 M points: the early vertex exits skip more work than divergence costs.)
template <typename T>
__device__ __forceinline__ float sdf_closest(const v3<T> p, const T* __restrict__ r, v3<T>& cp) {
  const q4<T> A = ld4(r), C = ld4(r + 8), E12 = ld4(r + 12), E31 = ld4(r + 20);
  const v3<T> p1 = {p.x - A.x, p.y - A.y, p.z - A.z}, p3 = {p.x - C.x, p.y - C.y, p.z - C.z};
  const T dab = p1.x * E12.x + p1.y * E12.y + p1.z * E12.z;
  const T dca = p3.x * E31.x + p3.y * E31.y + p3.z * E31.z;
  if (dca > C.w && dab < (T)0) {                                       // :203
    cp = {A.x, A.y, A.z};
  } else {
    const q4<T> B = ld4(r + 4), E23 = ld4(r + 16);
    const v3<T> p2 = {p.x - B.x, p.y - B.y, p.z - B.z};
    const T dbc = p2.x * E23.x + p2.y * E23.y + p2.z * E23.z;
    if (dab > A.w && dbc < (T)0) {                                     // :207
      cp = {B.x, B.y, B.z};
    } else if (dbc > B.w && dca < (T)0) {                              // :209
      cp = {C.x, C.y, C.z};
    } else {
      const q4<T> N12 = ld4(r + 24), N23 = ld4(r + 28), N31 = ld4(r + 32);
      if (dab <= A.w && dab >= (T)0 && (N12.x * p1.x + N12.y * p1.y + N12.z * p1.z) <= (T)0) {          // :212
        const float t = (float)(dab / A.w);                            // point_at(vertex, edge, float t), :170
        cp = {A.x + E12.x * t, A.y + E12.y * t, A.z + E12.z * t};
      } else if (dbc <= B.w && dbc >= (T)0 && (N23.x * p2.x + N23.y * p2.y + N23.z * p2.z) <= (T)0) {   // :214
        const float t = (float)(dbc / B.w);
        cp = {B.x + E23.x * t, B.y + E23.y * t, B.z + E23.z * t};
      } else if (dca <= C.w && dca >= (T)0 && (N31.x * p3.x + N31.y * p3.y + N31.z * p3.z) <= (T)0) {   // :216
        const float t = (float)(dca / C.w);
        cp = {C.x + E31.x * t, C.y + E31.y * t, C.z + E31.z * t};
      } else {                                                         // project_plane(v1, normal, p), :137-146
        const q4<T> U = ld4(r + 36);
        const T d = p1.x * U.x + p1.y * U.y + p1.z * U.z;
        cp = {p.x - U.x * d, p.y - U.y * d, p.z - U.z * d};
      }
    }
  }
  const v3<T> dv = vsub(p, cp);
  return (float)vnorm2(dv);
}

// keys[p] = min over faces of (float bits of the squared distance << 32 | face index); distances are >= 0, so their bit
// patterns order like the values and the packed minimum is the lowest-index face among the closest ones
template <typename T, int PPT>
__global__ void __launch_bounds__(SDF_THREADS)
sdf_pair_kernel(const T* __restrict__ points, int64_t P, const T* __restrict__ rec, int64_t F, int tiles_per_chunk,
                unsigned long long* __restrict__ keys) {
  __shared__ __align__(128) T stage[2][SDF_FT * SDF_REC];
  __shared__ uint64_t bar[2];
  const int t = threadIdx.x;
  const int64_t ntiles = (F + SDF_FT - 1) / SDF_FT;
  const int64_t tile0 = (int64_t)blockIdx.y * tiles_per_chunk;
  const int nt = (int)min((int64_t)tiles_per_chunk, ntiles - tile0);
  if (nt <= 0) return;
  if (t == 0) {
    sdf_mbar_init(&bar[0], 1);
    sdf_mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  constexpr unsigned TILE_BYTES = SDF_FT * SDF_REC * sizeof(T);
  auto issue = [&](int i) {   // records are padded to whole tiles, so every stage is one full bulk copy
    sdf_mbar_expect_tx(&bar[i & 1], TILE_BYTES);
    sdf_bulk_g2s(stage[i & 1], rec + (tile0 + i) * (int64_t)(SDF_FT * SDF_REC), TILE_BYTES, &bar[i & 1]);
  };
  if (t == 0) {
    issue(0);
    if (nt > 1) issue(1);
  }
  v3<T> p[PPT];
  float best[PPT];
  unsigned bidx[PPT];
  const int64_t base = (int64_t)blockIdx.x * (SDF_THREADS * PPT) + t;
#pragma unroll
  for (int j = 0; j < PPT; j++) {
    const int64_t i = min(base + (int64_t)j * SDF_THREADS, P - 1);     // tail threads redo the last point (no store below)
    p[j] = {points[i * 3], points[i * 3 + 1], points[i * 3 + 2]};
    best[j] = INFINITY;
    bidx[j] = 0xFFFFFFFFu;
  }
  for (int i = 0; i < nt; i++) {
    sdf_mbar_wait(&bar[i & 1], (unsigned)((i >> 1) & 1));
    const T* s = stage[i & 1];
    const int64_t f0 = (tile0 + i) * SDF_FT;
    const int cnt = (int)min((int64_t)SDF_FT, F - f0);
    for (int f = 0; f < cnt; f++) {
#pragma unroll
      for (int j = 0; j < PPT; j++) {
        v3<T> cp;
        const float d = sdf_closest<T>(p[j], s + f * SDF_REC, cp);
        if (bidx[j] == 0xFFFFFFFFu || best[j] > d) { best[j] = d; bidx[j] = (unsigned)(f0 + f); }   // :236 (first face always taken)
      }
    }
    __syncthreads();                                   // everyone is done with this stage before it is refilled
    if (t == 0 && i + 2 < nt) issue(i + 2);
  }
#pragma unroll
  for (int j = 0; j < PPT; j++) {
    const int64_t i = base + (int64_t)j * SDF_THREADS;
    if (i < P) atomicMin(&keys[i], ((unsigned long long)__float_as_uint(best[j]) << 32) | bidx[j]);
  }
}

// ---- culled search -----------------------------------------------------------------------------------------------------
// Per 64-face tile a bounding sphere (float centre, radius rounded up).  A point needs a tile only if the sphere's lower bound
// on the distance can still beat (or tie) its current best; the first best is the tightest UPPER bound min_t (|p - c_t| + r_t).
// The surviving (point, tile) pairs of a warp are evaluated cooperatively - the point is broadcast, every lane takes two of the
// tile's faces, a 64-bit shuffle-min picks (distance, lowest face index) - so no coherence between the points of a warp is
// needed; when most lanes of a warp need the same tile each lane scans it for its own point instead (never worse than the
// exhaustive kernel).  Bounds are conservative (relative 1e-5 and absolute 1e-6 x coordinate-scale slack), faces that tie with
// the minimum are never skipped and the packed (distance, index) minimum is order independent, so the result is the same
// face, bit for bit, as the exhaustive search and as the reference.
constexpr int SDF_BCHUNK = 2048;   // tile bounds staged per pass (32 KB)

template <typename T>
__global__ void __launch_bounds__(32) sdf_tile_kernel(const T* __restrict__ rec, int64_t F, float4* __restrict__ bounds) {
  const int64_t t = blockIdx.x;
  const int lane = threadIdx.x;
  const int64_t f0 = t * SDF_FT;
  const int cnt = (int)min((int64_t)SDF_FT, F - f0);
  double sx = 0.0, sy = 0.0, sz = 0.0;
  for (int f = lane; f < cnt; f += 32) {
    const T* r = rec + (f0 + f) * SDF_REC;
    sx += (double)r[0] + (double)r[4] + (double)r[8];
    sy += (double)r[1] + (double)r[5] + (double)r[9];
    sz += (double)r[2] + (double)r[6] + (double)r[10];
  }
  for (int o = 16; o; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o); sz += __shfl_xor_sync(0xffffffffu, sz, o);
  }
  const float cx = (float)(sx / (3.0 * cnt)), cy = (float)(sy / (3.0 * cnt)), cz = (float)(sz / (3.0 * cnt));
  double r2 = 0.0;
  for (int f = lane; f < cnt; f += 32) {
    const T* r = rec + (f0 + f) * SDF_REC;
    for (int v = 0; v < 3; v++) {
      const double dx = (double)r[4 * v] - (double)cx, dy = (double)r[4 * v + 1] - (double)cy, dz = (double)r[4 * v + 2] - (double)cz;
      r2 = fmax(r2, dx * dx + dy * dy + dz * dz);
    }
  }
  for (int o = 16; o; o >>= 1) r2 = fmax(r2, __shfl_xor_sync(0xffffffffu, r2, o));
  // NaN / inf vertices: an infinite radius makes every point visit the tile
  float rad = (float)(sqrt(r2) * (1.0 + 1e-6)) * (1.f + 1e-6f) + 1e-37f;
  if (!(rad < INFINITY)) rad = INFINITY;
  if (lane == 0) bounds[t] = make_float4(cx, cy, cz, rad);
}

__device__ __forceinline__ unsigned long long sdf_pack(float d, unsigned f) {
  return ((unsigned long long)__float_as_uint(d) << 32) | f;
}

template <typename T>
__global__ void __launch_bounds__(256)
sdf_cull_kernel(const T* __restrict__ points, int64_t P, const T* __restrict__ rec, const float4* __restrict__ bounds,
                int64_t ntiles, int64_t F, unsigned long long* __restrict__ keys) {
  __shared__ float4 sb[SDF_BCHUNK];
  const int lane = threadIdx.x & 31;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = i < P;
  const int64_t ii = valid ? i : P - 1;
  const v3<T> p = {points[ii * 3], points[ii * 3 + 1], points[ii * 3 + 2]};
  const float px = (float)p.x, py = (float)p.y, pz = (float)p.z;
  const float pabs = fabsf(px) + fabsf(py) + fabsf(pz);
  // ---- pass 1: the tightest upper bound on the distance to the mesh
  float ubmin = INFINITY;
  int64_t tstar = 0;
  for (int64_t c0 = 0; c0 < ntiles; c0 += SDF_BCHUNK) {
    const int n = (int)min((int64_t)SDF_BCHUNK, ntiles - c0);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) sb[k] = bounds[c0 + k];
    __syncthreads();
    for (int k = 0; k < n; k++) {
      const float4 b = sb[k];
      const float dx = px - b.x, dy = py - b.y, dz = pz - b.z;
      const float d = sqrtf(dx * dx + dy * dy + dz * dz);
      const float slack = 1e-6f * (pabs + fabsf(b.x) + fabsf(b.y) + fabsf(b.z) + b.w);
      const float ub = d + b.w + slack;
      if (ub < ubmin) { ubmin = ub; tstar = c0 + k; }
    }
  }
  float best = ubmin * ubmin * (1.f + 1e-5f);
  if (!(best < INFINITY)) best = INFINITY;           // NaN coordinates: no culling at all
  unsigned long long key = sdf_pack(best, 0xFFFFFFFFu);
  // the tile that gave the upper bound first (cooperatively, one point at a time): it usually holds the closest face or a
  // neighbour of it, so pass 2 starts from a near-final best and visits ~half as many tiles
  {
    unsigned m = __ballot_sync(0xffffffffu, valid);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const v3<T> q = {__shfl_sync(0xffffffffu, p.x, src), __shfl_sync(0xffffffffu, p.y, src), __shfl_sync(0xffffffffu, p.z, src)};
      const int64_t f0 = __shfl_sync(0xffffffffu, tstar, src) * SDF_FT;
      const int cnt = (int)min((int64_t)SDF_FT, F - f0);
      unsigned long long kk = ~0ull;
#pragma unroll
      for (int h = 0; h < SDF_FT / 32; h++) {
        const int f = lane + 32 * h;
        if (f < cnt) {
          v3<T> cp;
          const float dd = sdf_closest<T>(q, rec + (f0 + f) * SDF_REC, cp);
          const unsigned long long k2 = sdf_pack(dd, (unsigned)(f0 + f));
          kk = k2 < kk ? k2 : kk;
        }
      }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, kk, o);
        kk = other < kk ? other : kk;
      }
      if (lane == src) key = kk < key ? kk : key;
    }
    best = __uint_as_float((unsigned)(key >> 32));
  }
  // ---- pass 2: visit the tiles whose lower bound can still beat or tie the best
  for (int64_t c0 = 0; c0 < ntiles; c0 += SDF_BCHUNK) {
    const int n = (int)min((int64_t)SDF_BCHUNK, ntiles - c0);
    __syncthreads();
    for (int k = threadIdx.x; k < n; k += blockDim.x) sb[k] = bounds[c0 + k];
    __syncthreads();
    for (int k = 0; k < n; k++) {
      const float4 b = sb[k];
      const float dx = px - b.x, dy = py - b.y, dz = pz - b.z;
      const float d = sqrtf(dx * dx + dy * dy + dz * dz);
      const float slack = 1e-6f * (pabs + fabsf(b.x) + fabsf(b.y) + fabsf(b.z) + b.w);
      const float lb = d - b.w - slack;
      const bool skip = lb > 0.f && lb * lb * (1.f - 1e-5f) > best * (1.f + 1e-5f);     // false for NaN: visit
      const bool need = valid && !skip;
      const unsigned mask = __ballot_sync(0xffffffffu, need);
      if (mask == 0u) continue;
      const int64_t f0 = (c0 + k) * SDF_FT;
      const int cnt = (int)min((int64_t)SDF_FT, F - f0);
      if (__popc(mask) >= 16) {                       // most of the warp wants this tile: every lane scans it for its own point
        if (need) {
          for (int f = 0; f < cnt; f++) {
            v3<T> cp;
            const float dd = sdf_closest<T>(p, rec + (f0 + f) * SDF_REC, cp);
            const unsigned long long kk = sdf_pack(dd, (unsigned)(f0 + f));
            key = kk < key ? kk : key;
          }
        }
      } else {                                        // few lanes: the warp evaluates the tile for one point at a time
        unsigned m = mask;
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const v3<T> q = {__shfl_sync(0xffffffffu, p.x, src), __shfl_sync(0xffffffffu, p.y, src), __shfl_sync(0xffffffffu, p.z, src)};
          unsigned long long kk = ~0ull;
#pragma unroll
          for (int h = 0; h < SDF_FT / 32; h++) {
            const int f = lane + 32 * h;
            if (f < cnt) {
              v3<T> cp;
              const float dd = sdf_closest<T>(q, rec + (f0 + f) * SDF_REC, cp);
              const unsigned long long k2 = sdf_pack(dd, (unsigned)(f0 + f));
              kk = k2 < kk ? k2 : kk;
            }
          }
#pragma unroll
          for (int o = 16; o; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, kk, o);
            kk = other < kk ? other : kk;
          }
          if (lane == src) key = kk < key ? kk : key;
        }
      }
      best = __uint_as_float((unsigned)(key >> 32));
    }
  }
  if (valid) {
    if ((unsigned)(key & 0xFFFFFFFFull) == 0xFFFFFFFFu) {     // safety net (NaN bounds): exhaustive scan
      key = ~0ull;
      for (int64_t f = 0; f < F; f++) {
        v3<T> cp;
        const float dd = sdf_closest<T>(p, rec + f * SDF_REC, cp);
        const unsigned long long kk = sdf_pack(dd, (unsigned)f);
        if (f == 0 || kk < key) key = kk;
      }
    }
    keys[i] = key;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
sdf_final_kernel(const T* __restrict__ points, int64_t P, const T* __restrict__ rec, const unsigned long long* __restrict__ keys,
                 T* __restrict__ dist, int* __restrict__ sgn, T* __restrict__ normals, T* __restrict__ closest) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const unsigned f = (unsigned)(keys[i] & 0xFFFFFFFFull);
  const T* r = rec + (int64_t)f * SDF_REC;
  const v3<T> p = {points[i * 3], points[i * 3 + 1], points[i * 3 + 2]};
  v3<T> cp;
  const float d = sdf_closest<T>(p, r, cp);
  const v3<T> dv = vsub(p, cp);
  const T inv = sdf_rsqrt(1e-16f + vnorm2(dv));                          // grad_normal, :233
  const v3<T> n = {r[15], r[19], r[23]};
  if (dist) dist[i] = (T)d;
  if (sgn) sgn[i] = vdot(dv, n) >= (T)0 ? 1 : -1;                          // :234
  if (normals) { normals[i * 3] = dv.x * inv; normals[i * 3 + 1] = dv.y * inv; normals[i * 3 + 2] = dv.z * inv; }
  if (closest) { closest[i * 3] = cp.x; closest[i * 3 + 1] = cp.y; closest[i * 3 + 2] = cp.z; }
}

// unbatched_triangle_distance_backward_cuda_kernel (:256-270): grad_points = (p - closest) * (2 grad_dist)
template <typename T>
__global__ void __launch_bounds__(256)
sdf_backward_kernel(const T* __restrict__ gd, const T* __restrict__ points, const T* __restrict__ closest, int64_t P,
                    T* __restrict__ gp) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const T g = (T)(2. * gd[i]);
#pragma unroll
  for (int r = 0; r < 3; r++) gp[i * 3 + r] = (points[i * 3 + r] - closest[i * 3 + r]) * g;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
struct cdx_sdf_state {
  void* rec = nullptr;                  // [Fpad][SDF_REC] records of the current call
  size_t rec_bytes = 0;
  unsigned long long* keys = nullptr;   // [P]
  int64_t keys_cap = 0;
  float4* bounds = nullptr;             // [ntiles] bounding spheres of the face tiles
  int64_t bounds_cap = 0;
  int mode = 0;                         // cdx_sdf_set_mode: 0 auto (culled search for F >= 512 and P >= 128 x SMs), 1 exhaustive, 2 culled
};

void cdx_sdf_free(cdx_ctx* ctx) {
  if (!ctx || !ctx->sdf) return;
  cudaFree(ctx->sdf->rec);
  cudaFree(ctx->sdf->keys);
  cudaFree(ctx->sdf->bounds);
  delete ctx->sdf;
  ctx->sdf = nullptr;
}

template <typename T>
static int sdf_forward_t(cdx_ctx* ctx, cdx_sdf_state* s, const T* points, int64_t P, const T* fv, int64_t F, T* dist, int* sgn,
                         T* normals, T* closest, cudaStream_t st) {
  const int64_t ntiles = (F + SDF_FT - 1) / SDF_FT, Fpad = ntiles * SDF_FT;
  const size_t need = (size_t)Fpad * SDF_REC * sizeof(T);
  if (need > s->rec_bytes) {
    CDX_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->rec);
    s->rec = nullptr; s->rec_bytes = 0;
    CDX_CUDA(cudaMalloc(&s->rec, need));
    s->rec_bytes = need;
  }
  if (P > s->keys_cap) {
    CDX_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->keys);
    s->keys = nullptr; s->keys_cap = 0;
    CDX_CUDA(cudaMalloc((void**)&s->keys, sizeof(unsigned long long) * (size_t)P));
    s->keys_cap = P;
  }
  T* rec = reinterpret_cast<T*>(s->rec);
  sdf_face_kernel<T><<<(unsigned)((Fpad + 127) / 128), 128, 0, st>>>(fv, F, Fpad, rec);
  CDX_CHECK_LAUNCH(ctx);
  // the culled search is one serial walk per point: it pays once there are enough points to fill the chip (measured: 4,096
  // points x 20,480 faces 1.18 ms culled against 0.29 ms exhaustive; 1 M points 14.4 ms against 52.4 ms)
  const bool cull = s->mode == 2 || (s->mode == 0 && F >= 512 && P >= (int64_t)ctx->sms * 128);
  if (cull) {
    if (ntiles > s->bounds_cap) {
      CDX_CUDA(cudaStreamSynchronize(st));
      cudaFree(s->bounds);
      s->bounds = nullptr; s->bounds_cap = 0;
      CDX_CUDA(cudaMalloc((void**)&s->bounds, sizeof(float4) * (size_t)ntiles));
      s->bounds_cap = ntiles;
    }
    sdf_tile_kernel<T><<<(unsigned)ntiles, 32, 0, st>>>(rec, F, s->bounds);
    CDX_CHECK_LAUNCH(ctx);
    // small batches: 64-thread CTAs so that the points spread over the SMs
    const int bt = P >= (int64_t)ctx->sms * 512 ? 256 : 64;
    sdf_cull_kernel<T><<<(unsigned)((P + bt - 1) / bt), bt, 0, st>>>(points, P, rec, s->bounds, ntiles, F, s->keys);
    CDX_CHECK_LAUNCH(ctx);
  } else {
    CDX_CUDA(cudaMemsetAsync(s->keys, 0xFF, sizeof(unsigned long long) * (size_t)P, st));
    // two points per thread (each record load serves two evaluations) whenever point tiles x face tiles can still give ~4 CTAs
    // per SM; the faces are split over blockIdx.y until they do
    const bool two = ((P + SDF_THREADS * 2 - 1) / (SDF_THREADS * 2)) * ntiles >= (int64_t)ctx->sms * 4;
    const int ppt = two ? 2 : 1;
    const int64_t nx = (P + SDF_THREADS * ppt - 1) / (SDF_THREADS * ppt);
    int64_t ny = (ctx->sms * 4 + nx - 1) / nx;
    if (ny > ntiles) ny = ntiles;
    if (ny > 65535) ny = 65535;
    if (ny < 1) ny = 1;
    const int tpc = (int)((ntiles + ny - 1) / ny);
    ny = (ntiles + tpc - 1) / tpc;
    dim3 grid((unsigned)nx, (unsigned)ny);
    if (two) sdf_pair_kernel<T, 2><<<grid, SDF_THREADS, 0, st>>>(points, P, rec, F, tpc, s->keys);
    else sdf_pair_kernel<T, 1><<<grid, SDF_THREADS, 0, st>>>(points, P, rec, F, tpc, s->keys);
    CDX_CHECK_LAUNCH(ctx);
  }
  sdf_final_kernel<T><<<(unsigned)((P + 255) / 256), 256, 0, st>>>(points, P, rec, s->keys, dist, sgn, normals, closest);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}

extern "C" int cdx_sdf_forward(cdx_ctx* ctx, const void* points, int64_t P, const void* face_vertices, int64_t F, int dtype,
                               void* dist, int32_t* sign, void* normals, void* closest, void* stream) {
  CDX_NVTX("cdx_sdf_forward");
  if (!ctx || P < 0 || F < 0 || (dtype != 0 && dtype != 1) || (P > 0 && !points) || (F > 0 && !face_vertices)) {
    cdx_set_error("cdx_sdf_forward: bad argument (dtype 0 = float32, 1 = float64)");
    return CDX_ERR_ARG;
  }
  if (F > 0x7FFFFFFF) { cdx_set_error("cdx_sdf_forward: more than 2^31 - 1 faces"); return CDX_ERR_ARG; }
  if (P == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t es = dtype ? 8 : 4;
  if (F == 0) {   // the reference's kernel never writes: outputs keep the zeros of torch.zeros (sdf.py:42-49)
    if (dist) CDX_CUDA(cudaMemsetAsync(dist, 0, es * P, st));
    if (sign) CDX_CUDA(cudaMemsetAsync(sign, 0, 4 * (size_t)P, st));
    if (normals) CDX_CUDA(cudaMemsetAsync(normals, 0, es * P * 3, st));
    if (closest) CDX_CUDA(cudaMemsetAsync(closest, 0, es * P * 3, st));
    return CDX_OK;
  }
  if (!ctx->sdf) ctx->sdf = new cdx_sdf_state();
  if (dtype) return sdf_forward_t<double>(ctx, ctx->sdf, (const double*)points, P, (const double*)face_vertices, F, (double*)dist,
                                          sign, (double*)normals, (double*)closest, st);
  return sdf_forward_t<float>(ctx, ctx->sdf, (const float*)points, P, (const float*)face_vertices, F, (float*)dist, sign,
                              (float*)normals, (float*)closest, st);
}

extern "C" int cdx_sdf_set_mode(cdx_ctx* ctx, int mode) {
  if (!ctx || mode < 0 || mode > 2) { cdx_set_error("cdx_sdf_set_mode: mode must be 0 (auto), 1 (exhaustive) or 2 (culled)"); return CDX_ERR_ARG; }
  if (!ctx->sdf) ctx->sdf = new cdx_sdf_state();
  ctx->sdf->mode = mode;
  return CDX_OK;
}

extern "C" int cdx_sdf_backward(cdx_ctx* ctx, const void* grad_dist, const void* points, const void* closest, int64_t P,
                                int dtype, void* grad_points, void* stream) {
  CDX_NVTX("cdx_sdf_backward");
  if (!ctx || P < 0 || (dtype != 0 && dtype != 1) || (P > 0 && (!grad_dist || !points || !closest || !grad_points))) {
    cdx_set_error("cdx_sdf_backward: bad argument");
    return CDX_ERR_ARG;
  }
  if (P == 0) return CDX_OK;
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned nb = (unsigned)((P + 255) / 256);
  if (dtype) sdf_backward_kernel<double><<<nb, 256, 0, st>>>((const double*)grad_dist, (const double*)points, (const double*)closest, P, (double*)grad_points);
  else sdf_backward_kernel<float><<<nb, 256, 0, st>>>((const float*)grad_dist, (const float*)points, (const float*)closest, P, (float*)grad_points);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}
