// Surface extraction from a GPIS mean grid (GPIS.topcd, gpis.py:127-150): a voxel belongs to the surface when its mean is
// negative and at least one of its 6 face neighbours is not; only interior voxels (1..steps-2 per axis) are tested.  The
// reference walks the steps^3 grid in a triple Python loop; here one stencil pass counts per 256-voxel block, a
// single-CTA scan turns the counts into offsets, and a second pass writes the LINEAR indices in ascending (C) order -
// the order numpy's boolean-mask indexing returns.
#include "cdx_internal.cuh"

#include <mutex>

// Scratch of cdx_grid_surface_static (below): the per-block surface counts / offsets of a grid of at most 256^3 voxels
// (65,536 blocks of 256 voxels, 256 KB).  Lives in every GPU that loads the library until the process exits.  In host code
// this NAME is only a stand-in for the symbol - its device address comes from cudaGetSymbolAddress (grid_static_scratch).
#define CDX_GRID_STATIC_MAX_STEPS 256
#define CDX_GRID_STATIC_MAX_BLOCKS (CDX_GRID_STATIC_MAX_STEPS * CDX_GRID_STATIC_MAX_STEPS * CDX_GRID_STATIC_MAX_STEPS / 256)
static __device__ __align__(256) unsigned cdx_grid_static_counts[CDX_GRID_STATIC_MAX_BLOCKS];

namespace {

__device__ __forceinline__ bool grid_is_surface(const float* __restrict__ m, int steps, int64_t lin) {
  const int64_t s2 = (int64_t)steps * steps;
  const int i = (int)(lin / s2), j = (int)((lin / steps) % steps), k = (int)(lin % steps);
  if (i < 1 || j < 1 || k < 1 || i > steps - 2 || j > steps - 2 || k > steps - 2) return false;
  if (!(m[lin] < 0.0f)) return false;
  return !(m[lin - s2] < 0.0f) || !(m[lin + s2] < 0.0f) || !(m[lin - steps] < 0.0f) || !(m[lin + steps] < 0.0f) ||
         !(m[lin - 1] < 0.0f) || !(m[lin + 1] < 0.0f);
}

__global__ void __launch_bounds__(256) grid_count_kernel(const float* __restrict__ m, int steps, int64_t total,
                                                         unsigned* __restrict__ counts) {
  const int64_t lin = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const int c = __syncthreads_count(lin < total && grid_is_surface(m, steps, lin));
  if (threadIdx.x == 0) counts[blockIdx.x] = (unsigned)c;
}

// exclusive scan of the block counts (one CTA; nb <= 2^23 for steps <= 1290)
__global__ void __launch_bounds__(1024) grid_scan_kernel(unsigned* __restrict__ counts, int nb, long long* __restrict__ total_out) {
  __shared__ unsigned part[1024];
  const int per = (nb + 1023) / 1024, b0 = threadIdx.x * per, b1 = min(nb, b0 + per);
  unsigned s = 0;
  for (int b = b0; b < b1; b++) s += counts[b];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    unsigned v = (int)threadIdx.x >= o ? part[threadIdx.x - o] : 0u;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  unsigned run = part[threadIdx.x] - s;   // exclusive prefix of this thread's segment
  for (int b = b0; b < b1; b++) { const unsigned c = counts[b]; counts[b] = run; run += c; }
  if (threadIdx.x == 1023) *total_out = (long long)part[1023];
}

__global__ void __launch_bounds__(256) grid_scatter_kernel(const float* __restrict__ m, int steps, int64_t total,
                                                           const unsigned* __restrict__ offsets, long long* __restrict__ idx) {
  __shared__ unsigned wsum[8];
  const int64_t lin = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const bool on = lin < total && grid_is_surface(m, steps, lin);
  const unsigned bal = __ballot_sync(0xffffffffu, on);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) wsum[w] = __popc(bal);
  __syncthreads();
  unsigned base = offsets[blockIdx.x];
  for (int i = 0; i < w; i++) base += wsum[i];
  if (on) idx[base + __popc(bal & ((1u << lane) - 1u))] = (long long)lin;
}

// device address of cdx_grid_static_counts on `device` (which must be current): looked up once per device
int grid_static_scratch(int device, unsigned** out) {
  static std::mutex mu;
  static unsigned* addr[64] = {};
  if (device < 0 || device >= 64) { cdx_set_error("cdx_grid_surface_static: device ordinal out of range"); return CDX_ERR_ARG; }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    unsigned* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, cdx_grid_static_counts);
    if (e != cudaSuccess || !p) {
      cdx_set_error("cdx_grid_surface_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return CDX_ERR_CUDA;
    }
    addr[device] = p;
  }
  *out = addr[device];
  return CDX_OK;
}

}  // namespace

extern "C" int cdx_grid_surface(cdx_ctx* ctx, const float* mean, int steps, long long* idx, long long* count, void* stream) {
  CDX_NVTX("cdx_grid_surface");
  if (!ctx || !mean || !idx || !count || steps < 1 || steps > 1290) {
    cdx_set_error("cdx_grid_surface: bad argument (1 <= steps <= 1290)");
    return CDX_ERR_ARG;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = (int64_t)steps * steps * steps;
  const int nb = (int)((total + 255) / 256);
  unsigned* counts = nullptr;
  CDX_CUDA(cudaMallocAsync((void**)&counts, sizeof(unsigned) * nb, st));
  grid_count_kernel<<<nb, 256, 0, st>>>(mean, steps, total, counts);
  CDX_CHECK_LAUNCH(ctx);
  grid_scan_kernel<<<1, 1024, 0, st>>>(counts, nb, count);
  CDX_CHECK_LAUNCH(ctx);
  grid_scatter_kernel<<<nb, 256, 0, st>>>(mean, steps, total, counts, idx);
  CDX_CHECK_LAUNCH(ctx);
  CDX_CUDA(cudaFreeAsync(counts, st));
  return CDX_OK;
}

// cdx_grid_surface for callers that must not allocate (CUDA graph capture, serving loops): same three kernels, the block
// counts in the file-scope array above instead of a stream-ordered allocation.  No cudaMalloc / cudaFree of any kind.
// Every call on a device shares that array: calls that overlap on the GPU (different streams) race.  Nothing in it has to
// start at zero - grid_count_kernel overwrites entries [0, nb) before the scan reads them.
extern "C" int cdx_grid_surface_static(cdx_ctx* ctx, const float* mean, int steps, long long* idx, long long* count, void* stream) {
  if (!ctx || !mean || !idx || !count || steps < 1) {
    cdx_set_error("cdx_grid_surface_static: bad argument");
    return CDX_ERR_ARG;
  }
  if (steps > CDX_GRID_STATIC_MAX_STEPS) {
    cdx_set_error("cdx_grid_surface_static: steps = %d exceeds the static scratch (steps <= %d); use cdx_grid_surface", steps,
                  CDX_GRID_STATIC_MAX_STEPS);
    return CDX_ERR_ARG;
  }
  CDX_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  unsigned* counts = nullptr;
  if (int rc = grid_static_scratch(ctx->device, &counts)) return rc;
  const int64_t total = (int64_t)steps * steps * steps;
  const int nb = (int)((total + 255) / 256);
  grid_count_kernel<<<nb, 256, 0, st>>>(mean, steps, total, counts);
  CDX_CHECK_LAUNCH(ctx);
  grid_scan_kernel<<<1, 1024, 0, st>>>(counts, nb, count);
  CDX_CHECK_LAUNCH(ctx);
  grid_scatter_kernel<<<nb, 256, 0, st>>>(mean, steps, total, counts, idx);
  CDX_CHECK_LAUNCH(ctx);
  return CDX_OK;
}
