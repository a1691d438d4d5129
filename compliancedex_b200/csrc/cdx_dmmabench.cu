// How the FP64 DMMA (mma.sync m8n8k4) rate depends on the launch shape: warps per SM sub-partition, independent accumulators
// per warp, and distinct A/B operand registers (the shape of a register-tiled GEMM warp: MT x NT accumulator blocks fed by
// MT + NT fragment registers).  Prints one JSON object; the evidence for the warp-tile choices in cdx_query.cu / cdx_gemm.cu.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/cdx_dmmabench cdx_dmmabench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// MT x NT accumulator blocks, MT + NT operand registers refreshed from shared memory every step when LDS is set
template <int MT, int NT, bool LDS, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) tile_kernel(double* out, int iters) {
  __shared__ double sh[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sh[i] = 1.0 + 1e-9 * i;
  __syncthreads();
  double acc[MT][NT][2];
#pragma unroll
  for (int i = 0; i < MT; i++)
#pragma unroll
    for (int j = 0; j < NT; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
  double a[MT], b[NT];
#pragma unroll
  for (int i = 0; i < MT; i++) a[i] = 1.0 + threadIdx.x * 1e-6 * (i + 1);
#pragma unroll
  for (int j = 0; j < NT; j++) b[j] = 1.0 - threadIdx.x * 1e-7 * (j + 1);
  const int lane = threadIdx.x & 31;
  for (int it = 0; it < iters; it++) {
    if (LDS) {
      const int o = ((it & 7) * 64) + (lane >> 2) * 20 + (lane & 3);      // < 512 + 143; every index below stays < 2048
#pragma unroll
      for (int i = 0; i < MT; i++) a[i] = sh[(o + i * 160) & 2047];
#pragma unroll
      for (int j = 0; j < NT; j++) b[j] = sh[(o + 700 + j * 160) & 2047];
    }
#pragma unroll
    for (int i = 0; i < MT; i++)
#pragma unroll
      for (int j = 0; j < NT; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < MT; i++)
#pragma unroll
    for (int j = 0; j < NT; j++) s += acc[i][j][0] + acc[i][j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MT, int NT, bool LDS, int WARPS>
static double run(int iters) {   // -> TF/s of the chip
  const int warps_per_sm = WARPS;
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int threads = warps_per_sm * 32;
  double* out = nullptr;
  if (cudaMalloc(&out, sizeof(double) * sms * threads) != cudaSuccess) { fprintf(stderr, "cudaMalloc failed\n"); exit(1); }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  tile_kernel<MT, NT, LDS, WARPS><<<sms, threads>>>(out, iters / 10);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 3; r++) {
    cudaEventRecord(e0);
    tile_kernel<MT, NT, LDS, WARPS><<<sms, threads>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess) { fprintf(stderr, "kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); exit(1); }
    if (ms < best) best = ms;
  }
  cudaFree(out);
  const double flop = (double)sms * warps_per_sm * (double)iters * MT * NT * 512.0;
  return flop / (best * 1e-3) * 1e-12;
}

int main() {
  const int it = 4000;
  printf("{");
#define ROW(MT, NT, LDS, TAG)                                                                                   \
  printf("\"" TAG "\": {\"w4\": %.2f, \"w8\": %.2f, \"w16\": %.2f, \"w32\": %.2f}, ", run<MT, NT, LDS, 4>(it), \
         run<MT, NT, LDS, 8>(it), run<MT, NT, LDS, 16>(it), run<MT, NT, LDS, 32>(it))
  ROW(1, 8, false, "tile_1x8_regs");
  ROW(2, 5, false, "tile_2x5_regs");
  ROW(2, 5, true, "tile_2x5_lds");
  ROW(4, 4, false, "tile_4x4_regs");
  ROW(4, 4, true, "tile_4x4_lds");
  ROW(2, 8, true, "tile_2x8_lds");
  ROW(4, 8, true, "tile_4x8_lds");
  printf("\"unit\": \"TFLOP/s, whole chip; wN = N warps per SM (N/4 per sub-partition)\"}\n");
  return 0;
}
