// Pipe-rate microbenchmarks for the FP64 units this path runs on (MEASURED_PEAKS.json has only HBM and bf16).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/cdx_pipebench cdx_pipebench.cu
// Prints one JSON object: DFMA TF/s, DMMA(m8n8k4) TF/s, both interleaved, FP32 FFMA TF/s, MUFU.RSQ rate.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int MODE>  // 0 = DFMA, 1 = DMMA, 2 = both, 3 = FFMA, 4 = rsqrtf + 2 DFMA
__global__ void __launch_bounds__(256) pipe_kernel(double* out, int iters) {
  double a = threadIdx.x * 1e-3 + 1.0, b = 1.0 - threadIdx.x * 1e-6;
  double acc[16];
  double d[16];
  float f[16];
#pragma unroll
  for (int i = 0; i < 16; i++) { acc[i] = i; d[i] = 0.0; f[i] = i; }
  for (int it = 0; it < iters; it++) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; i++) acc[i] = fma(acc[i], b, a);
    }
    if (MODE == 1 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; i++) dmma884(d[2 * i], d[2 * i + 1], a, b);
    }
    if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 16; i++) f[i] = fmaf(f[i], (float)b, (float)a);
    }
    if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 16; i++) {
        float y = rsqrtf((float)acc[i]);
        acc[i] = fma(acc[i], (double)y, a);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i] + d[i] + f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static double run(int iters, int blocks_per_sm) {
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int grid = sms * blocks_per_sm;
  double* out;
  cudaMalloc(&out, sizeof(double) * grid * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  pipe_kernel<MODE><<<grid, 256>>>(out, iters / 10);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0);
    pipe_kernel<MODE><<<grid, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaFree(out);
  return (double)best * 1e-3 / ((double)grid * 256.0 * iters);  // seconds per thread-iteration
}

int main() {
  const int iters = 20000;
  double t_dfma = run<0>(iters, 4), t_dmma = run<1>(iters, 4), t_both = run<2>(iters, 4), t_ffma = run<3>(iters, 4),
         t_rsq = run<4>(iters, 4);
  // per thread-iteration: DFMA 16 fma = 32 flop; DMMA 8 x (512/32) = 128 flop; FFMA 32 flop
  printf("{\"dfma_tflops\": %.2f, \"dmma_m8n8k4_tflops\": %.2f, \"dfma_plus_dmma_tflops\": %.2f, "
         "\"dfma_plus_dmma_time_vs_sum\": %.3f, \"ffma_tflops\": %.2f, \"rsqrtf_cvt_dfma_gops\": %.1f}\n",
         32.0 / t_dfma * 1e-12, 128.0 / t_dmma * 1e-12, 160.0 / t_both * 1e-12, t_both / (t_dfma + t_dmma),
         32.0 / t_ffma * 1e-12, 16.0 / t_rsq * 1e-9);
  return 0;
}
