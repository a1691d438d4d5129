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

// The k* generator of gpis_std_i8_kernel in isolation (cdx_query_i8.cu, generator warps): per (query, training point) pair
// 4 FP64 for the expanded squared distance, float rsqrt seed by bit moves, 1 Newton step (3 FP64), cubic + fixed-point magic
// (2 FP64), then the 4x4 byte transposes.  No tensor memory, no barriers: measures what the FP64 / ALU / XU pipes sustain for
// this instruction mix at the kernel's occupancy (512 generator threads per SM).
__device__ __forceinline__ float pb_rsqrt_ftz(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x)); return y; }
template <int CAP>   // CAP = launch bound in threads: 704 reproduces the register cap (88) of the 22-warp query kernel
__global__ void __launch_bounds__(CAP) gen_kernel(const double* __restrict__ pts, unsigned* __restrict__ out, int stages) {
  __shared__ double2 P2[2 * 64];
  for (int i = threadIdx.x; i < 128; i += blockDim.x) P2[i] = reinterpret_cast<const double2*>(pts)[i];
  __syncthreads();
  const double xq = 0.01 * (threadIdx.x & 127), yq = 0.02, zq = -0.03 + 1e-4 * blockIdx.x;
  const double qq = fma(xq, xq, fma(yq, yq, zq * zq)) + 1e-16;
  const double c2 = 2.0 * 1048576.0, c3 = -3.0 * 0.35 * 1048576.0, c0m = 0.04 * 1048576.0 + 6755399441055744.0;
  const int kg = (threadIdx.x >> 7) & 3;
  unsigned acc = 0;
  for (int st = 0; st < stages; st++) {
    const double2* Pk = P2 + kg * 32;
#pragma unroll
    for (int gq = 0; gq < 4; gq++) {
      unsigned L[4], H[4];
#pragma unroll
      for (int p = 0; p < 4; p++) {
        const int j = gq * 4 + p;
        const double2 a = Pk[2 * j], b = Pk[2 * j + 1];
        const double s = fma(xq, a.x, fma(yq, a.y, fma(zq, b.x, b.y + qq))) + 1e-9 * st;
        const unsigned shi = (unsigned)__double2hiint(s), slo = (unsigned)__double2loint(s);
        const unsigned fb = __funnelshift_l(slo, shi - 0x38000000u, 3);
        const unsigned yb = __float_as_uint(pb_rsqrt_ftz(__uint_as_float(fb)));
        const double yd = __hiloint2double((int)((yb >> 3) + 0x38000000u), (int)(yb << 29));
        const double hy = __hiloint2double((int)((yb >> 3) + 0x37F00000u), (int)(yb << 29));
        double d = s * yd;
        d = fma(fma(-d, d, s), hy, d);
        const double res = fma(s, fma(c2, d, c3), c0m);
        L[p] = (unsigned)__double2loint(res);
        H[p] = (unsigned)__double2hiint(res);
      }
      const unsigned t0 = __byte_perm(L[0], L[1], 0x5140), t1 = __byte_perm(L[2], L[3], 0x5140);
      const unsigned t2 = __byte_perm(L[0], L[1], 0x7362), t3 = __byte_perm(L[2], L[3], 0x7362);
      const unsigned t4 = __byte_perm(H[0], H[1], 0x5140), t5 = __byte_perm(H[2], H[3], 0x5140);
      acc ^= __byte_perm(t0, t1, 0x5410) ^ __byte_perm(t0, t1, 0x7632) ^ __byte_perm(t2, t3, 0x5410) ^
             __byte_perm(t2, t3, 0x7632) ^ __byte_perm(t4, t5, 0x5410);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int CAP>
static double run_gen(int stages, int blocks_per_sm) {   // -> cycles-equivalent: seconds per (SM, stage of 128x64 pairs)
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = sms * blocks_per_sm;
  double* pts;
  unsigned* out;
  cudaMalloc(&pts, sizeof(double) * 256);
  cudaMalloc(&out, sizeof(unsigned) * grid * 512);
  double h[256];
  for (int i = 0; i < 256; i++) h[i] = (i % 4 == 3) ? 0.01 + 1e-4 * i : 0.02 * ((i * 7) % 11 - 5);
  cudaMemcpy(pts, h, sizeof(h), cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  gen_kernel<CAP><<<grid, 512>>>(pts, out, stages / 10);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0);
    gen_kernel<CAP><<<grid, 512>>>(pts, out, stages);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaFree(pts);
  cudaFree(out);
  return (double)best * 1e-3 / ((double)stages * blocks_per_sm);   // each block-stage = 512 threads x 16 pairs = 128 x 64 pairs
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
  const double g1 = run_gen<512>(20000, 1), g2 = run_gen<512>(20000, 2), g3 = run_gen<704>(20000, 1);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  // per thread-iteration: DFMA 16 fma = 32 flop; DMMA 8 x (512/32) = 128 flop; FFMA 32 flop
  printf("{\"dfma_tflops\": %.2f, \"dmma_m8n8k4_tflops\": %.2f, \"dfma_plus_dmma_tflops\": %.2f, "
         "\"dfma_plus_dmma_time_vs_sum\": %.3f, \"ffma_tflops\": %.2f, \"rsqrtf_cvt_dfma_gops\": %.1f, "
         "\"kstar_gen_us_per_128x64_stage_1cta\": %.4f, \"kstar_gen_us_per_128x64_stage_2cta\": %.4f, "
         "\"kstar_gen_us_per_128x64_stage_1cta_88regs\": %.4f, \"max_sm_khz\": %d}\n",
         32.0 / t_dfma * 1e-12, 128.0 / t_dmma * 1e-12, 160.0 / t_both * 1e-12, t_both / (t_dfma + t_dmma),
         32.0 / t_ffma * 1e-12, 16.0 / t_rsq * 1e-9, g1 * 1e6, g2 * 1e6, g3 * 1e6, khz);
  return 0;
}
