// Pipe-rate microbenchmarks for the FP64 units this path runs on (MEASURED_PEAKS.json has only HBM and bf16).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/cdx_pipebench cdx_pipebench.cu
// Prints one JSON object: DFMA TF/s, DMMA(m8n8k4) TF/s, both interleaved, FP32 FFMA TF/s, MUFU.RSQ rate, and the
// tcgen05.mma kind::i8 issue rate (TOP/s) for the operand shapes of gpis_std_i8x_kernel.
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

// The k* generator of gpis_std_i8x_kernel in isolation (cdx_query_i8x.cu, generator warps): per (query, training point) pair
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

// ---- tcgen05 kind::i8 issue rate (the roof of gpis_std_i8x_kernel): one CTA per SM, one thread issues `count` MMAs of
// M=128 x N x K=32 onto two alternating TMEM accumulators, operands constant in shared memory (SS) or A in tensor memory (TS)
__device__ __forceinline__ unsigned pb_s32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long pb_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
  return (unsigned long long)((saddr >> 4) & 0x3FFFu) | ((unsigned long long)((lbo >> 4) & 0x3FFFu) << 16) |
         ((unsigned long long)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) i8_peak_kernel(int count, unsigned* out) {
  extern __shared__ __align__(128) unsigned char pb_smem[];
  __shared__ unsigned long long bar;
  __shared__ unsigned tmem_slot;
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < (128 * 32 + N * 32) / 4; i += 128)      // pseudo-random digits: the data the MMAs toggle on is part of their power draw
    reinterpret_cast<unsigned*>(pb_smem)[i] = (unsigned)(i + 1) * 2654435761u ^ ((unsigned)(i + 7) * 40503u << 11);
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(pb_s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(pb_s32(&tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const unsigned tmem = tmem_slot;
  if (t == 0) {
    const unsigned idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((unsigned)(N >> 3) << 17) | (8u << 24);
    const unsigned long long da = pb_desc(pb_s32(pb_smem), 128 * 16, 128);
    const unsigned long long db = pb_desc(pb_s32(pb_smem) + 128 * 32, N * 16, 128);
    const int nacc = 512 / N >= 2 ? 2 : 1;
    for (int i = 0; i < count; i++) {
      const unsigned d = tmem + (unsigned)((i % nacc) * N);
      if (TS)
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %3, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %4, {%5,%5,%5,%5}, p;\n}\n" ::"r"(d),
                     "r"(tmem + 2u * N), "l"(db), "r"((unsigned)(i >= nacc)), "r"(idesc), "r"(0u) : "memory");
      else
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %3, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %4, {%5,%5,%5,%5}, p;\n}\n" ::"r"(d),
                     "l"(da), "l"(db), "r"((unsigned)(i >= nacc)), "r"(idesc), "r"(0u) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(pb_s32(&bar)) : "memory");
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(pb_s32(&bar)) : "memory");
    out[blockIdx.x] = count;
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem));
}

template <int N, bool TS>
static double run_i8(int count) {   // -> TOP/s (2 x multiply-adds) of the whole chip
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  unsigned* out;
  cudaMalloc(&out, sizeof(unsigned) * sms);
  const size_t smem = 128 * 32 + N * 32;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  i8_peak_kernel<N, TS><<<sms, 128, smem>>>(count / 10, out);
  if (cudaDeviceSynchronize() != cudaSuccess) { cudaGetLastError(); cudaFree(out); return 0.0; }
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0);
    i8_peak_kernel<N, TS><<<sms, 128, smem>>>(count, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaFree(out);
  return 2.0 * 128.0 * N * 32.0 * (double)count * sms / ((double)best * 1e-3) * 1e-12;
}

// the same loop back to back for ~3 s; the rate of the second half: what the chip sustains under its power cap (the std
// kernel of the 10 M-query step is timed inside ~1 s of uninterrupted tensor work, so this - not the burst figure - is its roof)
template <int N>
static double run_i8_sustained(int count) {
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  unsigned* out;
  cudaMalloc(&out, sizeof(unsigned) * sms);
  const size_t smem = 128 * 32 + N * 32;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int launches = 500;           // ~2.8 ms each at the burst rate
  for (int r = 0; r < launches; r++) i8_peak_kernel<N, false><<<sms, 128, smem>>>(count, out);
  cudaEventRecord(e0);
  for (int r = 0; r < launches; r++) i8_peak_kernel<N, false><<<sms, 128, smem>>>(count, out);
  cudaEventRecord(e1);
  if (cudaEventSynchronize(e1) != cudaSuccess) { cudaGetLastError(); cudaFree(out); return 0.0; }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaFree(out);
  return 2.0 * 128.0 * N * 32.0 * (double)count * sms * launches / ((double)ms * 1e-3) * 1e-12;
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
  const double i8_ss256 = run_i8<256, false>(40000), i8_ss128 = run_i8<128, false>(40000), i8_ss96 = run_i8<96, false>(40000),
               i8_ss64 = run_i8<64, false>(40000), i8_ts64 = run_i8<64, true>(40000);
  const double i8_sus256 = run_i8_sustained<256>(40000), i8_sus96 = run_i8_sustained<96>(40000);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  // per thread-iteration: DFMA 16 fma = 32 flop; DMMA 8 x (512/32) = 128 flop; FFMA 32 flop
  printf("{\"dfma_tflops\": %.2f, \"dmma_m8n8k4_tflops\": %.2f, \"dfma_plus_dmma_tflops\": %.2f, "
         "\"dfma_plus_dmma_time_vs_sum\": %.3f, \"ffma_tflops\": %.2f, \"rsqrtf_cvt_dfma_gops\": %.1f, "
         "\"kstar_gen_us_per_128x64_stage_1cta\": %.4f, \"kstar_gen_us_per_128x64_stage_2cta\": %.4f, "
         "\"kstar_gen_us_per_128x64_stage_1cta_88regs\": %.4f, \"i8_mma_ss_m128n256k32_tops\": %.1f, "
         "\"i8_mma_ss_m128n128k32_tops\": %.1f, \"i8_mma_ss_m128n96k32_tops\": %.1f, \"i8_mma_ss_m128n64k32_tops\": %.1f, "
         "\"i8_mma_ts_m128n64k32_tops\": %.1f, \"i8_mma_ss_m128n256k32_tops_sustained\": %.1f, "
         "\"i8_mma_ss_m128n96k32_tops_sustained\": %.1f, \"max_sm_khz\": %d}\n",
         32.0 / t_dfma * 1e-12, 128.0 / t_dmma * 1e-12, 160.0 / t_both * 1e-12, t_both / (t_dfma + t_dmma),
         32.0 / t_ffma * 1e-12, 16.0 / t_rsq * 1e-9, g1 * 1e6, g2 * 1e6, g3 * 1e6, i8_ss256, i8_ss128, i8_ss96, i8_ss64, i8_ts64, i8_sus256, i8_sus96, khz);
  return 0;
}
