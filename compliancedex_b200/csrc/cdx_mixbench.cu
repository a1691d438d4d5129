// Does FP64 arithmetic slow the tcgen05 INT8 MMAs down?  (gpis_std_i8x_kernel generates its k* operand in FP64 while the MMAs
// run: measured there, the two do not overlap.)  One CTA per SM: warp 0 issues `count` tcgen05.mma kind::i8 (M=128, N=96, K=32,
// operands constant in shared memory) and times them with clock64; NW other warps execute a chosen instruction mix until the
// MMAs are done and count what they executed.  Prints one JSON object: per mix the MMA cycles per instruction and the
// extra MMA cycles per FP64 warp-instruction executed on the SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/cdx_mixbench cdx_mixbench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned mb_s32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long mb_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
  return (unsigned long long)((saddr >> 4) & 0x3FFFu) | ((unsigned long long)((lbo >> 4) & 0x3FFFu) << 16) |
         ((unsigned long long)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}

constexpr int N = 96, NW = 12;
// MIX: 0 idle, 1 DFMA back to back (8 chains), 2 one DFMA then 4 FFMA, 3 one DFMA then 16 FFMA, 4 FFMA only, 5 DFMA on ONE
// dependent chain (latency bound: sparse issue), 6 DMUL/DADD mix like the k* generator, 7 IMAD only
template <int MIX>
__global__ void __launch_bounds__(32 * (1 + NW), 1) mix_kernel(int count, long long* out) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ unsigned long long bar;
  __shared__ unsigned tmem_slot;
  __shared__ volatile int done;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  for (int i = t; i < (128 * 32 + N * 32) / 4; i += blockDim.x) reinterpret_cast<unsigned*>(smem)[i] = 0x01010101u * (i & 3);
  if (t == 0) {
    done = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mb_s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(mb_s32(&tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const unsigned tmem = tmem_slot;
  if (warp == 0) {
    if (lane == 0) {
      const unsigned idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((unsigned)(N >> 3) << 17) | (8u << 24);
      const unsigned long long da = mb_desc(mb_s32(smem), 128 * 16, 128), db = mb_desc(mb_s32(smem) + 128 * 32, N * 16, 128);
      const long long t0 = clock64();
      for (int i = 0; i < count; i++) {
        const unsigned d = tmem + (unsigned)((i % 5) * N);
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %3, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %4, {%5,%5,%5,%5}, p;\n}\n" ::"r"(d),
                     "l"(da), "l"(db), "r"((unsigned)(i >= 5)), "r"(idesc), "r"(0u) : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mb_s32(&bar)) : "memory");
      asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(mb_s32(&bar)) : "memory");
      const long long t1 = clock64();
      done = 1;
      if (blockIdx.x == 0) out[0] = t1 - t0;
    }
  } else if (MIX != 0) {
    double a[8];
    float f[16];
    unsigned u[8];
    const double b = 1.0 - lane * 1e-9, c = 1e-12;
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = 1.0 + i + lane; u[i] = i + lane; }
#pragma unroll
    for (int i = 0; i < 16; i++) f[i] = i + lane;
    long long n = 0;
    while (!done) {
#pragma unroll 1
      for (int rep = 0; rep < 8; rep++) {
        if (MIX == 1) {
#pragma unroll
          for (int i = 0; i < 8; i++) a[i] = fma(a[i], b, c);
        } else if (MIX == 2) {
#pragma unroll
          for (int i = 0; i < 8; i++) { a[i] = fma(a[i], b, c); f[0] = fmaf(f[0], 0.999f, 1.f); f[1] = fmaf(f[1], 0.999f, 1.f); f[2] = fmaf(f[2], 0.999f, 1.f); f[3] = fmaf(f[3], 0.999f, 1.f); }
        } else if (MIX == 3) {
#pragma unroll
          for (int i = 0; i < 8; i++) {
            a[i] = fma(a[i], b, c);
#pragma unroll
            for (int j = 0; j < 16; j++) f[j] = fmaf(f[j], 0.999f, 1.f);
          }
        } else if (MIX == 4) {
#pragma unroll
          for (int i = 0; i < 8; i++)
#pragma unroll
            for (int j = 0; j < 16; j++) f[j] = fmaf(f[j], 0.999f, 1.f);
        } else if (MIX == 5) {
#pragma unroll
          for (int i = 0; i < 8; i++) a[0] = fma(a[0], b, c);
        } else if (MIX == 6) {
#pragma unroll
          for (int i = 0; i < 8; i += 2) { a[i] = a[i] * b; a[i + 1] = a[i + 1] + c; }
        } else {
#pragma unroll
          for (int i = 0; i < 8; i++) u[i] = u[i] * 1664525u + 1013904223u;
        }
      }
      n += 64;
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += a[i] + u[i];
    for (int i = 0; i < 16; i++) s += f[i];
    if (s == 1.2345) out[3] = 1;
    if (blockIdx.x == 0 && lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(out + 1), (unsigned long long)n);
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem));
}

template <int MIX>
static void run(int count, const char* name, double base_cycles, double* cycles_out) {
  int dev, sms;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long* out;
  cudaMalloc(&out, 4 * sizeof(long long));
  const size_t smem = 128 * 32 + N * 32;
  mix_kernel<MIX><<<sms, 32 * (1 + NW), smem>>>(count / 10, out);
  cudaDeviceSynchronize();
  cudaMemset(out, 0, 4 * sizeof(long long));
  mix_kernel<MIX><<<sms, 32 * (1 + NW), smem>>>(count, out);
  cudaDeviceSynchronize();
  long long h[4];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(out);
  const double cyc = (double)h[0] / count;
  *cycles_out = cyc;
  // h[1]: warp-instructions of the mix's main operation executed by the other warps of SM 0 while the MMAs ran
  printf("\"%s\": {\"mma_cycles_per_instr\": %.2f, \"other_warp_instr_per_mma\": %.2f, \"extra_mma_cycles_per_other_warp_instr\": %.3f}%s\n", name, cyc,
         (double)h[1] / count, h[1] > 0 && base_cycles > 0 ? (cyc - base_cycles) * count / (double)h[1] : 0.0, MIX == 7 ? "" : ",");
}

int main() {
  const int count = 20000;
  double base = 0, c = 0;
  printf("{\n");
  run<0>(count, "idle", 0, &base);
  run<1>(count, "dfma_dense_8_chains", base, &c);
  run<2>(count, "dfma_then_4_ffma", base, &c);
  run<3>(count, "dfma_then_16_ffma", base, &c);
  run<4>(count, "ffma_only", base, &c);
  run<5>(count, "dfma_one_dependent_chain", base, &c);
  run<6>(count, "dmul_dadd", base, &c);
  run<7>(count, "imad_only", base, &c);
  printf("}\n");
  return 0;
}
