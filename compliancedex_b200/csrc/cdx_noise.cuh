// Keyed uniform stream for the Kabsch perturbation - device twin of oracle/noise.py (same constants, same order).
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#define CDX_NOISE_HD __host__ __device__ __forceinline__
#else
#define CDX_NOISE_HD inline
#endif
CDX_NOISE_HD uint64_t cdx_mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// u in [0,1): seed, iteration, copy p, GLOBAL candidate b, element 0..8 of the 3x3
CDX_NOISE_HD double cdx_kabsch_noise(uint64_t seed, int it, int p, int64_t b, int elem) {
  uint64_t h0 = cdx_mix64(seed ^ (0x9E3779B97F4A7C15ull * (uint64_t)(it + 1)));
  uint64_t h1 = cdx_mix64(h0 ^ ((uint64_t)b * 0xBF58476D1CE4E5B9ull));
  uint64_t h2 = cdx_mix64(h1 ^ ((uint64_t)p * 16ull + (uint64_t)elem));
  return (double)(h2 >> 11) * (1.0 / 9007199254740992.0);
}
