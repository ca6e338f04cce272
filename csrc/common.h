// Shared host/device helpers for the attackfl_b200 native library.
//
// Every numerical routine of the library is written once, as a function of (tid, nthreads), in
// headers that compile both with nvcc for sm_100a (tid = threadIdx.x) and with g++ (tid = 0,
// nthreads = 1).  The g++ build ("CPU twin", csrc/cpu_backend.cpp) is what the non-GPU test suite
// runs, so the kernel source itself is exercised on machines without a GPU; it is also the
// `--device cpu` backend of the framework.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define AFL_DEV __device__ __forceinline__
#define AFL_HD __host__ __device__ __forceinline__
#define AFL_SYNC() __syncthreads()
#else
#define AFL_DEV inline
#define AFL_HD inline
#define AFL_SYNC() ((void)0)
struct alignas(16) float4 { float x, y, z, w; };
#endif

// L2 -> L1 prefetch of one 128-byte line (device only; a no-op in the CPU twin)
#if defined(__CUDA_ARCH__)
#define AFL_PREFETCH_L1(ptr) asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr))
#else
#define AFL_PREFETCH_L1(ptr) ((void)0)
#endif

#define AFL_MIN(a, b) ((a) < (b) ? (a) : (b))
#define AFL_MAX(a, b) ((a) > (b) ? (a) : (b))

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al.), counter-based: identical streams on host and device, so dropout
// masks are a pure function of (seed, client, step, sample, op, element) and never stored.
// ---------------------------------------------------------------------------------------------
struct AflU4 { uint32_t x, y, z, w; };

AFL_HD uint32_t afl_mulhi(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

// Philox4x32-R (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3"): R = 10 is the standard generator, R = 7 the
// smallest round count that the paper reports as passing BigCrush (used only for the attention-weight dropout bits).
template <int R>
AFL_HD AflU4 afl_philox_r(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int r = 0; r < R; ++r) {
    uint32_t hi0 = afl_mulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = afl_mulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += W0; k1 += W1;
  }
  AflU4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}

AFL_HD AflU4 afl_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  return afl_philox_r<10>(c0, c1, c2, c3, k0, k1);
}

// Uniform in [0,1) for element `elem` of stream (client, step, sample, rng_id).
AFL_HD float afl_uniform(uint32_t seed_lo, uint32_t seed_hi, uint32_t client, uint32_t step,
                         uint32_t sample, uint32_t rng_id, uint32_t elem) {
  AflU4 r = afl_philox(elem >> 2, (rng_id & 0xFFFFu) | (client << 16), step, sample, seed_lo, seed_hi);
  uint32_t v = (elem & 3u) == 0 ? r.x : (elem & 3u) == 1 ? r.y : (elem & 3u) == 2 ? r.z : r.w;
  return (float)(v >> 8) * (1.0f / 16777216.0f);
}

// Dropout keep decision: keep iff u >= p.
AFL_HD bool afl_keep(uint32_t seed_lo, uint32_t seed_hi, uint32_t client, uint32_t step,
                     uint32_t sample, uint32_t rng_id, uint32_t elem, float p) {
  return afl_uniform(seed_lo, seed_hi, client, step, sample, rng_id, elem) >= p;
}

// Standard normal via Box-Muller on two Philox words (Random attack, kernel K4).
AFL_HD float afl_normal(uint32_t seed_lo, uint32_t seed_hi, uint32_t stream, uint32_t idx) {
  AflU4 r = afl_philox(idx, stream, 0x5EEDu, 0u, seed_lo, seed_hi);
  float u1 = ((float)(r.x >> 8) + 1.0f) * (1.0f / 16777216.0f);   // (0,1]
  float u2 = (float)(r.y >> 8) * (1.0f / 16777216.0f);
  return sqrtf(-2.0f * logf(u1)) * cosf(6.28318530717958647692f * u2);
}

AFL_HD float afl_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }
AFL_HD float afl_gelu(float z) { return 0.5f * z * (1.0f + erff(z * 0.70710678118654752440f)); }
AFL_HD float afl_gelu_grad(float z) {
  return 0.5f * (1.0f + erff(z * 0.70710678118654752440f)) + z * 0.39894228040143267794f * expf(-0.5f * z * z);
}
