// Dropout random streams shared by the HAR kernels (har_kernels.cu, attn_tcgen05.cu).
#pragma once
#include <stdint.h>

#include "common.h"

namespace afl_har {

// Row stride of the per-(sample, head) arrays lse / delta / keep-mask words: T rounded up to 64 so that every 64-query
// block starts on a 16-byte boundary (bulk-copy alignment) and reads past T stay inside the row.
__host__ __device__ inline int lse_stride(int T) { return (T + 63) / 64 * 64; }

struct Rng { uint32_t s0, s1, client, step, site; float p; };

// Element-wise dropout (LayerNorm residuals, FFN, head): Philox call (row, col >> 2), component col & 3, 24-bit uniform.
__device__ __forceinline__ AflU4 rng4(const Rng& r, uint32_t row, uint32_t col) {
  return afl_philox(col >> 2, (r.site & 0xFFFFu) | (r.client << 16), r.step, row, r.s0, r.s1);
}
__device__ __forceinline__ float u01(uint32_t v) { return (float)(v >> 8) * (1.0f / 16777216.0f); }

// Attention-weight dropout (Philox4x32-7): 16 bits per decision, so one Philox call covers the 8 keys (key & ~7 .. key | 7) of a query
// row: component (key >> 1) & 3, half key & 1; keep iff the 16-bit value >= thr = round(p * 65536).
__device__ __forceinline__ AflU4 attn_call(const Rng& r, uint32_t row, uint32_t key) {
  return afl_philox_r<7>(key >> 3, (r.site & 0xFFFFu) | (r.client << 16), r.step, row, r.s0, r.s1);
}
__device__ __forceinline__ uint32_t attn_bits(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t key) {
  const uint32_t c = (key >> 1) & 3u;
  const uint32_t word = c == 0 ? x : c == 1 ? y : c == 2 ? z : w;
  return (word >> (16u * (key & 1u))) & 0xFFFFu;
}
__device__ __forceinline__ uint32_t attn_thr(float p) { return (uint32_t)(p * 65536.0f + 0.5f); }

}  // namespace afl_har
