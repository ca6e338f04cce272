// HAR multi-head attention forward on the tensor cores (T <= 600 keys, d_head = 16): S = Q K^T and O = P V are
// tcgen05.mma kind::tf32 instructions with the scores living in TMEM; the softmax reads them with tcgen05.ld, applies
// the attention-weight dropout and writes P into shared memory in the swizzled operand layout of the second MMA, so
// neither S nor P ever reaches global memory ("flash-style single-CTA path for T <= 600").
//
// Reference: nn.MultiheadAttention inside nn.TransformerEncoderLayer (src/Model.py:441-446), i.e.
// softmax(q k^T / sqrt(16)) -> dropout(0.1) -> times v, for every (sample, head).
//
// One CTA = 128 queries of one (sample, head).  qkv is [B*T, 192] row-major; a TMA box {32 floats, 128 rows} holds the
// q (or k, or v) vectors of TWO heads, and head h uses the 64-byte half (h & 1) of every 128-byte row:
//   S block [128 q x 128 keys]  : A = Q tile, B = K block, both K-major SWIZZLE_128B, reduction = the 16 floats of head h
//   O      [128 q x 32]         : A = P (written by the softmax threads, K-major SWIZZLE_128B, reduction = keys),
//                                 B = V block as MN-major operand (TMA box with the 128B_ATOM_32B swizzle); the 32
//                                 output columns are both heads of the pair, the 16 of head h are kept.
// Two passes over the key blocks (row maximum first, then exp / sum / P.V), which costs a second, very cheap, S MMA
// but removes every rescale of the TMEM accumulator.  K and V stay resident in shared memory for both passes.
// 320 threads: warp 0 TMA, warp 1 MMA issue, warps 2-9 softmax (two warps per TMEM lane quarter, 64 keys of a block each).
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "common.h"
#include "har_rng.cuh"
#include "runtime.h"
#include "tc_common.cuh"

namespace {

using namespace afl_tc;
using namespace afl_har;

constexpr int kThreads = 320;
constexpr int kD = 64, kNH = 4, kDH = 16;
constexpr int kBlk = 128;                       // queries per CTA and keys per block
constexpr int kBlkBytes = kBlk * 128;           // one {32 floats, 128 rows} box

struct AttnParams {
  int T;               // tokens per sample
  int n_kb;            // key blocks = ceil(T / 128)
  int tail_rows;       // rows of the last key block actually loaded (multiple of 8)
  Rng rng;
  float* out;          // [B*T][64]
  float* lse;          // [B][4][lse_stride(T)], base-2 log-sum-exp of the scaled scores
  uint32_t* mask;      // [B][4][4*n_kb][lse_stride(T)] keep bits (NULL: not wanted), see attn_bwd_tcgen05.cu
};

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// v[16] (raw scores of keys kc .. kc+15) -> probabilities exp2(s c - m) with dropped entries zeroed (the kept ones are
// scaled by 1/(1-p) once, on the output row); returns the sum of the un-dropped probabilities.
template <bool MASK, bool DROP>
__device__ __forceinline__ float probs16(float* v, float c, float mx, int kc, int T, const Rng& hr, uint32_t row, uint32_t thr,
                                         uint32_t& keep_bits) {
  float sum = 0.f;
  keep_bits = 0xFFFFu;
#pragma unroll
  for (int g8 = 0; g8 < 2; ++g8) {
    AflU4 rw; rw.x = rw.y = rw.z = rw.w = 0u;
    if (DROP) rw = attn_call(hr, row, (uint32_t)(kc + 8 * g8));
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float p = ex2_approx(fmaf(v[8 * g8 + j], c, -mx));
      if (MASK && kc + 8 * g8 + j >= T) p = 0.f;
      sum += p;
      if (DROP && attn_bits(rw.x, rw.y, rw.z, rw.w, (uint32_t)j) < thr) { p = 0.f; keep_bits &= ~(1u << (8 * g8 + j)); }
      v[8 * g8 + j] = p;
    }
  }
  return sum;
}

__global__ void __launch_bounds__(kThreads, 1)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_kt,
                   const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_vt, AttnParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - smem_u32(smem_raw));
  const int n_kb = P.n_kb;
  const uint32_t kv_bytes = (uint32_t)(n_kb - 1) * kBlkBytes + (uint32_t)P.tail_rows * 128u;
  const uint32_t q_base = base;
  const uint32_t k_base = q_base + kBlkBytes;
  const uint32_t v_base = k_base + kv_bytes;
  const uint32_t p_base = (v_base + kv_bytes + 1023u) & ~1023u;          // 4 k-blocks of [128 rows][32 keys]
  uint8_t* p_ptr = aligned + (p_base - base);
  uint64_t* bars = reinterpret_cast<uint64_t*>(p_ptr + 4 * kBlkBytes);
  uint64_t* qk_full = bars;            // Q + all K blocks landed
  uint64_t* v_full = bars + 1;
  uint64_t* s_full = bars + 2;         // [2]
  uint64_t* s_empty = bars + 4;        // [2]
  uint64_t* p_full = bars + 6;
  uint64_t* p_empty = bars + 7;
  uint64_t* o_full = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
  float* xchg = reinterpret_cast<float*>(bars + 10);                     // [2][128] row max / row sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kBlk;
  const int hp = h >> 1;
  const uint32_t hoff = (uint32_t)(h & 1) * 64u;                         // byte offset of head h inside a 128-byte row
  const int row0 = b * P.T;
  constexpr uint32_t kTmemCols = 512;                                    // S x2 (256) + O (32) -> next power of two
  constexpr uint32_t kOCol = 256;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_k)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_v)) : "memory");
    mbar_init(qk_full, 1); mbar_init(v_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(s_full + i, 1); mbar_init(s_empty + i, 8); }
    mbar_init(p_full, 8); mbar_init(p_empty, 1); mbar_init(o_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA: Q tile + K blocks (SWIZZLE_128B), V blocks (128B_ATOM_32B) =====
      mbar_expect_tx(qk_full, kBlkBytes + kv_bytes);
      tma_load_2d(q_base, &tmap_k, hp * 32, row0 + q0, qk_full);
      for (int kb = 0; kb < n_kb; ++kb)
        tma_load_2d(k_base + (uint32_t)kb * kBlkBytes, kb + 1 < n_kb ? &tmap_k : &tmap_kt, kD + hp * 32, row0 + kb * kBlk, qk_full);
      mbar_expect_tx(v_full, kv_bytes);
      for (int kb = 0; kb < n_kb; ++kb)
        tma_load_2d(v_base + (uint32_t)kb * kBlkBytes, kb + 1 < n_kb ? &tmap_v : &tmap_vt, 2 * kD + hp * 32, row0 + kb * kBlk, v_full);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      const uint32_t idesc_s = idesc_tf32(128, kBlk, 0, 0);              // S: M = 128 queries, N = 128 keys
      const uint32_t idesc_o = idesc_tf32(128, 32, 0, 1);                // O: N = 32 (head pair), B MN-major
      const int n_s = 2 * n_kb;
      const uint64_t qd0 = desc_k_sw128(q_base + hoff), qd1 = desc_add(qd0, 32);
      const uint64_t kd0 = desc_k_sw128(k_base + hoff), kd1 = desc_add(kd0, 32);
      const uint64_t pd0 = desc_k_sw128(p_base);
      const uint64_t vd0 = desc_mn_sw128(v_base, kBlkBytes);
      auto issue_s = [&](int n) {
        const int buf = n & 1, kb = n % n_kb;
        mbar_wait(s_empty + buf, (((uint32_t)n >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t koff = (uint32_t)kb * kBlkBytes;
        umma_tf32(tmem_base + (uint32_t)buf * kBlk, qd0, desc_add(kd0, koff), idesc_s, 0u);
        umma_tf32(tmem_base + (uint32_t)buf * kBlk, qd1, desc_add(kd1, koff), idesc_s, 1u);
        umma_commit(s_full + buf);
      };
      mbar_wait(qk_full, 0);
      tc_fence_after();
      issue_s(0);
      for (int n = 0; n < n_s; ++n) {
        if (n + 1 < n_s) issue_s(n + 1);
        if (n >= n_kb) {
          const int kb = n - n_kb;
          if (kb == 0) { mbar_wait(v_full, 0); tc_fence_after(); }
          mbar_wait(p_full, (uint32_t)kb & 1u);
          tc_fence_after();
          const int nk8 = (kb + 1 < n_kb ? kBlk : P.tail_rows) / 8;
          const uint64_t vd = desc_add(vd0, (uint32_t)kb * kBlkBytes);
#pragma unroll
          for (int k8 = 0; k8 < kBlk / 8; ++k8)
            if (k8 < nk8)
              umma_tf32(tmem_base + kOCol, desc_add(pd0, (uint32_t)(k8 >> 2) * kBlkBytes + (uint32_t)(k8 & 3) * 32u),
                        desc_add(vd, (uint32_t)k8 * 1024u), idesc_o, (kb > 0 || k8 > 0) ? 1u : 0u);
          umma_commit(p_empty);
        }
      }
      umma_commit(o_full);
    }
  } else {
    // ===== softmax warps: row = TMEM lane, key half = (warp - 2) / 4 =====
    const int sw = warp - 2;
    const int q = warp & 3;                    // TMEM lane quarter of this warp (warp id % 4)
    const int half = sw >> 2;                  // which 64 keys of a 128-key block
    const int r = q * 32 + lane;               // query row inside the tile
    const int qi = q0 + r;
    const float c = 0.25f * 1.4426950408889634f;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t thr = attn_thr(P.rng.p);
    const float inv_keep = P.rng.p > 0.f ? 1.0f / (1.0f - P.rng.p) : 1.f;
    Rng hr = P.rng; hr.site = P.rng.site + (uint32_t)h;
    // ---- pass 1: row maximum
    float mx = -INFINITY;
    for (int n = 0; n < n_kb; ++n) {
      const int buf = n & 1;
      mbar_wait(s_full + buf, ((uint32_t)n >> 1) & 1u);
      tc_fence_after();
      const int key0 = n * kBlk + half * 64;
      float v[64];
      tmem_ld32_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(half * 64), v);
      tmem_ld32_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(half * 64 + 32), v + 32);
      tmem_ld_wait();
      if (key0 + 64 <= P.T) {                   // every key of this half is real: no masking
#pragma unroll
        for (int j = 0; j < 64; ++j) mx = fmaxf(mx, v[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 64; ++j) if (key0 + j < P.T) mx = fmaxf(mx, v[j]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty + buf);
    }
    xchg[half * 128 + r] = mx;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    mx = fmaxf(mx, xchg[(half ^ 1) * 128 + r]) * c;      // row maximum in scaled base-2 units
    asm volatile("bar.sync 1, 256;" ::: "memory");
    // ---- pass 2: p = exp2(s c - m), row sum, dropout, P -> shared memory (operand layout), P V on the tensor core
    float l = 0.f;
    for (int kb = 0; kb < n_kb; ++kb) {
      const int n = n_kb + kb, buf = n & 1;
      mbar_wait(s_full + buf, ((uint32_t)n >> 1) & 1u);
      tc_fence_after();
      const int key0 = kb * kBlk + half * 64;
      const bool full = key0 + 64 <= P.T;
      uint32_t keep_lo = 0;
      float v64[64];
      tmem_ld32_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(half * 64), v64);
      tmem_ld32_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(half * 64 + 32), v64 + 32);
      tmem_ld_wait();
      // all 64 probabilities first (registers), so that this math overlaps the previous block's P V; only then wait for
      // the P buffer and write the operand tile
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float* v = v64 + c0;
        const int kc = key0 + c0;
        uint32_t kb16;
        if (full) {
          if (P.rng.p > 0.f) l += probs16<false, true>(v, c, mx, kc, P.T, hr, (uint32_t)(row0 + qi), thr, kb16);
          else l += probs16<false, false>(v, c, mx, kc, P.T, hr, (uint32_t)(row0 + qi), thr, kb16);
        } else {
          if (P.rng.p > 0.f) l += probs16<true, true>(v, c, mx, kc, P.T, hr, (uint32_t)(row0 + qi), thr, kb16);
          else l += probs16<true, false>(v, c, mx, kc, P.T, hr, (uint32_t)(row0 + qi), thr, kb16);
        }
        // keep-mask words for the backward kernels: bit (key & 31) of word [b][h][key >> 5][query]
        if (c0 & 16) {
          if (P.mask && qi < P.T)
            P.mask[(((size_t)b * kNH + h) * (size_t)(n_kb * 4) + (size_t)(kc >> 5)) * lse_stride(P.T) + qi] = keep_lo | (kb16 << 16);
        } else {
          keep_lo = kb16;
        }
      }
      mbar_wait(p_empty, ((uint32_t)kb & 1u) ^ 1u);       // previous block's P V has consumed the P buffer
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        const float* v = v64 + c0;
        // keys (half*64 + c0 .. +15) of this block: k-block (32 keys) index and 16-byte chunk inside the 128-byte row
        const int kcol = half * 64 + c0;
        uint8_t* rowp = p_ptr + (size_t)(kcol >> 5) * kBlkBytes + (size_t)r * 128;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = ((kcol & 31) >> 2) + g;
          *reinterpret_cast<float4*>(rowp + ((chunk ^ (r & 7)) << 4)) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
      }
      tc_fence_before();
      fence_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(s_empty + buf); mbar_arrive(p_full); }
    }
    xchg[half * 128 + r] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    l += xchg[(half ^ 1) * 128 + r];
    // ---- epilogue: O / l  (half 0 warps write the output row, half 1 warps the log-sum-exp)
    mbar_wait(o_full, 0);
    tc_fence_after();
    if (half == 0) {
      float o[16];
      tmem_ld16(lane_addr + kOCol + (uint32_t)(h & 1) * 16u, o);
      if (qi < P.T) {
        const float il = inv_keep / l;
        float4* dst = reinterpret_cast<float4*>(P.out + (size_t)(row0 + qi) * kD + h * kDH);
#pragma unroll
        for (int g = 0; g < 4; ++g) dst[g] = make_float4(o[4 * g] * il, o[4 * g + 1] * il, o[4 * g + 2] * il, o[4 * g + 3] * il);
      }
    } else if (qi < P.T) {
      P.lse[((size_t)b * kNH + h) * lse_stride(P.T) + qi] = mx + log2f(l);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

extern "C" {

// Returns 1 when the tensor-core kernel can take (T): K and V of one (sample, head pair) must fit in shared memory.
int afl_har_attn_tc_supported(int T) {
  if (T < 1) return 0;
  const int n_kb = (T + kBlk - 1) / kBlk;
  const int tail = ((T - (n_kb - 1) * kBlk) + 7) / 8 * 8;
  const size_t kv = (size_t)(n_kb - 1) * kBlkBytes + (size_t)tail * 128;
  const size_t need = 1024 + kBlkBytes + 2 * kv + 1024 + 4 * (size_t)kBlkBytes + 10 * 8 + 2 * 128 * 4 + 64;
  return need <= (size_t)afl_rt::max_smem_optin() ? 1 : 0;
}

int afl_har_attn_fwd_tc(const float* qkv, int B, int T, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                        uint32_t site, float* out, float* lse, uint32_t* keep_mask, void* stream) {
  AFL_CHECK_ARG(B >= 1 && afl_har_attn_tc_supported(T), "afl_har_attn_fwd_tc: T too large for the shared-memory resident K/V");
  AFL_CHECK_ARG((uintptr_t)qkv % 16 == 0, "afl_har_attn_fwd_tc: qkv must be 16-byte aligned");
  const int n_kb = (T + kBlk - 1) / kBlk;
  const int tail = ((T - (n_kb - 1) * kBlk) + 7) / 8 * 8;
  const int64_t rows = (int64_t)B * T;
  CUtensorMap mk, mkt, mv, mvt;
  int rc = make_tmap_f32(&mk, qkv, 3 * kD, rows, 3 * kD, kBlk, false);
  if (!rc) rc = make_tmap_f32(&mkt, qkv, 3 * kD, rows, 3 * kD, tail, false);
  if (!rc) rc = make_tmap_f32(&mv, qkv, 3 * kD, rows, 3 * kD, kBlk, true);
  if (!rc) rc = make_tmap_f32(&mvt, qkv, 3 * kD, rows, 3 * kD, tail, true);
  if (rc) return rc;
  AttnParams P{};
  P.T = T; P.n_kb = n_kb; P.tail_rows = tail; P.out = out; P.lse = lse; P.mask = drop_p > 0.f ? keep_mask : nullptr;
  P.rng.s0 = (uint32_t)seed; P.rng.s1 = (uint32_t)(seed >> 32); P.rng.client = client; P.rng.step = step; P.rng.site = site;
  P.rng.p = drop_p;
  const size_t kv = (size_t)(n_kb - 1) * kBlkBytes + (size_t)tail * 128;
  const size_t smem = 1024 + kBlkBytes + 2 * kv + 1024 + 4 * (size_t)kBlkBytes + 10 * 8 + 2 * 128 * 4 + 64;
  AFL_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((T + kBlk - 1) / kBlk, kNH, B);
  attn_fwd_tc_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(mk, mkt, mv, mvt, P);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_attn_fwd_tc");
}

}  // extern "C"
