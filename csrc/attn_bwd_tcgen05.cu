// HAR multi-head attention backward on the tensor cores (companion of attn_tcgen05.cu; d_head = 16, T tokens/sample).
//
// With S = q k^T / 4, P = softmax(S), Pd = dropout(P), O = Pd v:
//   dPd = dO v^T,  dS = P * (keep/(1-p) * dPd - delta),  delta_i = dO_i . O_i,
//   dQ = dS k / 4,  dK = dS^T q / 4,  dV = Pd^T dO.
// Two kernels, every product a tcgen05.mma kind::tf32 with TMEM accumulators, all operands either TMA tiles of qkv /
// dout (K-major: SWIZZLE_128B box; MN-major: 128B_ATOM_32B box of the same rows — no transposes in memory) or the dS /
// Pd tiles the element-wise warps write into shared memory in the K-major operand layout:
//   dq  kernel: CTA = 128 queries (thread row = query), streams 64-key blocks:
//               S = Q K^T, dPd = dO V^T (TMEM) -> dS -> smem -> dQ += dS K (B = K block MN-major)
//   dkv kernel: CTA = 128 keys (thread row = key), streams 64-query blocks:
//               S^T = K Q^T, dPd^T = V dO^T (TMEM) -> Pd^T, dS^T -> smem -> dV += Pd^T dO, dK += dS^T Q (B MN-major)
// P is recomputed from the forward's base-2 log-sum-exp; the dropout keep bits come from the mask words the forward
// kernel stored (bit (key & 31) of word [key >> 5][query]), so no random numbers are regenerated here.
// 64 + 128 * kParts threads per CTA: warp 0 TMA, warp 1 MMA issue, the rest element-wise (kParts warps per TMEM lane
// quarter, 64 / kParts columns of a streamed block each).
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "har_rng.cuh"
#include "runtime.h"
#include "tc_common.cuh"

namespace {

using namespace afl_tc;
using namespace afl_har;

// Element-wise warps per TMEM lane quarter: each owns 64 / kParts columns of a streamed block.  Measured (B200, batch 128,
// dropout 0.1): 2 (8 warps, 320 threads) 468 us per layer, 4 (16 warps, 576 threads) 499 us — the per-block fixed costs of
// a warp (TMEM load latency, proxy fence, barrier round trips) do not shrink with its column count, so more warps only
// add arrivals; 2 is the default.
#ifndef AFL_ATTN_PARTS
#define AFL_ATTN_PARTS 2
#endif
constexpr int kParts = AFL_ATTN_PARTS;
constexpr int kCW = 64 / kParts;                // columns per element-wise warp
constexpr int kThreads = 64 + 128 * kParts;
constexpr int kD = 64, kNH = 4, kDH = 16;
constexpr int kTile = 128;                      // rows per CTA (queries or keys)
constexpr int kBlk = 64;                        // streamed block (keys or queries)
constexpr int kTileBytes = kTile * 128;         // {32 floats, 128 rows}
constexpr int kBlkBytes = kBlk * 128;           // {32 floats, 64 rows}
constexpr float kScaleLog2e = 0.25f * 1.4426950408889634f;

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_dst), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// this thread's 16-float row segment `seg` (0..1) of k-block `kblk` in a K-major SWIZZLE_128B operand tile [rows][32]
__device__ __forceinline__ void store16_sw128(uint8_t* tile, int kblk, int kblk_bytes, int row, int seg, const float* v) {
  uint8_t* rowp = tile + (size_t)kblk * kblk_bytes + (size_t)row * 128;
#pragma unroll
  for (int g = 0; g < 4; ++g)
    *reinterpret_cast<float4*>(rowp + (((seg * 4 + g) ^ (row & 7)) << 4)) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
}

// delta[b][h][i] = dO_i . O_i over the 16 columns of head h
__global__ void __launch_bounds__(256) attn_delta_kernel(const float* __restrict__ out, const float* __restrict__ dout, int B, int T,
                                                         float* __restrict__ delta) {
  const int64_t n = (int64_t)B * T * kNH;
  const int Tp = lse_stride(T);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    const int64_t row = i >> 2;
    const int h = (int)(i & 3);
    const float4* o = reinterpret_cast<const float4*>(out + row * kD + h * kDH);
    const float4* g = reinterpret_cast<const float4*>(dout + row * kD + h * kDH);
    float a = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 x = o[k], y = g[k];
      a += x.x * y.x + x.y * y.y + x.z * y.z + x.w * y.w;
    }
    const int b = (int)(row / T), t = (int)(row - (int64_t)b * T);
    delta[((int64_t)b * kNH + h) * Tp + t] = a;
  }
}

struct BwdParams {
  int T, n_blk;              // tokens per sample, streamed 64-row blocks = ceil(T / 64)
  int mask_words;            // words per (sample, head, query) row group in keep_mask (multiple of 4)
  float drop_p;
  const float* lse;          // [B][4][Tp]
  const float* delta;        // [B][4][Tp]
  const uint32_t* mask;      // [B][4][mask_words][Tp] or NULL
  float* dqkv;               // [B*T][192]
};

// ------------------------------------------------------------------------------------------------------------ dQ
// stage = {K block K-major, K block MN-major, V block K-major}
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tm_q128, const __grid_constant__ CUtensorMap tm_o128,
                      const __grid_constant__ CUtensorMap tm_k64, const __grid_constant__ CUtensorMap tm_k64t, BwdParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int kStages = 3;
  constexpr uint32_t kStageBytes = 3 * kBlkBytes;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t q_base = base, do_base = base + kTileBytes, st_base = do_base + kTileBytes;
  const uint32_t ds_base = st_base + kStages * kStageBytes;               // 2 buffers of [128][64 keys] = 2 k-blocks each
  uint8_t* ds_ptr = aligned + (ds_base - base);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ds_ptr + 2 * 2 * kTileBytes);
  uint64_t* qd_full = bars;
  uint64_t* kv_full = bars + 1;          // [3]
  uint64_t* kv_empty = bars + 4;         // [3]
  uint64_t* sd_full = bars + 7;          // [2]  S / dPd accumulators ready
  uint64_t* sd_empty = bars + 9;         // [2]
  uint64_t* ds_full = bars + 11;         // [2]  dS tile written
  uint64_t* ds_empty = bars + 13;        // [2]
  uint64_t* dq_full = bars + 15;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kTile;
  const int hp = h >> 1;
  const uint32_t hoff = (uint32_t)(h & 1) * 64u;
  const int row0 = b * P.T;
  const int n_blk = P.n_blk;
  constexpr uint32_t kTmemCols = 512;        // S[2] 0,64 | dPd[2] 128,192 | dQ 256
  constexpr uint32_t kDqCol = 256;

  if (warp == 0 && lane == 0) {
    mbar_init(qd_full, 1);
    for (int i = 0; i < kStages; ++i) { mbar_init(kv_full + i, 1); mbar_init(kv_empty + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(sd_full + i, 1); mbar_init(sd_empty + i, 4 * kParts); mbar_init(ds_full + i, 4 * kParts); mbar_init(ds_empty + i, 1); }
    mbar_init(dq_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(qd_full, 2 * kTileBytes);
      tma_load_2d(q_base, &tm_q128, hp * 32, row0 + q0, qd_full);
      tma_load_2d(do_base, &tm_o128, hp * 32, row0 + q0, qd_full);
      for (int kb = 0; kb < n_blk; ++kb) {
        const int st = kb % kStages;
        mbar_wait(kv_empty + st, (((uint32_t)(kb / kStages)) & 1u) ^ 1u);
        mbar_expect_tx(kv_full + st, kStageBytes);
        const uint32_t s_addr = st_base + (uint32_t)st * kStageBytes;
        tma_load_2d(s_addr, &tm_k64, kD + hp * 32, row0 + kb * kBlk, kv_full + st);
        tma_load_2d(s_addr + kBlkBytes, &tm_k64t, kD + hp * 32, row0 + kb * kBlk, kv_full + st);
        tma_load_2d(s_addr + 2 * kBlkBytes, &tm_k64, 2 * kD + hp * 32, row0 + kb * kBlk, kv_full + st);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = idesc_tf32(128, kBlk, 0, 0);
      const uint32_t idesc_q = idesc_tf32(128, 32, 0, 1);
      const uint64_t qd0 = desc_k_sw128(q_base + hoff), qd1 = desc_add(qd0, 32);
      const uint64_t od0 = desc_k_sw128(do_base + hoff), od1 = desc_add(od0, 32);
      const uint64_t kd0 = desc_k_sw128(st_base + hoff);                       // + stage offset (+ 2 boxes for V)
      const uint64_t ad0 = desc_k_sw128(ds_base);                              // dS tile, + buffer / k-step offset
      const uint64_t bd0 = desc_mn_sw128(st_base + kBlkBytes, kBlkBytes);      // K block MN-major, + stage / k-step offset
      auto issue_dq = [&](int kb) {          // dQ += dS(kb) . K(kb)
        const int dbuf = kb & 1, st = kb % kStages;
        mbar_wait(ds_full + dbuf, ((uint32_t)kb >> 1) & 1u);
        tc_fence_after();
        const uint64_t ad = desc_add(ad0, (uint32_t)dbuf * 2 * kTileBytes);
        const uint64_t bd = desc_add(bd0, (uint32_t)st * kStageBytes);
#pragma unroll
        for (int k8 = 0; k8 < kBlk / 8; ++k8)
          umma_tf32(tmem_base + kDqCol, desc_add(ad, (uint32_t)(k8 >> 2) * kTileBytes + (uint32_t)(k8 & 3) * 32u),
                    desc_add(bd, (uint32_t)k8 * 1024u), idesc_q, (kb > 0 || k8 > 0) ? 1u : 0u);
        umma_commit(ds_empty + dbuf);
        umma_commit(kv_empty + st);
      };
      mbar_wait(qd_full, 0);
      for (int kb = 0; kb < n_blk; ++kb) {
        const int st = kb % kStages, buf = kb & 1;
        mbar_wait(kv_full + st, ((uint32_t)(kb / kStages)) & 1u);
        mbar_wait(sd_empty + buf, (((uint32_t)kb >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint64_t kd = desc_add(kd0, (uint32_t)st * kStageBytes);
        const uint64_t vd = desc_add(kd, 2 * kBlkBytes);
        const uint32_t s_col = tmem_base + (uint32_t)buf * kBlk;
        umma_tf32(s_col, qd0, kd, idesc_s, 0u);
        umma_tf32(s_col, qd1, desc_add(kd, 32), idesc_s, 1u);
        umma_tf32(s_col + 128u, od0, vd, idesc_s, 0u);
        umma_tf32(s_col + 128u, od1, desc_add(vd, 32), idesc_s, 1u);
        umma_commit(sd_full + buf);
        if (kb > 0) issue_dq(kb - 1);
      }
      issue_dq(n_blk - 1);
      umma_commit(dq_full);
    }
  } else {
    const int q = warp & 3, part = (warp - 2) >> 2;
    const int col0 = part * kCW;                            // this warp's columns of a 64-key block
    const int r = q * 32 + lane, qi = q0 + r;
    const bool live = qi < P.T;
    const int Tp = lse_stride(P.T);
    const size_t bh = (size_t)b * kNH + h;
    const float L = live ? P.lse[bh * Tp + qi] : 0.f;
    const float delta = live ? P.delta[bh * Tp + qi] : 0.f;
    const bool drop = P.drop_p > 0.f && P.mask != nullptr;
    const float inv_keep = P.drop_p > 0.f ? 1.0f / (1.0f - P.drop_p) : 1.f;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int wsel = col0 >> 5, wshift = col0 & 31;         // keep-mask word of the block's 2 and bit offset inside it
    uint32_t keep_next = 0xFFFFFFFFu;
    if (drop && live) keep_next = P.mask[(bh * P.mask_words + (size_t)wsel) * Tp + qi];
    for (int kb = 0; kb < n_blk; ++kb) {
      const int buf = kb & 1;
      const uint32_t keep = keep_next >> wshift;            // the next block's word is fetched one block ahead
      if (drop && live && kb + 1 < n_blk) keep_next = P.mask[(bh * P.mask_words + (size_t)((kb + 1) * 2 + wsel)) * Tp + qi];
      mbar_wait(sd_full + buf, ((uint32_t)kb >> 1) & 1u);
      tc_fence_after();
      const int key0 = kb * kBlk + col0;
      float ds[kCW], dd[kCW];                                 // scores -> dS in place; dPd
#pragma unroll
      for (int c = 0; c < kCW; c += 16) {
        tmem_ld16_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(col0 + c), ds + c);
        tmem_ld16_async(lane_addr + 128u + (uint32_t)buf * kBlk + (uint32_t)(col0 + c), dd + c);
      }
      tmem_ld_wait();
      const bool full = key0 + kCW <= P.T;
#pragma unroll
      for (int j = 0; j < kCW; ++j) {
        const float p = ex2_approx(fmaf(ds[j], kScaleLog2e, -L));
        const float kp = ((keep >> j) & 1u) ? inv_keep : 0.f;
        float v = p * fmaf(kp, dd[j], -delta);
        if (!full && key0 + j >= P.T) v = 0.f;
        ds[j] = v;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(sd_empty + buf);
      mbar_wait(ds_empty + buf, (((uint32_t)kb >> 1) & 1u) ^ 1u);
      uint8_t* tile = ds_ptr + (size_t)buf * 2 * kTileBytes;
#pragma unroll
      for (int c = 0; c < kCW; c += 16) store16_sw128(tile, (col0 + c) >> 5, kTileBytes, r, ((col0 + c) & 31) >> 4, ds + c);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_full + buf);
    }
    mbar_wait(dq_full, 0);
    tc_fence_after();
    if (part == 0) {
      float o[16];
      tmem_ld16(lane_addr + kDqCol + (uint32_t)(h & 1) * 16u, o);
      if (live) {
        float4* dst = reinterpret_cast<float4*>(P.dqkv + (size_t)(row0 + qi) * (3 * kD) + h * kDH);
#pragma unroll
        for (int g = 0; g < 4; ++g) dst[g] = make_float4(o[4 * g] * 0.25f, o[4 * g + 1] * 0.25f, o[4 * g + 2] * 0.25f, o[4 * g + 3] * 0.25f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------------------ dK, dV
// stage = {Q block K-major, dO block K-major, Q block MN-major, dO block MN-major, lse[64], delta[64], mask[4][64]}
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tm_k128, const __grid_constant__ CUtensorMap tm_q64,
                       const __grid_constant__ CUtensorMap tm_q64t, const __grid_constant__ CUtensorMap tm_o64,
                       const __grid_constant__ CUtensorMap tm_o64t, BwdParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int kStages = 3;
  constexpr uint32_t kAux = 2048;                                          // lse 256 + delta 256 + mask 1024 (+ pad)
  constexpr uint32_t kStageBytes = 4 * kBlkBytes + kAux;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t k_base = base, v_base = base + kTileBytes, st_base = v_base + kTileBytes;
  const uint32_t pd_base = st_base + kStages * kStageBytes;                // Pd^T [128 keys][64 q] = 2 k-blocks of 16 KB
  const uint32_t dst_base = pd_base + 2 * kTileBytes;                      // dS^T, same shape
  uint8_t* pd_ptr = aligned + (pd_base - base);
  uint8_t* dst_ptr = aligned + (dst_base - base);
  uint64_t* bars = reinterpret_cast<uint64_t*>(dst_ptr + 2 * kTileBytes);
  uint64_t* kv_full = bars;
  uint64_t* q_full = bars + 1;           // [3]
  uint64_t* q_empty = bars + 4;          // [3]
  uint64_t* sd_full = bars + 7;          // [2]
  uint64_t* sd_empty = bars + 9;         // [2]
  uint64_t* pd_full = bars + 11;
  uint64_t* pd_empty = bars + 12;
  uint64_t* acc_full = bars + 13;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * kTile;
  const int hp = h >> 1;
  const uint32_t hoff = (uint32_t)(h & 1) * 64u;
  const int row0 = b * P.T;
  const int n_blk = P.n_blk;
  const int Tp = lse_stride(P.T);
  const size_t bh = (size_t)b * kNH + h;
  const bool drop = P.drop_p > 0.f && P.mask != nullptr;
  constexpr uint32_t kTmemCols = 512;        // S^T[2] 0,64 | dPd^T[2] 128,192 | dK 256 | dV 288
  constexpr uint32_t kDkCol = 256, kDvCol = 288;

  if (warp == 0 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < kStages; ++i) { mbar_init(q_full + i, 1); mbar_init(q_empty + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(sd_full + i, 1); mbar_init(sd_empty + i, 4 * kParts); }
    mbar_init(pd_full, 4 * kParts); mbar_init(pd_empty, 1); mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * kTileBytes);
      tma_load_2d(k_base, &tm_k128, kD + hp * 32, row0 + k0, kv_full);
      tma_load_2d(v_base, &tm_k128, 2 * kD + hp * 32, row0 + k0, kv_full);
      for (int ib = 0; ib < n_blk; ++ib) {
        const int st = ib % kStages;
        mbar_wait(q_empty + st, (((uint32_t)(ib / kStages)) & 1u) ^ 1u);
        const uint32_t aux_bytes = 512u + (drop ? 1024u : 0u);
        mbar_expect_tx(q_full + st, 4 * kBlkBytes + aux_bytes);
        const uint32_t s_addr = st_base + (uint32_t)st * kStageBytes;
        const int r0 = row0 + ib * kBlk;
        tma_load_2d(s_addr, &tm_q64, hp * 32, r0, q_full + st);
        tma_load_2d(s_addr + kBlkBytes, &tm_o64, hp * 32, r0, q_full + st);
        tma_load_2d(s_addr + 2 * kBlkBytes, &tm_q64t, hp * 32, r0, q_full + st);
        tma_load_2d(s_addr + 3 * kBlkBytes, &tm_o64t, hp * 32, r0, q_full + st);
        const uint32_t aux = s_addr + 4 * kBlkBytes;
        bulk_load(aux, P.lse + bh * Tp + ib * kBlk, 256, q_full + st);
        bulk_load(aux + 256, P.delta + bh * Tp + ib * kBlk, 256, q_full + st);
        if (drop)
          for (int w = 0; w < 4; ++w)
            bulk_load(aux + 512 + (uint32_t)w * 256, P.mask + (bh * P.mask_words + (size_t)(k0 / 32 + w)) * Tp + ib * kBlk, 256,
                      q_full + st);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = idesc_tf32(128, kBlk, 0, 0);
      const uint32_t idesc_g = idesc_tf32(128, 32, 0, 1);
      const uint64_t kd0 = desc_k_sw128(k_base + hoff), kd1 = desc_add(kd0, 32);
      const uint64_t vd0 = desc_k_sw128(v_base + hoff), vd1 = desc_add(vd0, 32);
      const uint64_t qd0 = desc_k_sw128(st_base + hoff);                       // Q block K-major (+ stage; dO: + 1 box)
      const uint64_t pa0 = desc_k_sw128(pd_base), sa0 = desc_k_sw128(dst_base);
      const uint64_t qt0 = desc_mn_sw128(st_base + 2 * kBlkBytes, kBlkBytes);  // Q block MN-major (+ stage; dO: + 1 box)
      auto issue_grad = [&](int ib) {        // dV += Pd^T(ib) . dO(ib);  dK += dS^T(ib) . Q(ib)
        const int st = ib % kStages;
        mbar_wait(pd_full, (uint32_t)ib & 1u);
        tc_fence_after();
        const uint64_t qt = desc_add(qt0, (uint32_t)st * kStageBytes);
        const uint64_t ot = desc_add(qt, kBlkBytes);
#pragma unroll
        for (int k8 = 0; k8 < kBlk / 8; ++k8) {
          const uint32_t a_off = (uint32_t)(k8 >> 2) * kTileBytes + (uint32_t)(k8 & 3) * 32u;
          umma_tf32(tmem_base + kDvCol, desc_add(pa0, a_off), desc_add(ot, (uint32_t)k8 * 1024u), idesc_g, (ib > 0 || k8 > 0) ? 1u : 0u);
          umma_tf32(tmem_base + kDkCol, desc_add(sa0, a_off), desc_add(qt, (uint32_t)k8 * 1024u), idesc_g, (ib > 0 || k8 > 0) ? 1u : 0u);
        }
        umma_commit(pd_empty);
        umma_commit(q_empty + st);
      };
      mbar_wait(kv_full, 0);
      for (int ib = 0; ib < n_blk; ++ib) {
        const int st = ib % kStages, buf = ib & 1;
        mbar_wait(q_full + st, ((uint32_t)(ib / kStages)) & 1u);
        mbar_wait(sd_empty + buf, (((uint32_t)ib >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint64_t qd = desc_add(qd0, (uint32_t)st * kStageBytes);
        const uint64_t od = desc_add(qd, kBlkBytes);
        const uint32_t s_col = tmem_base + (uint32_t)buf * kBlk;
        umma_tf32(s_col, kd0, qd, idesc_s, 0u);
        umma_tf32(s_col, kd1, desc_add(qd, 32), idesc_s, 1u);
        umma_tf32(s_col + 128u, vd0, od, idesc_s, 0u);
        umma_tf32(s_col + 128u, vd1, desc_add(od, 32), idesc_s, 1u);
        umma_commit(sd_full + buf);
        if (ib > 0) issue_grad(ib - 1);
      }
      issue_grad(n_blk - 1);
      umma_commit(acc_full);
    }
  } else {
    const int q = warp & 3, part = (warp - 2) >> 2;
    const int col0 = part * kCW;                            // this warp's columns (queries) of a 64-query block
    const int r = q * 32 + lane, kj = k0 + r;
    const bool live = kj < P.T;
    const float inv_keep = P.drop_p > 0.f ? 1.0f / (1.0f - P.drop_p) : 1.f;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int ib = 0; ib < n_blk; ++ib) {
      const int buf = ib & 1, st = ib % kStages;
      mbar_wait(q_full + st, ((uint32_t)(ib / kStages)) & 1u); // lse / delta / mask rows of this block are in smem
      mbar_wait(sd_full + buf, ((uint32_t)ib >> 1) & 1u);
      tc_fence_after();
      const uint8_t* aux = aligned + (st_base - base) + (size_t)st * kStageBytes + 4 * kBlkBytes;
      const float* lse_s = reinterpret_cast<const float*>(aux) + col0;
      const float* del_s = reinterpret_cast<const float*>(aux + 256) + col0;
      const uint32_t* msk_s = reinterpret_cast<const uint32_t*>(aux + 512) + q * 64 + col0;
      const int i0 = ib * kBlk + col0;
      float pd[kCW], ds[kCW];                                 // S^T -> Pd^T in place; dPd^T -> dS^T in place
#pragma unroll
      for (int c = 0; c < kCW; c += 16) {
        tmem_ld16_async(lane_addr + (uint32_t)buf * kBlk + (uint32_t)(col0 + c), pd + c);
        tmem_ld16_async(lane_addr + 128u + (uint32_t)buf * kBlk + (uint32_t)(col0 + c), ds + c);
      }
      tmem_ld_wait();
      const bool full = i0 + kCW <= P.T;
#pragma unroll
      for (int i4 = 0; i4 < kCW; i4 += 4) {
        const float4 L4 = *reinterpret_cast<const float4*>(lse_s + i4);
        const float4 D4 = *reinterpret_cast<const float4*>(del_s + i4);
        uint4 M4 = make_uint4(~0u, ~0u, ~0u, ~0u);
        if (drop) M4 = *reinterpret_cast<const uint4*>(msk_s + i4);
        const float Lv[4] = {L4.x, L4.y, L4.z, L4.w}, Dv[4] = {D4.x, D4.y, D4.z, D4.w};
        const uint32_t Mv[4] = {M4.x, M4.y, M4.z, M4.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i4 + u;
          const float p = ex2_approx(fmaf(pd[i], kScaleLog2e, -Lv[u]));
          const float kp = ((Mv[u] >> lane) & 1u) ? inv_keep : 0.f;
          float a = p * kp, g = p * fmaf(kp, ds[i], -Dv[u]);
          if (!full && i0 + i >= P.T) { a = 0.f; g = 0.f; }
          pd[i] = a; ds[i] = g;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(sd_empty + buf);
      mbar_wait(pd_empty, ((uint32_t)ib & 1u) ^ 1u);
#pragma unroll
      for (int c = 0; c < kCW; c += 16) {
        store16_sw128(pd_ptr, (col0 + c) >> 5, kTileBytes, r, ((col0 + c) & 31) >> 4, pd + c);
        store16_sw128(dst_ptr, (col0 + c) >> 5, kTileBytes, r, ((col0 + c) & 31) >> 4, ds + c);
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(pd_full);
    }
    mbar_wait(acc_full, 0);
    tc_fence_after();
    if (part < 2) {                                         // part 0 writes dK, part 1 writes dV
      float o[16];
      tmem_ld16(lane_addr + (part == 0 ? kDkCol : kDvCol) + (uint32_t)(h & 1) * 16u, o);
      if (live) {
        const float sc = part == 0 ? 0.25f : 1.0f;
        float4* dst = reinterpret_cast<float4*>(P.dqkv + (size_t)(row0 + kj) * (3 * kD) + (part == 0 ? kD : 2 * kD) + h * kDH);
#pragma unroll
        for (int g = 0; g < 4; ++g) dst[g] = make_float4(o[4 * g] * sc, o[4 * g + 1] * sc, o[4 * g + 2] * sc, o[4 * g + 3] * sc);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

extern "C" {

int afl_har_attn_mask_words(int T) { return (T + 127) / 128 * 4; }

int afl_har_attn_bwd_tc(const float* qkv, const float* out, const float* dout, const float* lse, const uint32_t* keep_mask,
                        int B, int T, float drop_p, float* delta, float* dqkv, void* stream) {
  AFL_CHECK_ARG(B >= 1 && T >= 1, "afl_har_attn_bwd_tc: B, T >= 1");
  AFL_CHECK_ARG(drop_p == 0.f || keep_mask != nullptr, "afl_har_attn_bwd_tc: dropout needs the forward's keep mask");
  AFL_CHECK_ARG(((uintptr_t)qkv | (uintptr_t)dout | (uintptr_t)lse | (uintptr_t)delta | (uintptr_t)keep_mask) % 16 == 0,
                "afl_har_attn_bwd_tc: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = (int64_t)B * T;
  {
    int g = (int)((rows * kNH + 255) / 256);
    if (g > afl_rt::num_sms() * 8) g = afl_rt::num_sms() * 8;
    attn_delta_kernel<<<g, 256, 0, st>>>(out, dout, B, T, delta);
  }
  CUtensorMap q128, o128, k64, k64t, q64t, o64, o64t;
  int rc = make_tmap_f32(&q128, qkv, 3 * kD, rows, 3 * kD, kTile, false);
  if (!rc) rc = make_tmap_f32(&o128, dout, kD, rows, kD, kTile, false);
  if (!rc) rc = make_tmap_f32(&k64, qkv, 3 * kD, rows, 3 * kD, kBlk, false);
  if (!rc) rc = make_tmap_f32(&k64t, qkv, 3 * kD, rows, 3 * kD, kBlk, true);
  if (!rc) rc = make_tmap_f32(&q64t, qkv, 3 * kD, rows, 3 * kD, kBlk, true);
  if (!rc) rc = make_tmap_f32(&o64, dout, kD, rows, kD, kBlk, false);
  if (!rc) rc = make_tmap_f32(&o64t, dout, kD, rows, kD, kBlk, true);
  if (rc) return rc;
  BwdParams P{};
  P.T = T; P.n_blk = (T + kBlk - 1) / kBlk; P.mask_words = afl_har_attn_mask_words(T); P.drop_p = drop_p;
  P.lse = lse; P.delta = delta; P.mask = drop_p > 0.f ? keep_mask : nullptr; P.dqkv = dqkv;
  dim3 grid((T + kTile - 1) / kTile, kNH, B);
  const size_t smem_dq = 1024 + 2 * kTileBytes + 3 * 3 * kBlkBytes + 4 * kTileBytes + 17 * 8 + 64;
  AFL_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_dq));
  attn_bwd_dq_tc_kernel<<<grid, kThreads, smem_dq, st>>>(q128, o128, k64, k64t, P);
  const size_t smem_kv = 1024 + 2 * kTileBytes + 3 * (4 * kBlkBytes + 2048) + 4 * kTileBytes + 15 * 8 + 64;
  AFL_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_kv));
  attn_bwd_dkv_tc_kernel<<<grid, kThreads, smem_kv, st>>>(q128, k64, k64t, o64, o64t, P);
  afl_rt::count_launch(3);
  return afl_rt::check_last("afl_har_attn_bwd_tc");
}

}  // extern "C"
