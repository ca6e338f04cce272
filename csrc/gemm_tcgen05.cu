// Linear-layer GEMMs of the HAR TransformerClassifier (T = 561 tokens x batch 128 => M = 71,808 rows per step) on the
// 5th-gen tensor cores: TMA (SWIZZLE_128B) -> tcgen05.mma kind::tf32 on the fp32 activations / weights as they lie in
// memory (no cast pass) -> fp32 accumulators in TMEM -> tcgen05.ld epilogue with the bias / ReLU / dropout / mask fused.
//
// What it replaces: the cuBLAS sgemm calls + separate elementwise kernels behind nn.TransformerEncoderLayer's
// in_proj / out_proj / linear1 / linear2 and their autograd (reference src/Model.py:435-458, client.py:114-131;
// SURVEY 2.13 rows T3/T4 "mlp_ln (HAR)").
//
//   forward   Y[M,N]  = act(X[M,K] . W[N,K]^T + b)      A = X tile   (K-major)   B = W  (K-major)
//   input grad dX[M,K] = dY[M,N] . W[N,K]   (* mask)      A = dY tile  (K-major)   B = W  (MN-major: same smem image)
//   weight grad dW[N,K] = dY[M,N]^T . X[M,K]              A = dY chunk (MN-major)  B = X chunk (MN-major), split over M
//
// The transposed operands are never materialised: a TMA box {32 features, R rows} of a row-major matrix, loaded with the
// 128B_ATOM_32B swizzle, IS the canonical MN-major layout for 32-bit operands (tc_common.cuh), so the MMA descriptors
// just say "MN-major".
//
// Row-tile kernel (forward / input grad): persistent, one CTA per SM, W resident in shared memory (<= 64 KB), the
// activation tiles stream through a ring of 16 KB stages, two TMEM accumulators so the epilogue of tile i overlaps
// the MMAs of tile i+1.  192 threads: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "common.h"
#include "runtime.h"
#include "tc_common.cuh"

namespace {

using namespace afl_tc;

constexpr int kThreads = 192;
constexpr int kStageBytes = 128 * 128;       // one activation k-block: 128 rows x 32 fp32

struct RowGemmParams {
  int64_t M;             // rows
  int n_kblocks;         // reduction length / 32
  int n_out;             // output columns = UMMA N
  int mode;              // 0: B K-major (forward), 1: B MN-major (input grad)
  int w_rows;            // rows of W as stored ([w_rows][w_kblocks*32])
  int w_kblocks;
  int stages;
  float* out;            // [M][n_out]
  const float* bias;     // mode 0
  int relu;              // mode 0
  float drop_p;          // mode 0: inverted dropout after the activation (0 = off)
  uint32_t seed_lo, seed_hi, rng_client, rng_step, rng_id;
  const float* mask_src; // mode 1: out *= (mask_src[row][col] > 0 ? mask_scale : 0)
  float mask_scale;
  int accumulate;        // out += result
  // mode 0, n_out == 64: fused residual + dropout + LayerNorm epilogue (out = LN(resid + dropout(acc + bias)))
  int ln;
  const float* resid;    // [M][64]
  const float* gamma;    // [64]
  const float* beta;     // [64]
  float* rstd;           // [M]; xhat goes out through tmap_x
};

__global__ void __launch_bounds__(kThreads, 1)
row_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                const __grid_constant__ CUtensorMap tmap_o, const __grid_constant__ CUtensorMap tmap_x, RowGemmParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t w_box_bytes = (uint32_t)P.w_rows * 128u;
  const uint32_t w_bytes = w_box_bytes * (uint32_t)P.w_kblocks;
  const uint32_t w_base = base;
  const uint32_t a_base = base + w_bytes;
  const uint32_t o_base = a_base + (uint32_t)P.stages * kStageBytes;      // 2 (4 with LN) output staging tiles [128 rows][32 cols]
  uint8_t* o_ptr = aligned + (o_base - base);
  uint64_t* full = reinterpret_cast<uint64_t*>(o_ptr + (P.ln ? 4 : 2) * kStageBytes);
  uint64_t* empty = full + P.stages;
  uint64_t* w_full = empty + P.stages;
  uint64_t* acc_full = w_full + 1;             // [2]
  uint64_t* acc_empty = acc_full + 2;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* bias_s = reinterpret_cast<float*>(tmem_slot + 2);               // [n_out]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_tiles = (P.M + 127) / 128;
  const uint32_t buf_cols = tmem_cols_pow2(P.n_out);
  const uint32_t tmem_cols = buf_cols * 2;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_w)) : "memory");
    for (int s = 0; s < P.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(w_full, 1);
    for (int b = 0; b < 2; ++b) { mbar_init(acc_full + b, 1); mbar_init(acc_empty + b, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: W once, then the activation k-blocks of every tile of this CTA =====
      mbar_expect_tx(w_full, w_bytes);
      for (int b = 0; b < P.w_kblocks; ++b) tma_load_2d(w_base + (uint32_t)b * w_box_bytes, &tmap_w, b * 32, 0, w_full);
      int it = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < P.n_kblocks; ++kb, ++it) {
          const int st = it % P.stages;
          const uint32_t ph = (uint32_t)(it / P.stages) & 1u;
          mbar_wait(empty + st, ph ^ 1u);
          mbar_expect_tx(full + st, kStageBytes);
          tma_load_2d(a_base + (uint32_t)st * kStageBytes, &tmap_a, kb * 32, (int)(tile * 128), full + st);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      const uint32_t idesc = idesc_tf32(128, P.n_out, 0, P.mode);
      const uint64_t ad0 = desc_k_sw128(a_base), bk0 = desc_k_sw128(w_base), bm0 = desc_mn_sw128(w_base, w_box_bytes);
      mbar_wait(w_full, 0);
      int it = 0, t = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++t) {
        const int buf = t & 1;
        mbar_wait(acc_empty + buf, (((uint32_t)t >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * buf_cols;
        for (int kb = 0; kb < P.n_kblocks; ++kb, ++it) {
          const int st = it % P.stages;
          const uint32_t ph = (uint32_t)(it / P.stages) & 1u;
          mbar_wait(full + st, ph);
          tc_fence_after();
          const uint64_t ad = desc_add(ad0, (uint32_t)st * kStageBytes);
          const uint64_t bd = P.mode == 0 ? desc_add(bk0, (uint32_t)kb * w_box_bytes) : desc_add(bm0, (uint32_t)kb * 4096u);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)         // 8 tf32 (32 bytes) of the reduction per MMA
            umma_tf32(d_tmem, desc_add(ad, (uint32_t)ks * 32u),
                      P.mode == 0 ? desc_add(bd, (uint32_t)ks * 32u) : desc_add(bd, (uint32_t)ks * 1024u), idesc,
                      (kb > 0 || ks > 0) ? 1u : 0u);
          umma_commit(empty + st);
        }
        umma_commit(acc_full + buf);
      }
    }
  } else {
    // ===== epilogue warps 2..5: TMEM -> registers -> fused elementwise -> swizzled staging tile -> TMA store =====
    // The output leaves through 16 KB staging tiles ([128 rows][32 columns], SWIZZLE_128B) and cp.async.bulk.tensor
    // stores (reduce-add when accumulating), so global writes are full 128-byte lines and rows >= M are clipped by TMA.
    const int q = warp & 3;
    const int et = threadIdx.x - 64;                       // 0..127
    const int r = q * 32 + lane;                           // row of the tile = TMEM lane
    if (P.mode == 0 && P.bias)
      for (int i = et; i < P.n_out; i += 128) bias_s[i] = P.bias[i];
    if (P.ln && et < 64) { bias_s[64 + et] = P.gamma[et]; bias_s[128 + et] = P.beta[et]; }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    int t = 0, sc = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++t) {
      const int buf = t & 1;
      mbar_wait(acc_full + buf, ((uint32_t)t >> 1) & 1u);
      tc_fence_after();
      const int64_t row = tile * 128 + r;
      const bool live = row < P.M;
      if (P.ln) {
        // ---- fused epilogue: y = LayerNorm(resid + dropout(acc + bias)) * gamma + beta; the thread owns the whole row, so
        // the statistics need no shuffles.  y and xhat leave through four staging tiles and TMA stores, rstd directly.
        float v[64];
        tmem_ld32_async(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * buf_cols, v);
        tmem_ld32_async(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * buf_cols + 32u, v + 32);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty + buf);           // accumulator is in registers: the next tile's MMAs may start
        const float inv = P.drop_p > 0.f ? 1.0f / (1.0f - P.drop_p) : 1.f;
        float sum = 0.f;
#pragma unroll
        for (int g = 0; g < 16; ++g) {
          float a[4] = {v[4 * g] + bias_s[4 * g], v[4 * g + 1] + bias_s[4 * g + 1], v[4 * g + 2] + bias_s[4 * g + 2],
                        v[4 * g + 3] + bias_s[4 * g + 3]};
          if (P.drop_p > 0.f) {
            const AflU4 rr = afl_philox((uint32_t)g, (P.rng_id & 0xFFFFu) | (P.rng_client << 16), P.rng_step, (uint32_t)row,
                                        P.seed_lo, P.seed_hi);
            const uint32_t w[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) a[j] = ((float)(w[j] >> 8) * (1.0f / 16777216.0f) >= P.drop_p) ? a[j] * inv : 0.f;
          }
          float4 x4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (live) x4 = __ldg(reinterpret_cast<const float4*>(P.resid + row * 64) + g);
          v[4 * g] = a[0] + x4.x; v[4 * g + 1] = a[1] + x4.y; v[4 * g + 2] = a[2] + x4.z; v[4 * g + 3] = a[3] + x4.w;
          sum += (v[4 * g] + v[4 * g + 1]) + (v[4 * g + 2] + v[4 * g + 3]);
        }
        const float mean = sum * (1.0f / 64.0f);
        float sq = 0.f;
#pragma unroll
        for (int j = 0; j < 64; ++j) { v[j] -= mean; sq = fmaf(v[j], v[j], sq); }
        const float rs = rsqrtf(sq * (1.0f / 64.0f) + 1e-5f);
        if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // previous tile's stores have read the tiles
        asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint8_t* yrow = o_ptr + (size_t)half * kStageBytes + (size_t)r * 128;
          uint8_t* xrow = o_ptr + (size_t)(2 + half) * kStageBytes + (size_t)r * 128;
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const int c = 32 * half + 4 * g;
            const float4 xh = make_float4(v[c] * rs, v[c + 1] * rs, v[c + 2] * rs, v[c + 3] * rs);
            const float4 y = make_float4(fmaf(xh.x, bias_s[64 + c], bias_s[128 + c]), fmaf(xh.y, bias_s[65 + c], bias_s[129 + c]),
                                         fmaf(xh.z, bias_s[66 + c], bias_s[130 + c]), fmaf(xh.w, bias_s[67 + c], bias_s[131 + c]));
            *reinterpret_cast<float4*>(xrow + ((g ^ (r & 7)) << 4)) = xh;
            *reinterpret_cast<float4*>(yrow + ((g ^ (r & 7)) << 4)) = y;
          }
        }
        if (live) P.rstd[row] = rs;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         ::"l"(reinterpret_cast<uint64_t>(k < 2 ? &tmap_o : &tmap_x)), "r"((k & 1) * 32), "r"((int)(tile * 128)),
                           "r"(o_base + (uint32_t)k * kStageBytes) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        continue;
      }
      for (int c0 = 0; c0 < P.n_out; c0 += 32, ++sc) {
        uint8_t* stage = o_ptr + (size_t)(sc & 1) * kStageBytes;
        if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store that used this tile is done reading
        asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
        for (int hc = 0; hc < 2; ++hc) {
          const int cc = c0 + 16 * hc;
          float v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * buf_cols + (uint32_t)cc, v);
          if (P.mode == 0) {
            if (P.bias) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += bias_s[cc + j];
            }
            if (P.relu) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
            }
            if (P.drop_p > 0.f) {
              const float inv = 1.0f / (1.0f - P.drop_p);
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const AflU4 rr = afl_philox((uint32_t)(cc + 4 * g) >> 2, (P.rng_id & 0xFFFFu) | (P.rng_client << 16),
                                            P.rng_step, (uint32_t)row, P.seed_lo, P.seed_hi);
                const uint32_t w[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  v[4 * g + j] = ((float)(w[j] >> 8) * (1.0f / 16777216.0f) >= P.drop_p) ? v[4 * g + j] * inv : 0.f;
              }
            }
          } else if (P.mask_src && live) {
            const float4* m4 = reinterpret_cast<const float4*>(P.mask_src + row * P.n_out + cc);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const float4 m = __ldg(m4 + g);
              v[4 * g + 0] = m.x > 0.f ? v[4 * g + 0] * P.mask_scale : 0.f;
              v[4 * g + 1] = m.y > 0.f ? v[4 * g + 1] * P.mask_scale : 0.f;
              v[4 * g + 2] = m.z > 0.f ? v[4 * g + 2] * P.mask_scale : 0.f;
              v[4 * g + 3] = m.w > 0.f ? v[4 * g + 3] * P.mask_scale : 0.f;
            }
          }
          uint8_t* rowp = stage + (size_t)r * 128;
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *reinterpret_cast<float4*>(rowp + (((4 * hc + g) ^ (r & 7)) << 4)) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) {
          const uint32_t src = o_base + (uint32_t)(sc & 1) * kStageBytes;
          if (P.accumulate)
            asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
                         ::"l"(reinterpret_cast<uint64_t>(&tmap_o)), "r"(c0), "r"((int)(tile * 128)), "r"(src) : "memory");
          else
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         ::"l"(reinterpret_cast<uint64_t>(&tmap_o)), "r"(c0), "r"((int)(tile * 128)), "r"(src) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + buf);
    }
    if (et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------------------------------------------------------------------------- weight gradient
constexpr int kChunkRows = 32;                 // tokens per pipeline stage (4 MMAs of 8 tokens)
constexpr int kBoxBytes = kChunkRows * 128;    // one {32 features, 32 tokens} box

struct DwParams {
  int64_t M;
  int N, K;              // dW is [N][K]
  int stages;
  int64_t n_chunks;
  int want_db;           // also sum the columns of dY (bias gradient) from the staged tiles
  float* partials;       // [gridDim.x][N*K + N]: dW partial, then db partial
};

__global__ void __launch_bounds__(kThreads, 1)
dw_gemm_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x, DwParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* aligned = smem_raw + (base - smem_u32(smem_raw));
  const int a_boxes = P.N / 32, b_boxes = P.K / 32;
  const uint32_t stage_bytes = (uint32_t)(a_boxes + b_boxes) * kBoxBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(aligned + (size_t)P.stages * stage_bytes + 4 * kBoxBytes /* slack, see below */);
  uint64_t* empty = full + P.stages;
  uint64_t* acc_full = empty + P.stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (P.N + 127) / 128;
  const uint32_t tmem_cols = tmem_cols_pow2(m_tiles * P.K);
  const int64_t per = (P.n_chunks + gridDim.x - 1) / gridDim.x;
  const int64_t c0 = min(P.n_chunks, (int64_t)blockIdx.x * per);
  const int64_t c1 = min(P.n_chunks, c0 + per);
  const int nch = (int)(c1 - c0);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_dy)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");
    // a stage is free again when its MMAs have retired and (bias gradient wanted) the four epilogue warps have read it
    for (int s = 0; s < P.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, P.want_db ? 5 : 1); }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < nch; ++i) {
        const int st = i % P.stages;
        const uint32_t ph = (uint32_t)(i / P.stages) & 1u;
        mbar_wait(empty + st, ph ^ 1u);
        mbar_expect_tx(full + st, stage_bytes);
        const uint32_t s_addr = base + (uint32_t)st * stage_bytes;
        const int row0 = (int)((c0 + i) * kChunkRows);
        for (int a = 0; a < a_boxes; ++a) tma_load_2d(s_addr + (uint32_t)a * kBoxBytes, &tmap_dy, a * 32, row0, full + st);
        for (int b = 0; b < b_boxes; ++b)
          tma_load_2d(s_addr + (uint32_t)(a_boxes + b) * kBoxBytes, &tmap_x, b * 32, row0, full + st);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nch > 0) {
      // A = dY^T (M-dim = dY features), B = X^T (N-dim = X features), both MN-major; 8 tokens per MMA.
      // An M = 128 tile that covers fewer than 128 real features (N = 64, or the second tile of N = 192) reads the
      // next boxes of the stage (X data, or the slack after the last stage) as extra A rows: finite or not, those
      // accumulator rows are never read back.
      const uint32_t idesc = idesc_tf32(128, P.K, 1, 1);
      const uint64_t ad0 = desc_mn_sw128(base, kBoxBytes);
      for (int i = 0; i < nch; ++i) {
        const int st = i % P.stages;
        const uint32_t ph = (uint32_t)(i / P.stages) & 1u;
        mbar_wait(full + st, ph);
        tc_fence_after();
        const uint64_t ad = desc_add(ad0, (uint32_t)st * stage_bytes);
        const uint64_t bd = desc_add(ad, (uint32_t)a_boxes * kBoxBytes);
#pragma unroll
        for (int ks = 0; ks < kChunkRows / 8; ++ks)
          for (int mt = 0; mt < m_tiles; ++mt)
            umma_tf32(tmem_base + (uint32_t)(mt * P.K), desc_add(ad, (uint32_t)(mt * 4) * kBoxBytes + (uint32_t)ks * 1024u),
                      desc_add(bd, (uint32_t)ks * 1024u), idesc, (i > 0 || ks > 0) ? 1u : 0u);
        umma_commit(empty + st);
      }
      umma_commit(acc_full);
    }
  } else {
    const int q = warp & 3;
    float* out = P.partials + (size_t)blockIdx.x * ((size_t)P.N * P.K + P.N);
    if (P.want_db) {
      // bias gradient: column sums of the dY boxes while they sit in shared memory (MN-major 128B_ATOM_32B image:
      // token row r at r*128 bytes, 32-byte chunk index XOR-ed with r & 3); thread e owns columns e and e + 128.
      const int e = (warp - 2) * 32 + lane;
      float acc0 = 0.f, acc1 = 0.f;
      const uint32_t off0 = (uint32_t)(e >> 5) * kBoxBytes + (uint32_t)(e & 7) * 4u;
      const uint32_t c32 = (uint32_t)(e & 31) >> 3;
      for (int i = 0; i < nch; ++i) {
        const int st = i % P.stages;
        mbar_wait(full + st, (uint32_t)(i / P.stages) & 1u);
        const uint8_t* sp = aligned + (size_t)st * stage_bytes;
#pragma unroll 8
        for (int r = 0; r < kChunkRows; ++r) {
          const uint32_t o = off0 + (uint32_t)r * 128u + ((c32 ^ (uint32_t)(r & 3)) << 5);
          if (e < P.N) acc0 += *reinterpret_cast<const float*>(sp + o);
          if (e + 128 < P.N) acc1 += *reinterpret_cast<const float*>(sp + o + 4 * kBoxBytes);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + st);
      }
      float* ob = out + (size_t)P.N * P.K;
      if (e < P.N) ob[e] = acc0;
      if (e + 128 < P.N) ob[e + 128] = acc1;
    }
    if (nch > 0) {
      mbar_wait(acc_full, 0);
      tc_fence_after();
    }
    for (int mt = 0; mt < m_tiles; ++mt) {
      const int n = mt * 128 + q * 32 + lane;
      for (int k0 = 0; k0 < P.K; k0 += 16) {
        float v[16];
        if (nch > 0) {
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * P.K + k0), v);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = 0.f;
        }
        if (n < P.N) {
          float4* o4 = reinterpret_cast<float4*>(out + (size_t)n * P.K + k0);
#pragma unroll
          for (int g = 0; g < 4; ++g) o4[g] = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// dW[i] (+)= sum_p partials[p][i] for i < n_w, db[i - n_w] (+)= ... for the n_b entries after them.  A block owns 64
// consecutive outputs; its four 64-thread groups each sum every fourth partial (8 independent loads in flight per thread)
// and the groups are combined in a fixed order through shared memory: deterministic, and ~4x the memory parallelism of
// one thread per output walking all partials.
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float* __restrict__ partials, int n_parts, int64_t n_w,
                                                              int n_b, float* __restrict__ dw, float* __restrict__ db,
                                                              int accumulate) {
  __shared__ float part[4][64];
  const int o = threadIdx.x & 63, g = threadIdx.x >> 6;
  const int64_t i = (int64_t)blockIdx.x * 64 + o;
  const int64_t stride = n_w + n_b;
  float acc = 0.f;
  if (i < stride) {
    const float* src = partials + i;
    int p = g;
#pragma unroll 1
    for (; p + 28 < n_parts; p += 32) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = src[(size_t)(p + 4 * u) * stride];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u];
    }
    for (; p < n_parts; p += 4) acc += src[(size_t)p * stride];
  }
  part[g][o] = acc;
  __syncthreads();
  if (g == 0 && i < stride && (i < n_w || db != nullptr)) {
    const float sum = (part[0][o] + part[1][o]) + (part[2][o] + part[3][o]);
    float* dst = i < n_w ? dw + i : db + (i - n_w);
    *dst = accumulate ? *dst + sum : sum;
  }
}

int check_dims(int N, int K, const char* what) {
  if (N < 32 || N > 256 || K < 32 || K > 256 || (N % 32) || (K % 32)) {
    afl_rt::set_error("%s: N and K must be multiples of 32 in [32, 256] (got N=%d K=%d)", what, N, K);
    return -1;
  }
  return 0;
}

int launch_row_gemm(const float* A, int64_t M, int Kr, const float* W, int w_rows, int w_cols, int mode, RowGemmParams P,
                    cudaStream_t st, float* xhat = nullptr) {
  CUtensorMap ta, tw, to;
  int rc = make_tmap_f32(&ta, A, Kr, M, Kr, 128);
  if (rc) return rc;
  rc = make_tmap_f32(&tw, W, w_cols, w_rows, w_cols, w_rows, mode == 1);
  if (rc) return rc;
  const int n_out = mode == 0 ? w_rows : w_cols;
  rc = make_tmap_f32(&to, P.out, n_out, M, n_out, 128);
  if (rc) return rc;
  CUtensorMap tx = to;
  if (P.ln) {
    rc = make_tmap_f32(&tx, xhat, 64, M, 64, 128);
    if (rc) return rc;
  }
  P.M = M;
  P.n_kblocks = Kr / 32;
  P.mode = mode;
  P.w_rows = w_rows;
  P.w_kblocks = w_cols / 32;
  P.n_out = mode == 0 ? w_rows : w_cols;
  const size_t w_bytes = (size_t)w_rows * w_cols * 4;
  const size_t fixed = 1024 + (P.ln ? 4 : 2) * (size_t)kStageBytes + 2048 + w_bytes;   // alignment, output staging, barriers + bias
  const size_t budget = (size_t)afl_rt::max_smem_optin() - fixed;
  int stages = (int)(budget / kStageBytes);
  if (stages > 8) stages = 8;
  if (stages < 2) { afl_rt::set_error("row gemm: weight tile does not leave room for 2 pipeline stages"); return -2; }
  P.stages = stages;
  const size_t smem = fixed + (size_t)stages * kStageBytes;
  AFL_CUDA(cudaFuncSetAttribute(row_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t n_tiles = (M + 127) / 128;
  const int grid = (int)(n_tiles < afl_rt::num_sms() ? n_tiles : afl_rt::num_sms());
  row_gemm_kernel<<<grid, kThreads, smem, st>>>(ta, tw, to, tx, P);
  afl_rt::count_launch();
  return afl_rt::check_last("row_gemm_kernel");
}

}  // namespace

extern "C" {

int afl_tc_linear_fwd(const float* X, int64_t M, int K, const float* W, int N, const float* bias, int relu, float drop_p,
                      uint64_t seed, uint32_t rng_client, uint32_t rng_step, uint32_t rng_id, float* Y, void* stream) {
  if (check_dims(N, K, "afl_tc_linear_fwd")) return -1;
  AFL_CHECK_ARG(M >= 1 && drop_p >= 0.f && drop_p < 1.f, "afl_tc_linear_fwd: M >= 1, 0 <= p < 1");
  AFL_CHECK_ARG(((uintptr_t)X | (uintptr_t)W | (uintptr_t)Y) % 16 == 0, "afl_tc_linear_fwd: pointers must be 16-byte aligned");
  RowGemmParams P{};
  P.out = Y; P.bias = bias; P.relu = relu; P.drop_p = drop_p;
  P.seed_lo = (uint32_t)seed; P.seed_hi = (uint32_t)(seed >> 32);
  P.rng_client = rng_client; P.rng_step = rng_step; P.rng_id = rng_id;
  return launch_row_gemm(X, M, K, W, N, K, 0, P, (cudaStream_t)stream);
}

int afl_tc_linear_ln_fwd(const float* X, int64_t M, int K, const float* W, const float* bias, const float* resid, float drop_p,
                         uint64_t seed, uint32_t rng_client, uint32_t rng_step, uint32_t rng_id, const float* gamma,
                         const float* beta, float* Y, float* xhat, float* rstd, void* stream) {
  if (check_dims(64, K, "afl_tc_linear_ln_fwd")) return -1;
  AFL_CHECK_ARG(M >= 1 && drop_p >= 0.f && drop_p < 1.f && bias && resid && gamma && beta,
                "afl_tc_linear_ln_fwd: M >= 1, 0 <= p < 1, bias / resid / gamma / beta required");
  AFL_CHECK_ARG(((uintptr_t)X | (uintptr_t)W | (uintptr_t)Y | (uintptr_t)xhat | (uintptr_t)resid) % 16 == 0,
                "afl_tc_linear_ln_fwd: pointers must be 16-byte aligned");
  RowGemmParams P{};
  P.out = Y; P.bias = bias; P.drop_p = drop_p;
  P.seed_lo = (uint32_t)seed; P.seed_hi = (uint32_t)(seed >> 32);
  P.rng_client = rng_client; P.rng_step = rng_step; P.rng_id = rng_id;
  P.ln = 1; P.resid = resid; P.gamma = gamma; P.beta = beta; P.rstd = rstd;
  return launch_row_gemm(X, M, K, W, 64, K, 0, P, (cudaStream_t)stream, xhat);
}

int afl_tc_linear_dx(const float* dY, int64_t M, int N, const float* W, int K, const float* mask_src, float mask_scale,
                     int accumulate, float* dX, void* stream) {
  if (check_dims(N, K, "afl_tc_linear_dx")) return -1;
  AFL_CHECK_ARG(M >= 1, "afl_tc_linear_dx: M >= 1");
  AFL_CHECK_ARG(((uintptr_t)dY | (uintptr_t)W | (uintptr_t)dX | (uintptr_t)mask_src) % 16 == 0,
                "afl_tc_linear_dx: pointers must be 16-byte aligned");
  RowGemmParams P{};
  P.out = dX; P.mask_src = mask_src; P.mask_scale = mask_scale; P.accumulate = accumulate;
  return launch_row_gemm(dY, M, N, W, N, K, 1, P, (cudaStream_t)stream);
}

int afl_tc_linear_dw_workspace(int N, int K, int64_t* partial_floats) {
  if (check_dims(N, K, "afl_tc_linear_dw_workspace")) return -1;
  *partial_floats = (int64_t)afl_rt::num_sms() * ((int64_t)N * K + N);
  return 0;
}

int afl_tc_linear_dw(const float* dY, const float* X, int64_t M, int N, int K, float* partials, float* dW, float* db,
                     int accumulate, void* stream) {
  if (check_dims(N, K, "afl_tc_linear_dw")) return -1;
  AFL_CHECK_ARG(M >= 1, "afl_tc_linear_dw: M >= 1");
  AFL_CHECK_ARG(((uintptr_t)dY | (uintptr_t)X | (uintptr_t)dW | (uintptr_t)partials) % 16 == 0,
                "afl_tc_linear_dw: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap ty, tx;
  int rc = make_tmap_f32(&ty, dY, N, M, N, kChunkRows, true);
  if (rc) return rc;
  rc = make_tmap_f32(&tx, X, K, M, K, kChunkRows, true);
  if (rc) return rc;
  DwParams P{};
  P.M = M; P.N = N; P.K = K; P.partials = partials; P.want_db = db != nullptr ? 1 : 0;
  P.n_chunks = (M + kChunkRows - 1) / kChunkRows;
  const size_t stage_bytes = (size_t)(N / 32 + K / 32) * kBoxBytes;
  const size_t budget = (size_t)afl_rt::max_smem_optin() - 1024 - 256 - 4 * kBoxBytes;
  int stages = (int)(budget / stage_bytes);
  if (stages > 8) stages = 8;
  P.stages = stages;
  const size_t smem = 1024 + (size_t)stages * stage_bytes + 4 * kBoxBytes + (2 * stages + 1) * 8 + 16;
  AFL_CUDA(cudaFuncSetAttribute(dw_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = (int)(P.n_chunks < afl_rt::num_sms() ? P.n_chunks : afl_rt::num_sms());
  dw_gemm_kernel<<<grid, kThreads, smem, st>>>(ty, tx, P);
  afl_rt::count_launch();
  const int64_t n = (int64_t)N * K;
  reduce_partials_kernel<<<(int)((n + N + 63) / 64), 256, 0, st>>>(partials, grid, n, N, dW, db, accumulate);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_tc_linear_dw");
}

}  // extern "C"
