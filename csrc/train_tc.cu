// Tensor-core local training of the ICU TransformerModel: one FL round of one client in ONE CTA (sm_100a).
//
// What it replaces: the reference's per-client Python loop `train_ICU` (/root/reference/client.py:75-111) for
// `TransformerModel` (/root/reference/src/Model.py:184-191,225-246), ~525 optimiser steps per round.  The FFMA interpreter
// (train.cu) splits a batch over 16 blocks and meets at ~45 block barriers plus two client barriers per step; here the
// whole batch (M = 128 = one UMMA tile) stays in one CTA and every contraction is a tcgen05.mma (kind::tf32, fp32
// accumulate in TMEM):
//   * A operands never touch shared memory: the element-wise warps read an accumulator with tcgen05.ld, apply
//     bias / GELU / dropout / residual / LayerNorm, and write the next layer's A operand straight back to TMEM with
//     tcgen05.st (lane = sample, column = feature), so forward and dX chains are TMEM -> registers -> TMEM.
//   * B operands (weights) are streamed from the client's parameter row in L2 by TMA, K-major (SWIZZLE_128B) for the
//     forward and MN-major (SWIZZLE_128B_ATOM_32B) for dX = dY.W from the same bytes - no transposed copy exists.
//   * dW = dY^T.X runs as M = 64 MMAs with both operands MN-major: dY is written by the element-wise warps in the
//     32-byte-atom swizzle, X comes back from the activation stash (global, L2-resident) by TMA.  The 64 x N gradient
//     tile is drained from TMEM into the client's gradient row; biases / LayerNorm parameters are column sums done with
//     a 16-shuffle butterfly and shared-memory atomics.
//   * one producer warp issues every TMA and MMA from a static schedule; the vitals and labs branches are two
//     independent streams, so one branch's MMA latency hides behind the other's element-wise work.
//   * dropout masks are the same counter-based Philox functions of (seed, client, step, sample, site, element) as in the
//     interpreter, so both trainers draw identical masks.
// The optimiser step (fresh Adam per round / SGD-momentum, client.py:78,107) is afl_reduce_update on the gradient row.
#include <cuda.h>
#include <cuda_runtime.h>
#include <string.h>

#include "../include/attackfl_b200.h"
#include "runtime.h"
#include "tc_common.cuh"
#include "train_core.h"

namespace {

using namespace afl_tc;

constexpr int kEpi = 512;                 // element-wise threads: warp w -> TMEM lane quarter w & 3, 16-column chunk w >> 2
constexpr int kThreads = kEpi + 32;       // + one producer warp (TMA + MMA issue)
constexpr int kSlotB = 16384;             // one [128 x 32] fp32 block
constexpr int kSlotF = 4096;
constexpr int kNSlot = 48;                // stash slots per client

// ---- stash slots (per client); per-branch base = 19 * s
enum { SL_X = 0, SL_GP1 = 1, SL_H1 = 3, SL_VD = 5, SL_XH1 = 7, SL_N1 = 9, SL_GPF = 11, SL_GF = 12, SL_XH2 = 13, SL_XH3 = 15,
       SL_RES = 17, SL_BR = 19, SL_CAT = 38, SL_GPA = 42, SL_HD = 44, SL_H2 = 46 };

// ---- small parameters / their gradients mirrored in shared memory; per-branch base = 656 * s
enum { SP_BD = 0, SP_BV = 64, SP_BO = 128, SP_G1 = 192, SP_BE1 = 256, SP_BF0 = 320, SP_BF3 = 336, SP_G2 = 400, SP_BE2 = 464,
       SP_G3 = 528, SP_BE3 = 592, SP_BR = 656, SP_B1 = 1312, SP_B2 = 1376, SP_W3 = 1408, SP_B3 = 1440, SP_N = 1456 };

// ---- tensor maps
enum { TM_WV_K = 0, TM_WO_K = 2, TM_WF0_K = 4, TM_W1_K = 6, TM_W2_K = 7, TM_WV_M = 8, TM_WO_M = 10, TM_WF0_M = 12, TM_W1_M = 14,
       TM_W2_M = 15, TM_STASH = 16, TM_N = 17 };

// ---- mbarriers
enum { B_STEP = 0, B_A = 1 /*V,L,H*/, B_ACC = 4, B_DW = 7, B_WF_V = 10 /*Wv*/, B_WF_O = 12 /*Wo*/, B_WF_W1 = 14, B_WF_W2 = 15,
       B_WB = 16 /*[s][2]*/, B_X = 20 /*[s]*/, B_N = 22 };

// ---- TMEM columns
constexpr uint32_t kColA = 0;             // A operands: V 0..63, L 64..127 (cat = 0..127)
constexpr uint32_t kColAcc = 128;         // accumulators: V 128..191, L 192..255 (head: 128..)
constexpr uint32_t kColDw = 256;          // dW accumulators: V 256..383 (two 64-col slots), L 384..511

struct TcLay {            // offsets (floats) inside one parameter row, [2] = vitals / labs
  int wd[2], bd[2], wv[2], bv[2], wo[2], bo[2], g1[2], be1[2], wf0[2], bf0[2], wf3[2], bf3[2], g2[2], be2[2], g3[2], be3[2];
  int w1, b1, w2, b2, w3, b3;
  int kin[2], ff;
  int rng_head[2], rng_ln1[2], rng_f1[2], rng_ln2[2], rng_fc1;
};

struct TcParams {
  TcLay L;
  float* w; float* m; float* v; float* gw; float* stash; int64_t d_pad;
  const float* data; int rec;
  const int32_t* idx; int64_t idx_stride_c, idx_stride_e;
  const int32_t* n_data; const int32_t* client_ids;
  int batch, epochs, opt_kind;
  float lr, beta1, beta2, eps, momentum; int clip_proper; float clip;
  uint32_t seed_lo, seed_hi; int dropout; float p1, p3;
  int n_ranges; int range_lo[16], range_hi[16];
  float* loss_out; int32_t* steps_out;
};

struct TcMaps { CUtensorMap m[TM_N]; };

// ---- shared memory carve-up (bytes from a 1024-aligned base)
constexpr uint32_t kOffW = 0;                           // wbuf[s][2]: 4 x 16 KB
constexpr uint32_t kOffX = 4 * kSlotB;                  // xbuf[s]: 2 x 32 KB
constexpr uint32_t kOffDy = 8 * kSlotB;                 // dymn[s]: 2 x 32 KB
constexpr uint32_t kOffSm = 12 * kSlotB;                // small weights (no-swizzle K-major): wd V 2K, wd L 4K, wf3 V 2K, wf3 L 2K, w3 2K
constexpr uint32_t kSmWd0 = 0, kSmWd1 = 2048, kSmWf30 = 6144, kSmWf31 = 8192, kSmW3 = 10240, kSmBytes = 12288;
constexpr uint32_t kOffSp = kOffSm + kSmBytes;          // small parameters [SP_N]
constexpr uint32_t kOffSg = kOffSp + SP_N * 4;          // their gradients  [SP_N]
constexpr uint32_t kOffLnx = kOffSg + SP_N * 4;         // [2][128][4] float2: LayerNorm partial exchange (double buffered)
constexpr uint32_t kOffRow = kOffLnx + 2 * 128 * 4 * 8; // dz[128], y[128]
constexpr uint32_t kOffMisc = kOffRow + 2 * 128 * 4;    // loss[2], red[16], tmem slot
constexpr uint32_t kOffBar = kOffMisc + 128;
constexpr uint32_t kSmemBytes = kOffBar + B_N * 8;
static_assert(kSmemBytes <= 232448, "shared memory budget");

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void umma_tf32_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ uint64_t desc_nosw(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}

// mbarrier wait with a time-out: a protocol bug traps (the launch fails with an error) instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_to(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  const long long t0 = clock64();
  for (uint32_t it = 0;; ++it) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
        : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) return;
    if ((it & 255u) == 255u && clock64() - t0 > 4000000000LL) __trap();
  }
}

__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void bar_named(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ---- MMA job shapes (issued by one thread) --------------------------------------------------------------------------
// forward / dX: D[128 x N] = A(tmem, K columns) . B
__device__ __forceinline__ void job_ksw(uint32_t d, uint32_t a, uint32_t b_addr, int N, int K, uint32_t box_bytes) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 0);
  for (int ks = 0; ks < K / 8; ++ks)
    umma_tf32_ts(d, a + ks * 8, desc_add(desc_k_sw128(b_addr + (uint32_t)(ks >> 2) * box_bytes), (uint32_t)(ks & 3) * 32u), idesc, ks > 0);
}
__device__ __forceinline__ void job_knosw(uint32_t d, uint32_t a, uint32_t b_addr, int N, int K) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 0);
  const uint64_t bd = desc_nosw(b_addr, 128, (uint32_t)(K / 4) * 128u);
  for (int ks = 0; ks < K / 8; ++ks) umma_tf32_ts(d, a + ks * 8, desc_add(bd, (uint32_t)ks * 256u), idesc, ks > 0);
}
__device__ __forceinline__ void job_mn(uint32_t d, uint32_t a, uint32_t b_addr, int N, int K, uint32_t lbo) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 1);
  const uint64_t bd = desc_mn_sw128(b_addr, lbo);
  for (int ks = 0; ks < K / 8; ++ks) umma_tf32_ts(d, a + ks * 8, desc_add(bd, (uint32_t)ks * 1024u), idesc, ks > 0);
}
// dW: D[64 x N] = dY^T (smem MN-major, two 32-wide blocks 16 KB apart) . X (smem MN-major), K = 128 samples
__device__ __forceinline__ void job_dw(uint32_t d, uint32_t a_addr, uint32_t b_addr, int N, uint32_t b_lbo) {
  const uint32_t idesc = idesc_tf32(64, N, 1, 1);
  const uint64_t ad = desc_mn_sw128(a_addr, kSlotB), bd = desc_mn_sw128(b_addr, b_lbo);
  for (int ks = 0; ks < 16; ++ks) umma_tf32(d, desc_add(ad, (uint32_t)ks * 1024u), desc_add(bd, (uint32_t)ks * 1024u), idesc, ks > 0);
}

// Column sums over the 32 rows of a warp for 16 columns per lane: 16 shuffles, then one shared atomic per column pair.
__device__ __forceinline__ void colsum16(const float* v, float* dst, int lane) {
  float a[8], b[4], c[2];
  const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4, h1 = lane & 2;
#pragma unroll
  for (int j = 0; j < 8; ++j) { const float send = h4 ? v[j] : v[j + 8], keep = h4 ? v[j + 8] : v[j]; a[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16); }
#pragma unroll
  for (int j = 0; j < 4; ++j) { const float send = h3 ? a[j] : a[j + 4], keep = h3 ? a[j + 4] : a[j]; b[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8); }
#pragma unroll
  for (int j = 0; j < 2; ++j) { const float send = h2 ? b[j] : b[j + 2], keep = h2 ? b[j + 2] : b[j]; c[j] = keep + __shfl_xor_sync(0xffffffffu, send, 4); }
  float d = (h1 ? c[1] : c[0]) + __shfl_xor_sync(0xffffffffu, h1 ? c[0] : c[1], 2);
  d += __shfl_xor_sync(0xffffffffu, d, 1);
  const int col = (h4 ? 8 : 0) + (h3 ? 4 : 0) + (h2 ? 2 : 0) + (h1 ? 1 : 0);
  if (!(lane & 1)) atomicAdd(dst + col, d);
}

__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

struct Rng { uint32_t seed_lo, seed_hi, client, step; int on; };

// dropout scales of elements elem4 .. elem4+3 of site `rng` for sample `row` (same stream as afl_drop_scale4)
__device__ __forceinline__ void drop4(const Rng& r, int rng, uint32_t row, uint32_t elem4, float p, float* s) {
  if (!(r.on && p > 0.f)) { s[0] = s[1] = s[2] = s[3] = 1.f; return; }
  const AflU4 u = afl_philox(elem4 >> 2, ((uint32_t)rng & 0xFFFFu) | (r.client << 16), r.step, row, r.seed_lo, r.seed_hi);
  const float inv = 1.f / (1.f - p), sc = 1.0f / 16777216.0f;
  s[0] = ((float)(u.x >> 8) * sc >= p) ? inv : 0.f;
  s[1] = ((float)(u.y >> 8) * sc >= p) ? inv : 0.f;
  s[2] = ((float)(u.z >> 8) * sc >= p) ? inv : 0.f;
  s[3] = ((float)(u.w >> 8) * sc >= p) ? inv : 0.f;
}
__device__ __forceinline__ void drop16(const Rng& r, int rng, uint32_t row, uint32_t col0, float p, float* s) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) drop4(r, rng, row, col0 + j, p, s + j);
}
__device__ __forceinline__ float drop1(const Rng& r, int rng, uint32_t row, uint32_t elem, float p) {
  if (!(r.on && p > 0.f)) return 1.f;
  return afl_keep(r.seed_lo, r.seed_hi, r.client, r.step, row, (uint32_t)rng, elem, p) ? 1.f / (1.f - p) : 0.f;
}

__device__ __forceinline__ void ld16g(const float* p, float* v) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float4 t = __ldcg(reinterpret_cast<const float4*>(p) + j);
    v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
  }
}
__device__ __forceinline__ void st16g(float* p, const float* v) {
#pragma unroll
  for (int j = 0; j < 4; ++j) __stcg(reinterpret_cast<float4*>(p) + j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
}

// 16 values of (row, col0 .. col0+15) into an MN-major SWIZZLE_128B_ATOM_32B tile whose 32-column blocks are 16 KB apart
__device__ __forceinline__ void st_dymn(uint8_t* buf, int row, int col0, const float* v) {
  uint8_t* base = buf + (col0 >> 5) * kSlotB + row * 128;
  const int c8 = (col0 & 31) >> 3;                   // first 32-byte chunk (0 or 2)
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    float4* dst = reinterpret_cast<float4*>(base + (((c8 + h) ^ (row & 3)) << 5));
    dst[0] = make_float4(v[8 * h], v[8 * h + 1], v[8 * h + 2], v[8 * h + 3]);
    dst[1] = make_float4(v[8 * h + 4], v[8 * h + 5], v[8 * h + 6], v[8 * h + 7]);
  }
}

__global__ void __launch_bounds__(kThreads, 1)
tc_train_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  float* sp = reinterpret_cast<float*>(smem + kOffSp);
  float* sg = reinterpret_cast<float*>(smem + kOffSg);
  float2* lnx = reinterpret_cast<float2*>(smem + kOffLnx);
  float* dzrow = reinterpret_cast<float*>(smem + kOffRow);
  float* ylab = dzrow + 128;
  float* loss_acc = reinterpret_cast<float*>(smem + kOffMisc);      // [2]
  float* red = loss_acc + 2;                                        // [16]
  uint32_t* tslot = reinterpret_cast<uint32_t*>(smem + kOffMisc + 96);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = blockIdx.x;
  const TcLay& L = P.L;

  // ---- one-time set-up
  for (uint32_t i = tid; i < (kOffBar) / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  float* w = P.w + (int64_t)c * P.d_pad;
  float* mrow = P.m + (int64_t)c * P.d_pad;
  float* vrow = P.v ? P.v + (int64_t)c * P.d_pad : nullptr;
  float* gw = P.gw + (int64_t)c * P.d_pad;
  float* stash = P.stash + (int64_t)c * kNSlot * kSlotF;
  for (int64_t i = tid; i < P.d_pad; i += kThreads) gw[i] = 0.f;
  __syncthreads();
  if (tid == 0) {
    for (int b = 0; b < B_N; ++b) {
      const bool many = b == B_STEP || (b >= B_A && b < B_A + 3);
      mbar_init(&bars[b], many ? 16u : 1u);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;

  const int n = P.n_data[c];
  const int steps_per_epoch = (n + P.batch - 1) / P.batch;

  if (warp == kEpi / 32) {
    // =================================================================================== producer warp
    if (lane == 0) {
      uint32_t ph = 0;                          // parity toggles of the barriers this thread waits on every phase
      uint32_t n_acc[3] = {0, 0, 0}, n_dw[3] = {0, 0, 0};
      auto wait = [&](int b) { mbar_wait_to(&bars[b], (ph >> b) & 1u); ph ^= 1u << b; };
      auto wait_a = [&](int s) { wait(B_A + s); tc_fence_after(); };
      auto commit_acc = [&](int s) { umma_commit(&bars[B_ACC + s]); ++n_acc[s]; };
      auto commit_dw = [&](int s) { umma_commit(&bars[B_DW + s]); ++n_dw[s]; };
      auto wait_acc_last = [&](int s) { mbar_wait_to(&bars[B_ACC + s], (n_acc[s] - 1) & 1u); };
      auto wait_dw_last = [&](int s) { mbar_wait_to(&bars[B_DW + s], (n_dw[s] - 1) & 1u); };
      auto tma = [&](int map, uint32_t dst, int c0, int c1, int bar) { tma_load_3d(dst, &maps.m[map], c0, c1, c, &bars[bar]); };
      auto stash_ld = [&](int slot, int nslots, uint32_t dst, int bar) {
        mbar_expect_tx(&bars[bar], (uint32_t)nslots * kSlotB);
        for (int j = 0; j < nslots; ++j) tma_load_3d(dst + j * kSlotB, &maps.m[TM_STASH], 0, 0, c * kNSlot + slot + j, &bars[bar]);
      };
      const uint32_t wb[2][2] = {{sbase + kOffW, sbase + kOffW + kSlotB}, {sbase + kOffW + 2 * kSlotB, sbase + kOffW + 3 * kSlotB}};
      const uint32_t xb[2] = {sbase + kOffX, sbase + kOffX + 2 * kSlotB};
      const uint32_t dy[2] = {sbase + kOffDy, sbase + kOffDy + 2 * kSlotB};
      const uint32_t w2f = xb[1], wf0f[2] = {xb[1] + kSlotB, xb[1] + kSlotB + 4096};
      const uint32_t smw = sbase + kOffSm;
      const uint32_t accc[3] = {tmem + kColAcc, tmem + kColAcc + 64, tmem + kColAcc};
      const uint32_t acol[3] = {tmem + kColA, tmem + kColA + 64, tmem + kColA};
      int dwk[2] = {0, 0};                     // alternating dW accumulator slot per branch
      auto dwcol = [&](int s) { return tmem + kColDw + (uint32_t)s * 128u + (uint32_t)dwk[s] * 64u; };

      for (int e = 0; e < P.epochs; ++e) {
        for (int st = 0; st < steps_per_epoch; ++st) {
          const int b = min(P.batch, n - st * P.batch);
          if (b == 1) continue;
          // ---- forward weights (all buffers are free: the step barrier follows the last drain + optimiser step)
          wait(B_STEP);
          for (int s = 0; s < 2; ++s) {
            mbar_expect_tx(&bars[B_WF_V + s], 16384);
            tma(TM_WV_K + s, wb[s][0], 0, 0, B_WF_V + s);
            tma(TM_WV_K + s, wb[s][0] + 8192, 32, 0, B_WF_V + s);
            mbar_expect_tx(&bars[B_WF_O + s], 16384);
            tma(TM_WO_K + s, wb[s][1], 0, 0, B_WF_O + s);
            tma(TM_WO_K + s, wb[s][1] + 8192, 32, 0, B_WF_O + s);
          }
          mbar_expect_tx(&bars[B_WF_W2], 2 * 4096 + 4 * 2048);
          for (int kb = 0; kb < 2; ++kb) tma(TM_W2_K, w2f + kb * 4096, kb * 32, 0, B_WF_W2);
          for (int s = 0; s < 2; ++s)
            for (int kb = 0; kb < 2; ++kb) tma(TM_WF0_K + s, wf0f[s] + kb * 2048, kb * 32, 0, B_WF_W2);
          mbar_expect_tx(&bars[B_WF_W1], 32768);
          for (int kb = 0; kb < 4; ++kb) tma(TM_W1_K, xb[0] + kb * 8192, kb * 32, 0, B_WF_W1);

          // ---- forward
          for (int s = 0; s < 2; ++s) {   // P1: dense
            wait_a(s);
            job_knosw(accc[s], acol[s], smw + (s ? kSmWd1 : kSmWd0), 64, s ? 16 : 8);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P2: V projection
            wait_a(s);
            wait(B_WF_V + s);
            job_ksw(accc[s], acol[s], wb[s][0], 64, 64, 8192);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P3: out_proj
            wait_a(s);
            wait(B_WF_O + s);
            job_ksw(accc[s], acol[s], wb[s][1], 64, 64, 8192);
            commit_acc(s);
          }
          wait(B_WF_W2);
          for (int s = 0; s < 2; ++s) {   // P4: ffn.0 (N padded to 16)
            wait_a(s);
            job_ksw(accc[s], acol[s], wf0f[s], 16, 64, 2048);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P5: ffn.3 (K padded to 8)
            wait_a(s);
            job_knosw(accc[s], acol[s], smw + (s ? kSmWf31 : kSmWf30), 64, 8);
            commit_acc(s);
          }
          wait_a(0); wait_a(1);           // P6: fc1 on cat
          wait(B_WF_W1);
          job_ksw(accc[2], acol[2], xb[0], 64, 128, 8192);
          commit_acc(2);
          wait_a(2);                      // P7: fc2
          job_ksw(accc[2], acol[2], w2f, 32, 64, 4096);
          commit_acc(2);
          wait_a(2);                      // P8: output (N padded to 16)
          job_knosw(accc[2], acol[2], smw + kSmW3, 16, 32);
          commit_acc(2);

          // ---- backward, head
          wait_acc_last(2);               // every forward MMA has finished: weight / x buffers are free
          mbar_expect_tx(&bars[B_WB + 0], 8192);
          for (int nb = 0; nb < 2; ++nb) tma(TM_W2_M, wb[0][0] + nb * 4096, nb * 32, 0, B_WB + 0);
          stash_ld(SL_HD, 2, xb[0], B_X + 0);
          mbar_expect_tx(&bars[B_WB + 1], 32768);
          for (int nb = 0; nb < 4; ++nb) tma(TM_W1_M, wb[0][1] + nb * 8192, nb * 32, 0, B_WB + 1);
          wait_a(2);                      // B7: da2 in A[0..31] and dymn[V]
          wait(B_WB + 0);
          job_mn(accc[2], acol[2], wb[0][0], 64, 32, 4096);
          commit_acc(2);
          wait(B_X + 0);
          job_dw(tmem + kColDw, dy[0], xb[0], 64, kSlotB);                    // dW2 -> V slot 0
          commit_dw(2);
          wait_dw_last(2);
          stash_ld(SL_CAT, 4, xb[0], B_X + 1);
          wait_a(2);                      // B6: da in A[0..63] and dymn[L]
          wait(B_WB + 1);
          job_mn(accc[2], acol[2], wb[0][1], 128, 64, 8192);                  // dcat -> acc 128..255
          commit_acc(2);
          wait(B_X + 1);
          job_dw(tmem + kColDw + 128, dy[1], xb[0], 128, kSlotB);             // dW1 -> L slots (128 columns)
          commit_dw(2);
          // ---- backward, branches.  Weight double buffer: wb[s][0] / wb[s][1]; X single buffer xb[s]
          wait_acc_last(2);               // dX6 done: W1mn (wb[0][1], wb[1][0]) is free
          wait_dw_last(2);                // dW1 done: xbuf is free
          for (int s = 0; s < 2; ++s) {
            mbar_expect_tx(&bars[B_WB + 2 * s], 2048);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WF0_M + s, wb[s][0] + nb * 1024, nb * 32, 0, B_WB + 2 * s);
            stash_ld(SL_BR * s + SL_GF, 1, xb[s], B_X + s);
          }
          dwk[0] = 1; dwk[1] = 0;         // dW2 sits in V slot 0 until its drain; dW1 is drained before b_a[L]
          for (int s = 0; s < 2; ++s) {   // B5: ffn.3 (Wf3 MN-major written by the element-wise warps into wb[s][1])
            wait_a(s);
            job_mn(accc[s], acol[s], wb[s][1], 16, 64, 0);
            commit_acc(s);
            wait(B_X + s);
            job_dw(dwcol(s), dy[s], xb[s], 16, 0);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) {
            wait_dw_last(s);
            stash_ld(SL_BR * s + SL_N1, 2, xb[s], B_X + s);
            mbar_expect_tx(&bars[B_WB + 2 * s + 1], 16384);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WO_M + s, wb[s][1] + nb * 8192, nb * 32, 0, B_WB + 2 * s + 1);
          }
          for (int s = 0; s < 2; ++s) {   // B4: ffn.0
            wait_a(s);
            wait(B_WB + 2 * s);
            job_mn(accc[s], acol[s], wb[s][0], 64, 8, 1024);
            commit_acc(s);
            wait(B_X + s);
            job_dw(dwcol(s), dy[s], xb[s], 64, kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) {
            wait_dw_last(s);
            stash_ld(SL_BR * s + SL_VD, 2, xb[s], B_X + s);
            mbar_expect_tx(&bars[B_WB + 2 * s], 16384);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WV_M + s, wb[s][0] + nb * 8192, nb * 32, 0, B_WB + 2 * s);
          }
          for (int s = 0; s < 2; ++s) {   // B3: out_proj
            wait_a(s);
            wait(B_WB + 2 * s + 1);
            job_mn(accc[s], acol[s], wb[s][1], 64, 64, 8192);
            commit_acc(s);
            wait(B_X + s);
            job_dw(dwcol(s), dy[s], xb[s], 64, kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) { wait_dw_last(s); stash_ld(SL_BR * s + SL_H1, 2, xb[s], B_X + s); }
          for (int s = 0; s < 2; ++s) {   // B2: V projection
            wait_a(s);
            wait(B_WB + 2 * s);
            job_mn(accc[s], acol[s], wb[s][0], 64, 64, 8192);
            commit_acc(s);
            wait(B_X + s);
            job_dw(dwcol(s), dy[s], xb[s], 64, kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) { wait_dw_last(s); stash_ld(SL_BR * s + SL_X, 1, xb[s], B_X + s); }
          for (int s = 0; s < 2; ++s) {   // B1: dense (dW only)
            wait_a(s);
            wait(B_X + s);
            job_dw(dwcol(s), dy[s], xb[s], 16, 0);
            commit_dw(s);
            dwk[s] ^= 1;
          }
        }
      }
    }
  } else {
    // =================================================================================== element-wise warps
    const int q = warp & 3, chunk = warp >> 2, row = 32 * q + lane, col0 = 16 * chunk;
    const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
    uint8_t* dymn[2] = {smem + kOffDy, smem + kOffDy + 2 * kSlotB};
    uint32_t ph = 0;
    auto wait_b = [&](int b) { mbar_wait_to(&bars[b], (ph >> b) & 1u); ph ^= 1u << b; };
    auto wait_acc = [&](int s) { wait_b(B_ACC + s); tc_fence_after(); };
    auto wait_dw = [&](int s) { wait_b(B_DW + s); tc_fence_after(); };
    auto signal_a = [&](int s) {
      tmem_st_wait();
      fence_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[B_A + s]);
    };
    auto slot = [&](int sl) { return stash + (int64_t)sl * kSlotF + row * 32; };
    auto sl64 = [&](int sl) { return slot(sl + (col0 >> 5)) + (col0 & 31); };     // (row, col0..col0+15) of a 64-wide activation
    auto ld_acc16 = [&](uint32_t col, float* v) { tmem_ld16_async(tq + col, v); tmem_ld_wait(); };
    // drain a 64 x (16 per chunk) dW tile (M = 64: row r lives in lane (r / 16) * 32 + r % 16) into the gradient row
    auto drain = [&](uint32_t col, int w_off, int rows, int ld, int ncols) {
      float g[16];
      ld_acc16(col + col0, g);
      const int r = 16 * q + lane;
      if (lane < 16 && r < rows && col0 < ncols) {
        float* dst = gw + w_off + (int64_t)r * ld + col0;
        if ((ld & 3) == 0 && col0 + 16 <= ncols) st16g(dst, g);
        else for (int j = 0; j < 16; ++j) if (col0 + j < ncols) __stcg(dst + j, g[j]);
      }
    };
    // LayerNorm statistics of this thread's row across the 4 column chunks
    auto ln_stats = [&](int buf, float a, float b, float& sa, float& sb) {
      float2* x = lnx + (buf * 128 + row) * 4;
      x[chunk] = make_float2(a, b);
      bar_named(1 + q, 128);
      sa = (x[0].x + x[1].x) + (x[2].x + x[3].x);
      sb = (x[0].y + x[1].y) + (x[2].y + x[3].y);
    };

    Rng rng; rng.seed_lo = P.seed_lo; rng.seed_hi = P.seed_hi; rng.client = (uint32_t)P.client_ids[c]; rng.on = P.dropout;
    double b1t = 1.0, b2t = 1.0;
    int opt_steps = 0, gstep = 0;
    int dwk[2] = {0, 0};
    float rstd_r[2][3];                     // LayerNorm 1/std of this thread's row, kept for the backward
    auto dwcol = [&](int s) { return kColDw + (uint32_t)s * 128u + (uint32_t)dwk[s] * 64u; };

    // small weights + parameters -> shared memory (start of every step: after the optimiser changed them)
    auto load_small = [&]() {
      for (int s = 0; s < 2; ++s) {
        const int kin = L.kin[s], kp = s ? 16 : 8;
        uint8_t* wd = smem + kOffSm + (s ? kSmWd1 : kSmWd0);
        for (int i = tid; i < 64 * kp; i += kEpi) {
          const int nn = i / kp, k = i - nn * kp;
          const float val = k < kin ? __ldcg(w + L.wd[s] + nn * kin + k) : 0.f;
          *reinterpret_cast<float*>(wd + (nn >> 3) * (kp / 4 * 128) + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = val;
        }
        uint8_t* wf = smem + kOffSm + (s ? kSmWf31 : kSmWf30);
        for (int i = tid; i < 64 * 8; i += kEpi) {
          const int nn = i >> 3, k = i & 7;
          const float val = k < L.ff ? __ldcg(w + L.wf3[s] + nn * L.ff + k) : 0.f;
          *reinterpret_cast<float*>(wf + (nn >> 3) * 256 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = val;
        }
        const int B = SP_BR * s;
        for (int i = tid; i < 64; i += kEpi) {
          sp[B + SP_BD + i] = __ldcg(w + L.bd[s] + i); sp[B + SP_BV + i] = __ldcg(w + L.bv[s] + i);
          sp[B + SP_BO + i] = __ldcg(w + L.bo[s] + i); sp[B + SP_G1 + i] = __ldcg(w + L.g1[s] + i);
          sp[B + SP_BE1 + i] = __ldcg(w + L.be1[s] + i); sp[B + SP_BF3 + i] = __ldcg(w + L.bf3[s] + i);
          sp[B + SP_G2 + i] = __ldcg(w + L.g2[s] + i); sp[B + SP_BE2 + i] = __ldcg(w + L.be2[s] + i);
          sp[B + SP_G3 + i] = __ldcg(w + L.g3[s] + i); sp[B + SP_BE3 + i] = __ldcg(w + L.be3[s] + i);
          if (i < 16) sp[B + SP_BF0 + i] = i < L.ff ? __ldcg(w + L.bf0[s] + i) : 0.f;
        }
      }
      uint8_t* w3 = smem + kOffSm + kSmW3;
      for (int i = tid; i < 16 * 32; i += kEpi) {
        const int nn = i >> 5, k = i & 31;
        *reinterpret_cast<float*>(w3 + (nn >> 3) * 1024 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = nn == 0 ? __ldcg(w + L.w3 + k) : 0.f;
      }
      for (int i = tid; i < 64; i += kEpi) {
        sp[SP_B1 + i] = __ldcg(w + L.b1 + i);
        if (i < 32) { sp[SP_B2 + i] = __ldcg(w + L.b2 + i); sp[SP_W3 + i] = __ldcg(w + L.w3 + i); }
        if (i < 16) sp[SP_B3 + i] = i == 0 ? __ldcg(w + L.b3) : 0.f;
      }
      for (int i = tid; i < SP_N; i += kEpi) sg[i] = 0.f;
    };

    load_small();
    fence_async();
    bar_named(5, kEpi);
    if (lane == 0) mbar_arrive(&bars[B_STEP]);

    for (int e = 0; e < P.epochs; ++e) {
      const int32_t* idx_e = P.idx + (int64_t)c * P.idx_stride_c + (int64_t)e * P.idx_stride_e;
      for (int st = 0; st < steps_per_epoch; ++st, ++gstep) {
        const int b = min(P.batch, n - st * P.batch);
        if (b == 1) continue;                              // client.py:86-87
        rng.step = (uint32_t)gstep;
        const float inv_b = 1.0f / (float)b;
        const bool valid = row < b;

        // ---------------------------------------------------------------- E0: gather the batch
        if (chunk == 0) {
          float xv[8], xl[16], y = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) xv[j] = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) xl[j] = 0.f;
          if (valid) {
            const float* rec = P.data + (int64_t)idx_e[(int64_t)st * P.batch + row] * P.rec;
            for (int j = 0; j < L.kin[0]; ++j) xv[j] = rec[j];
            for (int j = 0; j < L.kin[1]; ++j) xl[j] = rec[L.kin[0] + j];
            y = rec[L.kin[0] + L.kin[1]];
          }
          ylab[row] = y;
          tmem_st8(tq + kColA, xv);
          tmem_st16(tq + kColA + 64, xl);
          float pad[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) pad[j] = j < 8 ? xv[j] : 0.f;
          st16g(slot(SL_X), pad);
          st16g(slot(SL_BR + SL_X), xl);
        }
        signal_a(0);
        signal_a(1);

        // ---------------------------------------------------------------- forward, branches
        for (int s = 0; s < 2; ++s) {   // E1: h1 = gelu(x Wd^T + bd)
          const int B = SP_BR * s, S0 = SL_BR * s;
          wait_acc(s);
          float a[16], g[16];
          ld_acc16(kColAcc + 64 * s + col0, a);
#pragma unroll
          for (int j = 0; j < 16; ++j) { const float z = a[j] + sp[B + SP_BD + col0 + j]; g[j] = afl_gelu_grad(z); a[j] = afl_gelu(z); }
          st16g(sl64(S0 + SL_GP1), g);
          st16g(sl64(S0 + SL_H1), a);
          tmem_st16(tq + kColA + 64 * s + col0, a);
          signal_a(s);
        }
        for (int s = 0; s < 2; ++s) {   // E2: vd = headdrop(h1 Wv^T + bv)
          const int B = SP_BR * s, S0 = SL_BR * s;
          wait_acc(s);
          float a[16];
          ld_acc16(kColAcc + 64 * s + col0, a);
          const float ds = drop1(rng, L.rng_head[s], row, chunk, P.p1);
#pragma unroll
          for (int j = 0; j < 16; ++j) a[j] = (a[j] + sp[B + SP_BV + col0 + j]) * ds;
          st16g(sl64(S0 + SL_VD), a);
          tmem_st16(tq + kColA + 64 * s + col0, a);
          signal_a(s);
        }
        for (int s = 0; s < 2; ++s) {   // E3: n1 = LN1(h1 + drop(vd Wo^T + bo))
          const int B = SP_BR * s, S0 = SL_BR * s;
          float h[16], ds[16];
          ld16g(sl64(S0 + SL_H1), h);
          drop16(rng, L.rng_ln1[s], row, col0, P.p1, ds);
          wait_acc(s);
          float a[16];
          ld_acc16(kColAcc + 64 * s + col0, a);
          float sum = 0.f, sq = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) { a[j] = h[j] + (a[j] + sp[B + SP_BO + col0 + j]) * ds[j]; sum += a[j]; sq += a[j] * a[j]; }
          float ts, tq2;
          ln_stats(s, sum, sq, ts, tq2);
          const float mean = ts * (1.f / 64.f);
          float var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          const float rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
#pragma unroll
          for (int j = 0; j < 16; ++j) { h[j] = (a[j] - mean) * rstd; a[j] = h[j] * sp[B + SP_G1 + col0 + j] + sp[B + SP_BE1 + col0 + j]; }
          rstd_r[s][0] = rstd;
          st16g(sl64(S0 + SL_XH1), h);
          st16g(sl64(S0 + SL_N1), a);
          tmem_st16(tq + kColA + 64 * s + col0, a);
          signal_a(s);
        }
        for (int s = 0; s < 2; ++s) {   // E4: gf = drop(gelu(n1 Wf0^T + bf0))   (ff <= 8 columns)
          const int B = SP_BR * s, S0 = SL_BR * s;
          wait_acc(s);
          if (chunk == 0) {
            float a[16], g[16], ds[8];
            ld_acc16(kColAcc + 64 * s, a);
            drop4(rng, L.rng_f1[s], row, 0, P.p1, ds);
            drop4(rng, L.rng_f1[s], row, 4, P.p1, ds + 4);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (j < 8 && j < L.ff) {
                const float z = a[j] + sp[B + SP_BF0 + j];
                g[j] = afl_gelu_grad(z) * ds[j]; a[j] = afl_gelu(z) * ds[j];
              } else { g[j] = 0.f; a[j] = 0.f; }
            }
            st16g(slot(S0 + SL_GPF), g);
            st16g(slot(S0 + SL_GF), a);
            tmem_st8(tq + kColA + 64 * s, a);
          }
          signal_a(s);
        }
        for (int s = 0; s < 2; ++s) {   // E5: n3 = LN3(LN2(n1 + drop(gf Wf3^T + bf3)))
          const int B = SP_BR * s, S0 = SL_BR * s;
          float h[16], ds[16];
          ld16g(sl64(S0 + SL_N1), h);
          drop16(rng, L.rng_ln2[s], row, col0, P.p1, ds);
          wait_acc(s);
          float a[16];
          ld_acc16(kColAcc + 64 * s + col0, a);
          float sum = 0.f, sq = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) { a[j] = h[j] + (a[j] + sp[B + SP_BF3 + col0 + j]) * ds[j]; sum += a[j]; sq += a[j] * a[j]; }
          float ts, tq2;
          ln_stats(0, sum, sq, ts, tq2);
          float mean = ts * (1.f / 64.f);
          float var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          float rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
          sum = 0.f; sq = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            h[j] = (a[j] - mean) * rstd; a[j] = h[j] * sp[B + SP_G2 + col0 + j] + sp[B + SP_BE2 + col0 + j];
            sum += a[j]; sq += a[j] * a[j];
          }
          rstd_r[s][1] = rstd;
          st16g(sl64(S0 + SL_XH2), h);
          ln_stats(1, sum, sq, ts, tq2);
          mean = ts * (1.f / 64.f);
          var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
#pragma unroll
          for (int j = 0; j < 16; ++j) { h[j] = (a[j] - mean) * rstd; a[j] = h[j] * sp[B + SP_G3 + col0 + j] + sp[B + SP_BE3 + col0 + j]; }
          rstd_r[s][2] = rstd;
          st16g(sl64(S0 + SL_XH3), h);
          st16g(sl64(SL_CAT + 2 * s), a);
          tmem_st16(tq + kColA + 64 * s + col0, a);
          signal_a(s);
        }
        // ---------------------------------------------------------------- forward, head
        {                               // E6: hd = drop(gelu(cat W1^T + b1))
          float ds[16];
          drop16(rng, L.rng_fc1, row, col0, P.p3, ds);
          wait_acc(2);
          float a[16], g[16];
          ld_acc16(kColAcc + col0, a);
#pragma unroll
          for (int j = 0; j < 16; ++j) { const float z = a[j] + sp[SP_B1 + col0 + j]; g[j] = afl_gelu_grad(z) * ds[j]; a[j] = afl_gelu(z) * ds[j]; }
          st16g(sl64(SL_GPA), g);
          st16g(sl64(SL_HD), a);
          tmem_st16(tq + kColA + col0, a);
          signal_a(2);
        }
        float h2[16], gp2[16];
        {                               // E7: h2 = gelu(hd W2^T + b2)   (32 columns: chunks 0, 1)
          wait_acc(2);
          if (chunk < 2) {
            ld_acc16(kColAcc + col0, h2);
#pragma unroll
            for (int j = 0; j < 16; ++j) { const float z = h2[j] + sp[SP_B2 + col0 + j]; gp2[j] = afl_gelu_grad(z); h2[j] = afl_gelu(z); }
            st16g(slot(SL_H2) + col0, h2);
            tmem_st16(tq + kColA + col0, h2);
          }
          signal_a(2);
        }
        {                               // E8 + B8: p = sigmoid(h2 w3 + b3), BCE, da2 = (dz w3) gelu'(a2)
          wait_acc(2);
          if (chunk == 0) {
            float a[16];
            ld_acc16(kColAcc, a);
            const float p = afl_sigmoid(a[0] + sp[SP_B3]);
            const float y = ylab[row];
            float dz = 0.f, loss = 0.f, bad = 0.f;
            if (valid) {
              float lp = logf(p), lq = logf(1.f - p);
              lp = lp < -100.f ? -100.f : lp;
              lq = lq < -100.f ? -100.f : lq;
              loss = -(y * lp + (1.f - y) * lq);
              dz = (p - y) * inv_b;
              if (loss != loss) { bad = 1.f; loss = 0.f; }
            }
            dzrow[row] = dz;
            loss = warp_sum(loss); bad = warp_sum(bad);
            const float dzs = warp_sum(dz);
            if (lane == 0) { atomicAdd(loss_acc, loss); atomicAdd(loss_acc + 1, bad); atomicAdd(sg + SP_B3, dzs); }
          }
          bar_named(1 + q, 128);
          if (chunk < 2) {
            const float dz = dzrow[row];
            float t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = dz * h2[j];
            colsum16(t, sg + SP_W3 + col0, lane);
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = dz * sp[SP_W3 + col0 + j] * gp2[j];
            colsum16(t, sg + SP_B2 + col0, lane);
            tmem_st16(tq + kColA + col0, t);
            st_dymn(dymn[0], row, col0, t);
          }
          signal_a(2);
        }
        // ---------------------------------------------------------------- backward, head
        {                               // B7: da = (da2 W2) gelu'(a) mask
          float g[16];
          ld16g(sl64(SL_GPA), g);
          wait_acc(2);
          float a[16];
          ld_acc16(kColAcc + col0, a);
#pragma unroll
          for (int j = 0; j < 16; ++j) a[j] *= g[j];
          colsum16(a, sg + SP_B1 + col0, lane);
          tmem_st16(tq + kColA + col0, a);
          st_dymn(dymn[1], row, col0, a);
          wait_dw(2);                                         // dW2 (reads dymn[V]) finished before it may be lapped
          signal_a(2);
          drain(kColDw, L.w2, 32, 64, 64);
        }
        {                               // B6: dcat -> LN3, LN2 backward per branch -> do2
          wait_acc(2);
          for (int s = 0; s < 2; ++s) {
            const int B = SP_BR * s, S0 = SL_BR * s;
            float xh[16], g[16], t[16], ds[16];
            ld16g(sl64(S0 + SL_XH3), xh);
            drop16(rng, L.rng_ln2[s], row, col0, P.p1, ds);
            ld_acc16(kColAcc + 64 * s + col0, g);
            float sa = 0.f, sb = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) { t[j] = g[j] * xh[j]; }
            colsum16(t, sg + B + SP_G3 + col0, lane);
            colsum16(g, sg + B + SP_BE3 + col0, lane);
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[j] *= sp[B + SP_G3 + col0 + j]; sa += g[j]; sb += g[j] * xh[j]; }
            float ta, tb;
            ln_stats(0, sa, sb, ta, tb);
            float rs = rstd_r[s][2];
            float m1 = ta * (1.f / 64.f), m2 = tb * (1.f / 64.f);
#pragma unroll
            for (int j = 0; j < 16; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);          // dL/dn2
            ld16g(sl64(S0 + SL_XH2), xh);
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = g[j] * xh[j];
            colsum16(t, sg + B + SP_G2 + col0, lane);
            colsum16(g, sg + B + SP_BE2 + col0, lane);
            sa = 0.f; sb = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[j] *= sp[B + SP_G2 + col0 + j]; sa += g[j]; sb += g[j] * xh[j]; }
            ln_stats(1, sa, sb, ta, tb);
            rs = rstd_r[s][1];
            m1 = ta * (1.f / 64.f); m2 = tb * (1.f / 64.f);
#pragma unroll
            for (int j = 0; j < 16; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);          // dL/dr2
            st16g(sl64(S0 + SL_RES), g);
#pragma unroll
            for (int j = 0; j < 16; ++j) g[j] *= ds[j];                                  // dL/do2
            colsum16(g, sg + B + SP_BF3 + col0, lane);
            // ffn.3 weights, MN-major, for dgf = do2 . Wf3 (N = in padded to 16, K = 64 outputs)
            uint8_t* wmn = smem + kOffW + (2 * s + 1) * kSlotB;
            for (int i = tid; i < 64 * 16; i += kEpi) {
              const int o = i >> 4, ii = i & 15;
              *reinterpret_cast<float*>(wmn + o * 128 + (((ii >> 3) ^ (o & 3)) << 5) + (ii & 7) * 4) =
                  ii < L.ff ? __ldcg(w + L.wf3[s] + o * L.ff + ii) : 0.f;
            }
            if (s == 1) wait_dw(2);                           // dW1 (reads dymn[L]) finished
            tmem_st16(tq + kColA + 64 * s + col0, g);
            st_dymn(dymn[s], row, col0, g);
            if (s == 1) {                                     // dW1: 64 x 128, 32 columns per chunk
              const int r = 16 * q + lane;
#pragma unroll
              for (int hf = 0; hf < 2; ++hf) {
                float g2[16];
                ld_acc16(kColDw + 128 + 32 * chunk + 16 * hf, g2);
                if (lane < 16) st16g(gw + L.w1 + (int64_t)r * 128 + 32 * chunk + 16 * hf, g2);
              }
            }
            signal_a(s);
          }
          dwk[0] = 1; dwk[1] = 0;
        }
        // ---------------------------------------------------------------- backward, branches
        for (int s = 0; s < 2; ++s) {   // B5: df = (do2 Wf3) gelu'(f) mask
          const int B = SP_BR * s, S0 = SL_BR * s;
          wait_acc(s);
          float a[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) a[j] = 0.f;
          if (chunk == 0) {
            float g[16];
            ld16g(slot(S0 + SL_GPF), g);
            ld_acc16(kColAcc + 64 * s, a);
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = (j < L.ff) ? a[j] * g[j] : 0.f;
            colsum16(a, sg + B + SP_BF0, lane);
          }
          wait_dw(s);                                           // dWf3 (reads dymn[s] = do2) finished
          if (chunk == 0) tmem_st8(tq + kColA + 64 * s, a);
          if (chunk < 2) st_dymn(dymn[s], row, col0, a);        // 32 columns: df (16) + zeros (16)
          signal_a(s);
          {                                                     // drain dWf3: [64 out] x [ff in]
            float g[16];
            ld_acc16(dwcol(s), g);
            const int r = 16 * q + lane;
            if (chunk == 0 && lane < 16) for (int j = 0; j < L.ff; ++j) __stcg(gw + L.wf3[s] + r * L.ff + j, g[j]);
          }
          dwk[s] ^= 1;
        }
        for (int s = 0; s < 2; ++s) {   // B4: dr1 = LN1'(df Wf0 + dr2); do = dr1 mask
          const int B = SP_BR * s, S0 = SL_BR * s;
          float res[16], xh[16], ds[16], t[16];
          ld16g(sl64(S0 + SL_RES), res);
          ld16g(sl64(S0 + SL_XH1), xh);
          drop16(rng, L.rng_ln1[s], row, col0, P.p1, ds);
          wait_acc(s);
          float g[16];
          ld_acc16(kColAcc + 64 * s + col0, g);
          float sa = 0.f, sb = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) { g[j] += res[j]; t[j] = g[j] * xh[j]; }
          colsum16(t, sg + B + SP_G1 + col0, lane);
          colsum16(g, sg + B + SP_BE1 + col0, lane);
#pragma unroll
          for (int j = 0; j < 16; ++j) { g[j] *= sp[B + SP_G1 + col0 + j]; sa += g[j]; sb += g[j] * xh[j]; }
          float ta, tb;
          ln_stats(s, sa, sb, ta, tb);
          const float rs = rstd_r[s][0];
          const float m1 = ta * (1.f / 64.f), m2 = tb * (1.f / 64.f);
#pragma unroll
          for (int j = 0; j < 16; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);            // dL/dr1
          st16g(sl64(S0 + SL_RES), g);
#pragma unroll
          for (int j = 0; j < 16; ++j) g[j] *= ds[j];                                    // dL/do
          colsum16(g, sg + B + SP_BO + col0, lane);
          wait_dw(s);                                           // dWf0 finished
          tmem_st16(tq + kColA + 64 * s + col0, g);
          st_dymn(dymn[s], row, col0, g);
          signal_a(s);
          drain(dwcol(s), L.wf0[s], L.ff, 64, 64);
          dwk[s] ^= 1;
        }
        for (int s = 0; s < 2; ++s) {   // B3: dv = (do Wo) headmask
          const int B = SP_BR * s;
          const float ds = drop1(rng, L.rng_head[s], row, chunk, P.p1);
          wait_acc(s);
          float g[16];
          ld_acc16(kColAcc + 64 * s + col0, g);
#pragma unroll
          for (int j = 0; j < 16; ++j) g[j] *= ds;
          colsum16(g, sg + B + SP_BV + col0, lane);
          wait_dw(s);                                           // dWo finished
          tmem_st16(tq + kColA + 64 * s + col0, g);
          st_dymn(dymn[s], row, col0, g);
          signal_a(s);
          drain(dwcol(s), L.wo[s], 64, 64, 64);
          dwk[s] ^= 1;
        }
        for (int s = 0; s < 2; ++s) {   // B2: da1 = (dv Wv + dr1) gelu'(a1)
          const int B = SP_BR * s, S0 = SL_BR * s;
          float res[16], gp[16];
          ld16g(sl64(S0 + SL_RES), res);
          ld16g(sl64(S0 + SL_GP1), gp);
          wait_acc(s);
          float g[16];
          ld_acc16(kColAcc + 64 * s + col0, g);
#pragma unroll
          for (int j = 0; j < 16; ++j) g[j] = (g[j] + res[j]) * gp[j];
          colsum16(g, sg + B + SP_BD + col0, lane);
          wait_dw(s);                                           // dWv finished
          st_dymn(dymn[s], row, col0, g);
          signal_a(s);
          drain(dwcol(s), L.wv[s], 64, 64, 64);
          dwk[s] ^= 1;
        }
        for (int s = 0; s < 2; ++s) {   // B1: drain dWd [64 x kin]
          wait_dw(s);
          float g[16];
          ld_acc16(dwcol(s), g);
          const int r = 16 * q + lane, kin = L.kin[s];
          if (chunk == 0 && lane < 16) for (int j = 0; j < kin; ++j) __stcg(gw + L.wd[s] + r * kin + j, g[j]);
          dwk[s] ^= 1;
        }
        tc_fence_before();
        bar_named(5, kEpi);
        // ---------------------------------------------------------------- small gradients -> gradient row
        for (int s = 0; s < 2; ++s) {
          const int B = SP_BR * s;
          for (int i = tid; i < 64; i += kEpi) {
            gw[L.bd[s] + i] = sg[B + SP_BD + i]; gw[L.bv[s] + i] = sg[B + SP_BV + i]; gw[L.bo[s] + i] = sg[B + SP_BO + i];
            gw[L.g1[s] + i] = sg[B + SP_G1 + i]; gw[L.be1[s] + i] = sg[B + SP_BE1 + i]; gw[L.bf3[s] + i] = sg[B + SP_BF3 + i];
            gw[L.g2[s] + i] = sg[B + SP_G2 + i]; gw[L.be2[s] + i] = sg[B + SP_BE2 + i];
            gw[L.g3[s] + i] = sg[B + SP_G3 + i]; gw[L.be3[s] + i] = sg[B + SP_BE3 + i];
            if (i < L.ff) gw[L.bf0[s] + i] = sg[B + SP_BF0 + i];
          }
        }
        for (int i = tid; i < 64; i += kEpi) {
          gw[L.b1 + i] = sg[SP_B1 + i];
          if (i < 32) { gw[L.b2 + i] = sg[SP_B2 + i]; gw[L.w3 + i] = sg[SP_W3 + i]; }
          if (i == 0) gw[L.b3] = sg[SP_B3];
        }
        __threadfence_block();
        bar_named(5, kEpi);
        // ---------------------------------------------------------------- optimiser step
        ++opt_steps;
        b1t *= (double)P.beta1; b2t *= (double)P.beta2;
        AflOpt o;
        o.kind = P.opt_kind; o.lr = P.lr; o.beta1 = P.beta1; o.beta2 = P.beta2; o.eps = P.eps;
        o.momentum = P.momentum; o.bc1 = (float)(1.0 - b1t); o.bc2 = (float)(1.0 - b2t);
        o.first_step = opt_steps == 1; o.clip_coef = 1.f;
        if (P.clip_proper) {
          float part = 0.f;
          for (int r = 0; r < P.n_ranges; ++r) part += afl_reduce_sqnorm(gw, 1, P.d_pad, P.range_lo[r], P.range_hi[r], tid, kEpi);
          part = warp_sum(part);
          if (lane == 0) red[warp] = part;
          bar_named(5, kEpi);
          float tot = 0.f;
          for (int i = 0; i < kEpi / 32; ++i) tot += red[i];
          const float coef = P.clip / (sqrtf(tot) + 1e-6f);
          o.clip_coef = coef < 1.f ? coef : 1.f;
          bar_named(5, kEpi);
        }
        for (int r = 0; r < P.n_ranges; ++r) afl_reduce_update(w, mrow, vrow, gw, 1, P.d_pad, P.range_lo[r], P.range_hi[r], o, tid, kEpi);
        __threadfence_block();
        bar_named(5, kEpi);
        load_small();
        fence_async();
        bar_named(5, kEpi);
        if (lane == 0) mbar_arrive(&bars[B_STEP]);
      }
    }
    if (tid == 0) {
      atomicAdd(P.loss_out + 2 * c, loss_acc[0]);
      atomicAdd(P.loss_out + 2 * c + 1, loss_acc[1]);
      P.steps_out[c] = opt_steps;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int make_map3(CUtensorMap* map, const void* base, int64_t d0, int64_t d1, int64_t d2, int64_t ld1_elems, int64_t ld2_elems,
              int box1, bool mn_major) {
  EncodeFn enc = get_encode();
  if (!enc) { afl_rt::set_error("cuTensorMapEncodeTiled not available"); return -5; }
  cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t strides[2] = {(cuuint64_t)ld1_elems * 4, (cuuint64_t)ld2_elems * 4};
  cuuint32_t box[3] = {32u, (cuuint32_t)box1, 1u};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { afl_rt::set_error("cuTensorMapEncodeTiled (3d) failed (%d)", (int)r); return -6; }
  return 0;
}

}  // namespace

extern "C" {

int afl_tc_stash_floats_per_client(void) { return kNSlot * kSlotF; }

int afl_tc_train_round(const int32_t* lay, int lay_words, float* w, float* m, float* v, float* gw, float* stash, int64_t d_pad,
                       const float* data, int rec, const int32_t* idx, int64_t idx_stride_c, int64_t idx_stride_e,
                       const int32_t* n_data, const int32_t* client_ids, int n_clients, int batch, int epochs, int opt_kind,
                       float lr, float beta1, float beta2, float eps, float momentum, int clip_proper, float clip, uint64_t seed,
                       int dropout, float p1, float p3, const int32_t* ranges, int n_ranges, float* loss_out,
                       int32_t* steps_out, void* stream) {
  AFL_CHECK_ARG(lay_words == (int)(sizeof(TcLay) / 4), "afl_tc_train_round: layout word count mismatch");
  AFL_CHECK_ARG(batch >= 2 && batch <= 128, "afl_tc_train_round: batch must be in [2, 128]");
  AFL_CHECK_ARG(n_ranges >= 1 && n_ranges <= 16, "afl_tc_train_round: 1..16 parameter ranges");
  AFL_CHECK_ARG(d_pad % 4 == 0, "afl_tc_train_round: d_pad must be a multiple of 4");
  TcParams P;
  memcpy(&P.L, lay, sizeof(TcLay));
  AFL_CHECK_ARG(P.L.ff >= 1 && P.L.ff <= 8 && P.L.kin[0] >= 1 && P.L.kin[0] <= 8 && P.L.kin[1] >= 1 && P.L.kin[1] <= 16,
                "afl_tc_train_round: model widths outside the tensor-core trainer's range");
  P.w = w; P.m = m; P.v = v; P.gw = gw; P.stash = stash; P.d_pad = d_pad; P.data = data; P.rec = rec;
  P.idx = idx; P.idx_stride_c = idx_stride_c; P.idx_stride_e = idx_stride_e; P.n_data = n_data; P.client_ids = client_ids;
  P.batch = batch; P.epochs = epochs; P.opt_kind = opt_kind; P.lr = lr; P.beta1 = beta1; P.beta2 = beta2; P.eps = eps;
  P.momentum = momentum; P.clip_proper = clip_proper; P.clip = clip; P.seed_lo = (uint32_t)seed; P.seed_hi = (uint32_t)(seed >> 32);
  P.dropout = dropout; P.p1 = p1; P.p3 = p3; P.n_ranges = n_ranges;
  for (int r = 0; r < n_ranges; ++r) {
    P.range_lo[r] = ranges[2 * r]; P.range_hi[r] = ranges[2 * r + 1];
    AFL_CHECK_ARG(P.range_lo[r] % 4 == 0 && P.range_hi[r] % 4 == 0 && P.range_hi[r] <= d_pad, "afl_tc_train_round: bad range");
  }
  P.loss_out = loss_out; P.steps_out = steps_out;
  TcMaps maps;
  const int C = n_clients;
  int rc = 0;
  for (int s = 0; s < 2 && !rc; ++s) {
    // K-major (forward) and MN-major (dX) views of the same weights: {in, out, client}
    rc |= make_map3(&maps.m[TM_WV_K + s], w + P.L.wv[s], 64, 64, C, 64, d_pad, 64, false);
    rc |= make_map3(&maps.m[TM_WO_K + s], w + P.L.wo[s], 64, 64, C, 64, d_pad, 64, false);
    rc |= make_map3(&maps.m[TM_WF0_K + s], w + P.L.wf0[s], 64, P.L.ff, C, 64, d_pad, 16, false);
    rc |= make_map3(&maps.m[TM_WV_M + s], w + P.L.wv[s], 64, 64, C, 64, d_pad, 64, true);
    rc |= make_map3(&maps.m[TM_WO_M + s], w + P.L.wo[s], 64, 64, C, 64, d_pad, 64, true);
    rc |= make_map3(&maps.m[TM_WF0_M + s], w + P.L.wf0[s], 64, P.L.ff, C, 64, d_pad, 8, true);
  }
  if (!rc) rc |= make_map3(&maps.m[TM_W1_K], w + P.L.w1, 128, 64, C, 128, d_pad, 64, false);
  if (!rc) rc |= make_map3(&maps.m[TM_W2_K], w + P.L.w2, 64, 32, C, 64, d_pad, 32, false);
  if (!rc) rc |= make_map3(&maps.m[TM_W1_M], w + P.L.w1, 128, 64, C, 128, d_pad, 64, true);
  if (!rc) rc |= make_map3(&maps.m[TM_W2_M], w + P.L.w2, 64, 32, C, 64, d_pad, 32, true);
  if (!rc) rc |= make_map3(&maps.m[TM_STASH], stash, 32, 128, (int64_t)C * kNSlot, 32, kSlotF, 128, true);
  if (rc) return rc;
  AFL_CHECK_ARG((int)kSmemBytes <= afl_rt::max_smem_optin(), "afl_tc_train_round: not enough shared memory on this device");
  AFL_CUDA(cudaFuncSetAttribute(tc_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
  cudaStream_t st = (cudaStream_t)stream;
  AFL_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float) * 2 * (size_t)n_clients, st));
  tc_train_kernel<<<n_clients, kThreads, kSmemBytes, st>>>(maps, P);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_tc_train_round");
}

}  // extern "C"
