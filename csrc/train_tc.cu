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

// ---- stash slots (per client); per-branch base = 8 * s
// Row-major [128 x 32] blocks that TMA reads back as MN-major MMA operands of the dW GEMMs:
enum { SL_X = 0, SL_H1 = 1, SL_VD = 3, SL_N1 = 5, SL_GF = 7, SL_BR = 8, SL_CAT = 16, SL_HD = 20, SL_H2 = 22 };
// Tensors that only the thread that wrote them reads back (activation derivatives, normalised rows, residual gradients):
// stored lane-interleaved, [warp][128-bit chunk][lane], so every warp access is 512 contiguous bytes.
enum { EL_GP1 = 24, EL_XH1 = 28, EL_XH2 = 32, EL_XH3 = 36, EL_RES = 40, EL_GPA = 44, EL_GPF = 46 };

// ---- small parameters / their gradients mirrored in shared memory; per-branch base = 656 * s
enum { SP_BD = 0, SP_BV = 64, SP_BO = 128, SP_G1 = 192, SP_BE1 = 256, SP_BF0 = 320, SP_BF3 = 336, SP_G2 = 400, SP_BE2 = 464,
       SP_G3 = 528, SP_BE3 = 592, SP_BR = 656, SP_B1 = 1312, SP_B2 = 1376, SP_W3 = 1408, SP_B3 = 1440, SP_N = 1456 };

// ---- tensor maps
enum { TM_WV_K = 0, TM_WO_K = 2, TM_WF0_K = 4, TM_W1_K = 6, TM_W2_K = 7, TM_WV_M = 8, TM_WO_M = 10, TM_WF0_M = 12, TM_W1_M = 14,
       TM_W2_M = 15, TM_STASH = 16, TM_N = 17 };

// ---- mbarriers
enum { B_STEP = 0, B_A = 1 /*V,L,H*/, B_ACC = 4, B_DW = 7, B_WF_V = 10 /*Wv*/, B_WF_O = 12 /*Wo*/, B_WF_W1 = 14, B_WF_W2 = 15,
       B_WB = 16 /*[s][2]*/, B_X = 20 /*[s]*/, B_GO = 22 /*helper CTAs: optimiser pass may start*/,
       B_DONE = 23 /*main CTA: every helper finished its share*/, B_N = 24 };

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
  long long* dbg;        // optional [64] clock64 stamps of thread 0 of client 0 at step 5 (profiling aid), or null
  int cluster;           // CTAs per client: rank 0 trains, ranks 1.. take a share of the optimiser pass (L2-bandwidth bound on one SM)
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

// ---- cluster helpers (optimiser pass shared between the CTAs of a client's cluster)
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(const void* smem_ptr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(smem_ptr)), "r"(rank));
  return r;
}
__device__ __forceinline__ void remote_arrive(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void remote_st_f32(uint32_t cluster_addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_addr), "f"(v) : "memory");
}
// wait on a barrier of this CTA that CTAs of the cluster arrive on (cluster-scope acquire); same time-out as mbar_wait_to
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  const long long t0 = clock64();
  for (uint32_t it = 0;; ++it) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
        : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) return;
    if ((it & 255u) == 255u && clock64() - t0 > 4000000000LL) __trap();
  }
}
__device__ __forceinline__ void bar_named(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// One lane of a converged warp.  ptxas recognises elect.sync, so a tcgen05 / TMA instruction guarded by it is issued once
// with uniform-register operands; guarded by `lane == 0` instead, every such instruction is wrapped in a
// vote-and-branch loop (ELECT ... BRA.U.ANY in SASS) with per-operand R2UR moves.
__device__ __forceinline__ bool elect_one() {
  uint32_t p;
  asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.u32 %0, 1, 0, P1;\n\t}" : "=r"(p));
  return p != 0;
}

// ---- MMA job shapes (issued by one elected lane of the converged producer warp) --------------------------------------------------------------------------
// forward / dX: D[128 x N] = A(tmem, K columns) . B.  Shapes are template parameters so that the issue loops unroll to one
// add per operand: the issuing thread is a serial resource on the critical path of every phase.
template <int N, int K>
__device__ __forceinline__ void job_ksw(uint32_t d, uint32_t a, uint32_t b_addr, uint32_t box_bytes) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 0);
  const uint64_t bd = desc_k_sw128(b_addr);
  if (elect_one()) {
#pragma unroll
    for (int ks = 0; ks < K / 8; ++ks)
      umma_tf32_ts(d, a + ks * 8, desc_add(bd, (uint32_t)(ks >> 2) * box_bytes + (uint32_t)(ks & 3) * 32u), idesc, ks > 0);
  }
}
template <int N, int K>
__device__ __forceinline__ void job_knosw(uint32_t d, uint32_t a, uint32_t b_addr) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 0);
  const uint64_t bd = desc_nosw(b_addr, 128, (uint32_t)(K / 4) * 128u);
  if (elect_one()) {
#pragma unroll
    for (int ks = 0; ks < K / 8; ++ks) umma_tf32_ts(d, a + ks * 8, desc_add(bd, (uint32_t)ks * 256u), idesc, ks > 0);
  }
}
template <int N, int K>
__device__ __forceinline__ void job_mn(uint32_t d, uint32_t a, uint32_t b_addr, uint32_t lbo) {
  const uint32_t idesc = idesc_tf32(128, N, 0, 1);
  const uint64_t bd = desc_mn_sw128(b_addr, lbo);
  if (elect_one()) {
#pragma unroll
    for (int ks = 0; ks < K / 8; ++ks) umma_tf32_ts(d, a + ks * 8, desc_add(bd, (uint32_t)ks * 1024u), idesc, ks > 0);
  }
}
// dW: D[64 x N] = dY^T (smem MN-major, two 32-wide blocks 16 KB apart) . X (smem MN-major), K = 128 samples
template <int N>
__device__ __forceinline__ void job_dw(uint32_t d, uint32_t a_addr, uint32_t b_addr, uint32_t b_lbo) {
  const uint32_t idesc = idesc_tf32(64, N, 1, 1);
  const uint64_t ad = desc_mn_sw128(a_addr, kSlotB), bd = desc_mn_sw128(b_addr, b_lbo);
  if (elect_one()) {
#pragma unroll
    for (int ks = 0; ks < 16; ++ks) umma_tf32(d, desc_add(ad, (uint32_t)ks * 1024u), desc_add(bd, (uint32_t)ks * 1024u), idesc, ks > 0);
  }
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

// gelu(z) and gelu'(z) from one exponential: erf by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7), whose exp(-x^2) with
// x = |z| / sqrt(2) is the Gaussian factor of the derivative as well.
__device__ __forceinline__ void gelu_both(float z, float& g, float& gp) {
  const float az = fabsf(z);
  const float e = __expf(-0.5f * z * z);
  const float t = __frcp_rn(1.0f + 0.3275911f * 0.70710678118654752440f * az);
  const float poly = t * (0.254829592f + t * (-0.284496736f + t * (1.421413741f + t * (-1.453152027f + t * 1.061405429f))));
  const float erf_abs = 1.0f - poly * e;
  const float cdf = 0.5f * (1.0f + copysignf(erf_abs, z));
  g = z * cdf;
  gp = cdf + z * e * 0.39894228040143267794f;
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

// ---- coalesced stash access ------------------------------------------------------------------------------------------
// A thread owns one sample row, i.e. 128 contiguous bytes of a [128 x 32] stash block, so a plain 128-bit access by the 32
// lanes of a warp touches 32 different lines (32 L1 transactions for 512 bytes).  An 8 x 8 transpose of float4 chunks
// inside each group of 8 lanes (3 butterfly stages) turns "lane = row" into "lane = chunk": every 128-bit instruction then
// covers 4 whole lines.  The transpose is its own inverse, so loads use it after the access and stores before it.
__device__ __forceinline__ void transpose8(float4* v, int lane) {
#pragma unroll
  for (int b = 4; b >= 1; b >>= 1) {
    const bool up = lane & b;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c & b) continue;
      float4 send = up ? v[c] : v[c ^ b];
      float4 recv;
      recv.x = __shfl_xor_sync(0xffffffffu, send.x, b);
      recv.y = __shfl_xor_sync(0xffffffffu, send.y, b);
      recv.z = __shfl_xor_sync(0xffffffffu, send.z, b);
      recv.w = __shfl_xor_sync(0xffffffffu, send.w, b);
      if (up) v[c] = recv; else v[c ^ b] = recv;
    }
  }
}
// 32 floats of this thread's row <-> one [128 x 32] stash block (blk: first float of the block, row0: first row of the warp)
__device__ __forceinline__ void st32t(float* blk, int row0, int lane, const float* x) {
  float4 v[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) v[c] = make_float4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
  transpose8(v, lane);
  float4* dst = reinterpret_cast<float4*>(blk + (row0 + (lane & ~7)) * 32) + (lane & 7);
#pragma unroll
  for (int k = 0; k < 8; ++k) __stcg(dst + k * 8, v[k]);
}
__device__ __forceinline__ void ld32t(const float* blk, int row0, int lane, float* x) {
  float4 v[8];
  const float4* src = reinterpret_cast<const float4*>(blk + (row0 + (lane & ~7)) * 32) + (lane & 7);
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = __ldcg(src + k * 8);
  transpose8(v, lane);
#pragma unroll
  for (int c = 0; c < 8; ++c) { x[4 * c] = v[c].x; x[4 * c + 1] = v[c].y; x[4 * c + 2] = v[c].z; x[4 * c + 3] = v[c].w; }
}
// 16-float variant (head phases: 16 columns per thread): 4 x 4 transpose inside groups of 4 lanes, 64-byte runs
__device__ __forceinline__ void transpose4(float4* v, int lane) {
#pragma unroll
  for (int b = 2; b >= 1; b >>= 1) {
    const bool up = lane & b;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c & b) continue;
      float4 send = up ? v[c] : v[c ^ b];
      float4 recv;
      recv.x = __shfl_xor_sync(0xffffffffu, send.x, b);
      recv.y = __shfl_xor_sync(0xffffffffu, send.y, b);
      recv.z = __shfl_xor_sync(0xffffffffu, send.z, b);
      recv.w = __shfl_xor_sync(0xffffffffu, send.w, b);
      if (up) v[c] = recv; else v[c ^ b] = recv;
    }
  }
}
__device__ __forceinline__ void st16t(float* blk, int row0, int lane, int colf, const float* x) {
  float4 v[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) v[c] = make_float4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
  transpose4(v, lane);
  float4* dst = reinterpret_cast<float4*>(blk + (row0 + (lane & ~3)) * 32 + colf) + (lane & 3);
#pragma unroll
  for (int k = 0; k < 4; ++k) __stcg(dst + k * 8, v[k]);
}
__device__ __forceinline__ void ld16t(const float* blk, int row0, int lane, int colf, float* x) {
  float4 v[4];
  const float4* src = reinterpret_cast<const float4*>(blk + (row0 + (lane & ~3)) * 32 + colf) + (lane & 3);
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = __ldcg(src + k * 8);
  transpose4(v, lane);
#pragma unroll
  for (int c = 0; c < 4; ++c) { x[4 * c] = v[c].x; x[4 * c + 1] = v[c].y; x[4 * c + 2] = v[c].z; x[4 * c + 3] = v[c].w; }
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
  const int CS = P.cluster;
  const uint32_t crank = CS > 1 ? cluster_rank() : 0u;
  const int c = blockIdx.x / CS;
  const TcLay& L = P.L;
  float* ccoef = reinterpret_cast<float*>(smem + kOffMisc + 100);   // clip coefficient handed to a helper CTA

  // ---- one-time set-up
  for (uint32_t i = tid; i < (kOffBar) / 4; i += kThreads) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  float* w = P.w + (int64_t)c * P.d_pad;
  float* mrow = P.m + (int64_t)c * P.d_pad;
  float* vrow = P.v ? P.v + (int64_t)c * P.d_pad : nullptr;
  float* gw = P.gw + (int64_t)c * P.d_pad;
  float* stash = P.stash + (int64_t)c * kNSlot * kSlotF;
  if (crank == 0) for (int64_t i = tid; i < P.d_pad; i += kThreads) gw[i] = 0.f;
  __syncthreads();
  if (tid == 0) {
    for (int b = 0; b < B_N; ++b) {
      const uint32_t cnt = (b == B_STEP || b == B_A + 2) ? 16u : (b == B_A || b == B_A + 1) ? 8u
                         : (b == B_DONE && CS > 1) ? (uint32_t)(CS - 1) : 1u;
      mbar_init(&bars[b], cnt);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0 && crank == 0) tmem_alloc(tslot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (CS > 1) cluster_sync_all();            // every CTA's barriers exist before a peer arrives on them
  const uint32_t tmem = *tslot;

  const int n = P.n_data[c];
  const int steps_per_epoch = (n + P.batch - 1) / P.batch;

  // This thread's share of one optimiser pass: the merged parameter ranges are dealt out in 128-bit chunks over the
  // element-wise threads of all CTAs of the cluster, two chunks per thread and pass (8 independent loads in flight).
  auto opt_share = [&](const AflOpt& o, int vt, int nvt) {
    for (int r = 0; r < P.n_ranges; ++r) {
      const int64_t lo = P.range_lo[r], hi = P.range_hi[r], span = 4 * (int64_t)nvt;
      for (int64_t i = lo + 4 * (int64_t)vt; i < hi; i += 2 * span) {
        const bool two = i + span < hi;
        afl_reduce_update(w, mrow, vrow, gw, 1, P.d_pad, i, i + 4, o, 0, 1);
        if (two) afl_reduce_update(w, mrow, vrow, gw, 1, P.d_pad, i + span, i + span + 4, o, 0, 1);
      }
    }
  };

  if (crank != 0) {
    // =================================================================================== helper CTA of the cluster
    // The optimiser pass reads g, w, m, v and writes w, m, v of every parameter: on one SM it is bound by that SM's L2
    // bandwidth (27 K of the step's 172 K cycles).  The helpers wait for the main CTA's gradients, update their share
    // of the row and report back; they own no TMEM, no tensor maps and no data.
    if (warp == 0) return;
    const int etid = tid - 32;
    const int skipped = (n % P.batch) == 1 ? 1 : 0;            // a trailing batch of one sample is skipped (client.py:86-87)
    const int total = P.batch == 1 ? 0 : P.epochs * (steps_per_epoch - skipped);
    const uint32_t done0 = map_to_rank(&bars[B_DONE], 0);
    // The merged ranges as one flat list of 128-bit chunks, dealt out over the threads of all helpers: chunk q of thread
    // vt in batch j is q = (4 j' + j) * NT + vt.  A thread's chunks never change, so the optimiser state of its first
    // batch is fetched BEFORE the hand-over (while the main CTA is still in its backward pass) and only the gradient is
    // read after it: one L2 round trip on the critical path instead of one per range.
    const int NT = (CS - 1) * kEpi, vt = ((int)crank - 1) * kEpi + etid;
    int total_chunks = 0;
    for (int r = 0; r < P.n_ranges; ++r) total_chunks += (P.range_hi[r] - P.range_lo[r]) >> 2;
    auto chunk_off = [&](int q) -> int {
      if (q >= total_chunks) return -1;
      for (int r = 0; r < P.n_ranges; ++r) {
        const int nr = (P.range_hi[r] - P.range_lo[r]) >> 2;
        if (q < nr) return P.range_lo[r] + 4 * q;
        q -= nr;
      }
      return -1;
    };
    int off0[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) off0[j] = chunk_off(j * NT + vt);
    const bool adam = P.opt_kind == 0;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    // With one batch per thread (clusters of 8 on this model) the share stays in registers for the whole round: no state
    // loads at all after the first step, and the moments are written back once at the end.
    const bool resident = total_chunks <= 4 * NT;
    double b1t = 1.0, b2t = 1.0;
    float4 W[4], M[4], V[4], G[4];
    for (int k = 0; k < total; ++k) {
      if (k == 0 || !resident) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          W[j] = M[j] = V[j] = z4;
          if (off0[j] >= 0) {
            W[j] = __ldcg(reinterpret_cast<const float4*>(w + off0[j]));
            M[j] = __ldcg(reinterpret_cast<const float4*>(mrow + off0[j]));
            if (adam) V[j] = __ldcg(reinterpret_cast<const float4*>(vrow + off0[j]));
          }
        }
      }
      const bool put_state = !resident || k == total - 1;
      b1t *= (double)P.beta1; b2t *= (double)P.beta2;
      AflOpt o;
      o.kind = P.opt_kind; o.lr = P.lr; o.beta1 = P.beta1; o.beta2 = P.beta2; o.eps = P.eps;
      o.momentum = P.momentum; o.bc1 = (float)(1.0 - b1t); o.bc2 = (float)(1.0 - b2t);
      o.first_step = k == 0;
      const float stepsz = o.lr / o.bc1, isq = 1.0f / sqrtf(o.bc2);
#define HS(i) do { if (P.dbg && c == 0 && crank == 1 && etid == 0 && k == 5) P.dbg[52 + (i)] = clock64(); } while (0)
      HS(0);
      mbar_wait_cluster(&bars[B_GO], (uint32_t)k & 1u);      // acquire: the gradient row (read below with L2 loads)
      HS(1);
      o.clip_coef = P.clip_proper ? *ccoef : 1.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) G[j] = off0[j] >= 0 ? __ldcg(reinterpret_cast<const float4*>(gw + off0[j])) : z4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (off0[j] >= 0) {
          afl_opt_apply4(W[j], M[j], V[j], G[j], o, stepsz, isq);
          __stcg(reinterpret_cast<float4*>(w + off0[j]), W[j]);
          if (put_state) {
            __stcg(reinterpret_cast<float4*>(mrow + off0[j]), M[j]);
            if (adam) __stcg(reinterpret_cast<float4*>(vrow + off0[j]), V[j]);
          }
        }
      }
      for (int base = 4 * NT; base < total_chunks; base += 4 * NT) {        // small clusters: further batches
        int off[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) off[j] = chunk_off(base + j * NT + vt);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          W[j] = M[j] = V[j] = G[j] = z4;
          if (off[j] >= 0) {
            G[j] = __ldcg(reinterpret_cast<const float4*>(gw + off[j]));
            W[j] = __ldcg(reinterpret_cast<const float4*>(w + off[j]));
            M[j] = __ldcg(reinterpret_cast<const float4*>(mrow + off[j]));
            if (adam) V[j] = __ldcg(reinterpret_cast<const float4*>(vrow + off[j]));
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (off[j] >= 0) {
            afl_opt_apply4(W[j], M[j], V[j], G[j], o, stepsz, isq);
            __stcg(reinterpret_cast<float4*>(w + off[j]), W[j]);
            __stcg(reinterpret_cast<float4*>(mrow + off[j]), M[j]);
            if (adam) __stcg(reinterpret_cast<float4*>(vrow + off[j]), V[j]);
          }
        }
      }
      HS(2);
      bar_named(13, kEpi);
      HS(3);
      if (etid == 0) {
        // one cumulative release for the whole CTA (the barrier ordered every thread's stores before this fence); the
        // proxy fence covers the main CTA's TMA re-read of the weights this CTA just wrote
        __threadfence();
        fence_async();
        remote_arrive(done0);
        HS(4);
      }
    }
    return;
  }

  if (warp == 0) {
    // =================================================================================== producer warp
    // (warp 0: the oldest warp of its scheduler, so the serial TMA / MMA issue stream is not starved by the math warps)
    {
      uint32_t ph = 0;                          // parity toggles of the barriers this thread waits on every phase
      uint32_t n_acc[3] = {0, 0, 0}, n_dw[3] = {0, 0, 0};
      auto wait = [&](int b) { mbar_wait_to(&bars[b], (ph >> b) & 1u); ph ^= 1u << b; };
      auto wait_a = [&](int s) { wait(B_A + s); tc_fence_after(); };
      auto commit_acc = [&](int s) { if (elect_one()) umma_commit(&bars[B_ACC + s]); ++n_acc[s]; };
      auto commit_dw = [&](int s) { if (elect_one()) umma_commit(&bars[B_DW + s]); ++n_dw[s]; };
      auto wait_acc_last = [&](int s) { mbar_wait_to(&bars[B_ACC + s], (n_acc[s] - 1) & 1u); };
      auto wait_dw_last = [&](int s) { mbar_wait_to(&bars[B_DW + s], (n_dw[s] - 1) & 1u); };
      auto tma = [&](int map, uint32_t dst, int c0, int c1, int bar) { if (elect_one()) tma_load_3d(dst, &maps.m[map], c0, c1, c, &bars[bar]); };
      auto expect = [&](int bar, uint32_t bytes) { if (elect_one()) mbar_expect_tx(&bars[bar], bytes); };
      auto stash_ld = [&](int slot, int nslots, uint32_t dst, int bar) {
        if (elect_one()) {
          mbar_expect_tx(&bars[bar], (uint32_t)nslots * kSlotB);
          for (int j = 0; j < nslots; ++j) tma_load_3d(dst + j * kSlotB, &maps.m[TM_STASH], 0, 0, c * kNSlot + slot + j, &bars[bar]);
        }
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
            expect(B_WF_V + s, 16384);
            tma(TM_WV_K + s, wb[s][0], 0, 0, B_WF_V + s);
            tma(TM_WV_K + s, wb[s][0] + 8192, 32, 0, B_WF_V + s);
            expect(B_WF_O + s, 16384);
            tma(TM_WO_K + s, wb[s][1], 0, 0, B_WF_O + s);
            tma(TM_WO_K + s, wb[s][1] + 8192, 32, 0, B_WF_O + s);
          }
          expect(B_WF_W2, 2 * 4096 + 4 * 2048);
          for (int kb = 0; kb < 2; ++kb) tma(TM_W2_K, w2f + kb * 4096, kb * 32, 0, B_WF_W2);
          for (int s = 0; s < 2; ++s)
            for (int kb = 0; kb < 2; ++kb) tma(TM_WF0_K + s, wf0f[s] + kb * 2048, kb * 32, 0, B_WF_W2);
          expect(B_WF_W1, 32768);
          for (int kb = 0; kb < 4; ++kb) tma(TM_W1_K, xb[0] + kb * 8192, kb * 32, 0, B_WF_W1);

          // ---- forward
          for (int s = 0; s < 2; ++s) {   // P1: dense
            wait_a(s);
            if (s) job_knosw<64, 16>(accc[s], acol[s], smw + kSmWd1); else job_knosw<64, 8>(accc[s], acol[s], smw + kSmWd0);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P2: V projection
            wait_a(s);
            wait(B_WF_V + s);
            job_ksw<64, 64>(accc[s], acol[s], wb[s][0], 8192);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P3: out_proj
            const bool dbgp = P.dbg && c == 0 && e == 0 && st == 5 && s == 0;
            if (dbgp && lane == 0) P.dbg[48] = clock64();
            wait_a(s);
            if (dbgp && lane == 0) P.dbg[49] = clock64();
            wait(B_WF_O + s);
            if (dbgp && lane == 0) P.dbg[50] = clock64();
            job_ksw<64, 64>(accc[s], acol[s], wb[s][1], 8192);
            commit_acc(s);
            if (dbgp && lane == 0) P.dbg[51] = clock64();
          }
          wait(B_WF_W2);
          for (int s = 0; s < 2; ++s) {   // P4: ffn.0 (N padded to 16)
            wait_a(s);
            job_ksw<16, 64>(accc[s], acol[s], wf0f[s], 2048);
            commit_acc(s);
          }
          for (int s = 0; s < 2; ++s) {   // P5: ffn.3 (K padded to 8)
            wait_a(s);
            job_knosw<64, 8>(accc[s], acol[s], smw + (s ? kSmWf31 : kSmWf30));
            commit_acc(s);
          }
          wait_a(0); wait_a(1);           // P6: fc1 on cat
          wait(B_WF_W1);
          job_ksw<64, 128>(accc[2], acol[2], xb[0], 8192);
          commit_acc(2);
          wait_a(2);                      // P7: fc2
          job_ksw<32, 64>(accc[2], acol[2], w2f, 4096);
          commit_acc(2);
          wait_a(2);                      // P8: output (N padded to 16)
          job_knosw<16, 32>(accc[2], acol[2], smw + kSmW3);
          commit_acc(2);

          // ---- backward, head
          wait_acc_last(2);               // every forward MMA has finished: weight / x buffers are free
          expect(B_WB + 0, 8192);
          for (int nb = 0; nb < 2; ++nb) tma(TM_W2_M, wb[0][0] + nb * 4096, nb * 32, 0, B_WB + 0);
          stash_ld(SL_HD, 2, xb[0], B_X + 0);
          expect(B_WB + 1, 32768);
          for (int nb = 0; nb < 4; ++nb) tma(TM_W1_M, wb[0][1] + nb * 8192, nb * 32, 0, B_WB + 1);
          wait_a(2);                      // B7: da2 in A[0..31] and dymn[V]
          wait(B_WB + 0);
          job_mn<64, 32>(accc[2], acol[2], wb[0][0], 4096);
          commit_acc(2);
          wait(B_X + 0);
          job_dw<64>(tmem + kColDw, dy[0], xb[0], kSlotB);                    // dW2 -> V slot 0
          commit_dw(2);
          wait_dw_last(2);
          stash_ld(SL_CAT, 4, xb[0], B_X + 1);
          wait_a(2);                      // B6: da in A[0..63] and dymn[L]
          wait(B_WB + 1);
          job_mn<128, 64>(accc[2], acol[2], wb[0][1], 8192);                  // dcat -> acc 128..255
          commit_acc(2);
          wait(B_X + 1);
          job_dw<128>(tmem + kColDw + 128, dy[1], xb[0], kSlotB);             // dW1 -> L slots (128 columns)
          commit_dw(2);
          // ---- backward, branches.  Weight double buffer: wb[s][0] / wb[s][1]; X single buffer xb[s]
          wait_acc_last(2);               // dX6 done: W1mn (wb[0][1], wb[1][0]) is free
          wait_dw_last(2);                // dW1 done: xbuf is free
          for (int s = 0; s < 2; ++s) {
            expect(B_WB + 2 * s, 2048);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WF0_M + s, wb[s][0] + nb * 1024, nb * 32, 0, B_WB + 2 * s);
            stash_ld(SL_BR * s + SL_GF, 1, xb[s], B_X + s);
          }
          dwk[0] = 1; dwk[1] = 0;         // dW2 sits in V slot 0 until its drain; dW1 is drained before b_a[L]
          for (int s = 0; s < 2; ++s) {   // B5: ffn.3 (Wf3 MN-major written by the element-wise warps into wb[s][1])
            wait_a(s);
            job_mn<16, 64>(accc[s], acol[s], wb[s][1], 0);
            commit_acc(s);
            wait(B_X + s);
            job_dw<16>(dwcol(s), dy[s], xb[s], 0);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) {
            wait_dw_last(s);
            stash_ld(SL_BR * s + SL_N1, 2, xb[s], B_X + s);
            expect(B_WB + 2 * s + 1, 16384);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WO_M + s, wb[s][1] + nb * 8192, nb * 32, 0, B_WB + 2 * s + 1);
          }
          for (int s = 0; s < 2; ++s) {   // B4: ffn.0
            wait_a(s);
            wait(B_WB + 2 * s);
            job_mn<64, 8>(accc[s], acol[s], wb[s][0], 1024);
            commit_acc(s);
            wait(B_X + s);
            job_dw<64>(dwcol(s), dy[s], xb[s], kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) {
            wait_dw_last(s);
            stash_ld(SL_BR * s + SL_VD, 2, xb[s], B_X + s);
            expect(B_WB + 2 * s, 16384);
            for (int nb = 0; nb < 2; ++nb) tma(TM_WV_M + s, wb[s][0] + nb * 8192, nb * 32, 0, B_WB + 2 * s);
          }
          for (int s = 0; s < 2; ++s) {   // B3: out_proj
            wait_a(s);
            wait(B_WB + 2 * s + 1);
            job_mn<64, 64>(accc[s], acol[s], wb[s][1], 8192);
            commit_acc(s);
            wait(B_X + s);
            job_dw<64>(dwcol(s), dy[s], xb[s], kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) { wait_dw_last(s); stash_ld(SL_BR * s + SL_H1, 2, xb[s], B_X + s); }
          for (int s = 0; s < 2; ++s) {   // B2: V projection
            wait_a(s);
            wait(B_WB + 2 * s);
            job_mn<64, 64>(accc[s], acol[s], wb[s][0], 8192);
            commit_acc(s);
            wait(B_X + s);
            job_dw<64>(dwcol(s), dy[s], xb[s], kSlotB);
            commit_dw(s);
            dwk[s] ^= 1;
          }
          for (int s = 0; s < 2; ++s) { wait_dw_last(s); stash_ld(SL_BR * s + SL_X, 1, xb[s], B_X + s); }
          for (int s = 0; s < 2; ++s) {   // B1: dense (dW only)
            wait_a(s);
            wait(B_X + s);
            job_dw<16>(dwcol(s), dy[s], xb[s], 0);
            commit_dw(s);
            dwk[s] ^= 1;
          }
        }
      }
    }
  } else {
    // =================================================================================== element-wise warps
    // Branch phases: warps 0-7 own the vitals stream and warps 8-15 the labs stream, so the two branches run concurrently
    // and each hides the other's latencies; a thread owns one sample (TMEM lane) and 32 of the 64 features, processed as
    // two 16-column halves.  Head phases: all 16 warps, 16 columns per thread.
    // warps 1..16; the TMEM lane quarter of a warp is fixed by the hardware (warp id % 4), its other role bits are free
    const int etid = tid - 32, ew = warp - 1;
    const int q = warp & 3, row = 32 * q + lane;
    const int s = ew >> 3, ch2 = (ew >> 2) & 1;               // branch phases: stream and 32-column half
    const int chunk = ew >> 2, col0 = 16 * chunk;             // head phases
    const int SB = SP_BR * s, S0 = SL_BR * s;
    const uint32_t tq = tmem + ((uint32_t)(32 * q) << 16);
    uint8_t* dymn[2] = {smem + kOffDy, smem + kOffDy + 2 * kSlotB};
    uint32_t ph = 0;
    auto wait_b = [&](int b) { mbar_wait_to(&bars[b], (ph >> b) & 1u); ph ^= 1u << b; };
    auto wait_acc = [&](int st) { wait_b(B_ACC + st); tc_fence_after(); };
    auto wait_dw = [&](int st) { wait_b(B_DW + st); tc_fence_after(); };
    // forward: the next MMA reads its A operand from TMEM only, no proxy fence needed
    auto signal_fwd = [&](int st) { tmem_st_wait(); tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(&bars[B_A + st]); };
    // backward: dY was also written to shared memory for the dW MMA (generic -> async proxy)
    auto signal_bwd = [&](int st) {
      tmem_st_wait();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[B_A + st]);
    };
    auto slot = [&](int sl) { return stash + (int64_t)sl * kSlotF + row * 32; };
    auto blk = [&](int sl) { return stash + (int64_t)sl * kSlotF; };
    // thread-private stash tensors, lane-interleaved (this thread's 128-bit chunk j of tensor `sl`)
    auto el32 = [&](int sl, int j) { return reinterpret_cast<float4*>(stash + (int64_t)sl * kSlotF) + (ew * 8 + j) * 32 + lane; };
    auto el16 = [&](int sl, int j) { return reinterpret_cast<float4*>(stash + (int64_t)sl * kSlotF) + (ew * 4 + j) * 32 + lane; };
    auto el_st32 = [&](int sl, const float* x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) __stcg(el32(sl, j), make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]));
    };
    auto el_ld32 = [&](int sl, float* x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float4 t = __ldcg(el32(sl, j)); x[4 * j] = t.x; x[4 * j + 1] = t.y; x[4 * j + 2] = t.z; x[4 * j + 3] = t.w; }
    };
    auto el_st16 = [&](int sl, const float* x) {
#pragma unroll
      for (int j = 0; j < 4; ++j) __stcg(el16(sl, j), make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]));
    };
    auto el_ld16 = [&](int sl, float* x) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float4 t = __ldcg(el16(sl, j)); x[4 * j] = t.x; x[4 * j + 1] = t.y; x[4 * j + 2] = t.z; x[4 * j + 3] = t.w; }
    };
    // forward-only TMEM scratch in the (then idle) dW accumulator columns: the residual inputs h1 and n1 of this branch
    const uint32_t tH1 = tq + kColDw + 64 * s + 32 * ch2, tN1 = tq + kColDw + 128 + 64 * s + 32 * ch2;
    const int row0 = 32 * q;
    auto ld_acc16 = [&](uint32_t col, float* v) { tmem_ld16_async(tq + col, v); tmem_ld_wait(); };
    int lnk = 0;
    // LayerNorm partial sums of this thread's row: exchanged between the two warps that share the row
    auto ln_stats2 = [&](float a, float b, float& sa, float& sb) {
      float2* x = lnx + ((s * 2 + lnk) * 128 + row) * 2;
      lnk ^= 1;
      x[ch2] = make_float2(a, b);
      bar_named(1 + q + 4 * s, 64);
      sa = x[0].x + x[1].x;
      sb = x[0].y + x[1].y;
    };
    // drain one 16-column piece of a dW tile (M = 64: row r lives in lane (r / 16) * 32 + r % 16) into the gradient row
    auto drain16 = [&](uint32_t col, float* dst_row0, int rows, int ld, int c0, int ncols) {
      float g[16];
      ld_acc16(col + c0, g);
      const int r = 16 * q + lane;
      if (lane < 16 && r < rows && c0 < ncols) {
        float* dst = dst_row0 + (int64_t)r * ld + c0;
        if ((ld & 3) == 0 && c0 + 16 <= ncols) st16g(dst, g);
        else for (int j = 0; j < 16; ++j) if (c0 + j < ncols) __stcg(dst + j, g[j]);
      }
    };

    int tsi = 0;
#define TS2(k) do { if (P.dbg && c == 0 && etid == 0 && gstep == 5) P.dbg[32 + (k)] = clock64(); } while (0)
#define TSTAMP() do { if (P.dbg && c == 0 && etid == 0 && gstep == 5 && tsi < 64) P.dbg[tsi++] = clock64(); } while (0)
    Rng rng; rng.seed_lo = P.seed_lo; rng.seed_hi = P.seed_hi; rng.client = (uint32_t)P.client_ids[c]; rng.on = P.dropout;
    const float inv1 = 1.f / (1.f - P.p1), inv3 = 1.f / (1.f - P.p3);
    double b1t = 1.0, b2t = 1.0;
    int opt_steps = 0, gstep = 0;
    int dwk = 0;
    float rstd_r[3];                          // LayerNorm 1/std of this thread's row (own branch), kept for the backward
    uint32_t mk1 = 0, mk2 = 0, mk3 = 0, mkh = 0, mkf = 0;   // dropout keep bits drawn in the forward, reused by the backward
    auto dwcol = [&]() { return kColDw + (uint32_t)s * 128u + (uint32_t)dwk * 64u; };
    // keep bits of 32 consecutive elements of a site (bit j = element e0 + j)
    auto keep32 = [&](int site, uint32_t e0, float p) -> uint32_t {
      if (!(rng.on && p > 0.f)) return 0xffffffffu;
      uint32_t bits = 0;
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const AflU4 u = afl_philox((e0 + j) >> 2, ((uint32_t)site & 0xFFFFu) | (rng.client << 16), rng.step, (uint32_t)row, rng.seed_lo, rng.seed_hi);
        const float sc = 1.0f / 16777216.0f;
        bits |= (((float)(u.x >> 8) * sc >= p) ? 1u : 0u) << j;
        bits |= (((float)(u.y >> 8) * sc >= p) ? 1u : 0u) << (j + 1);
        bits |= (((float)(u.z >> 8) * sc >= p) ? 1u : 0u) << (j + 2);
        bits |= (((float)(u.w >> 8) * sc >= p) ? 1u : 0u) << (j + 3);
      }
      return bits;
    };
    auto keep16 = [&](int site, uint32_t e0, float p) -> uint32_t {
      if (!(rng.on && p > 0.f)) return 0xffffu;
      uint32_t bits = 0;
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        const AflU4 u = afl_philox((e0 + j) >> 2, ((uint32_t)site & 0xFFFFu) | (rng.client << 16), rng.step, (uint32_t)row, rng.seed_lo, rng.seed_hi);
        const float sc = 1.0f / 16777216.0f;
        bits |= (((float)(u.x >> 8) * sc >= p) ? 1u : 0u) << j;
        bits |= (((float)(u.y >> 8) * sc >= p) ? 1u : 0u) << (j + 1);
        bits |= (((float)(u.z >> 8) * sc >= p) ? 1u : 0u) << (j + 2);
        bits |= (((float)(u.w >> 8) * sc >= p) ? 1u : 0u) << (j + 3);
      }
      return bits;
    };

    // small weights + parameters -> shared memory (start of every step: after the optimiser changed them).  All global
    // loads are issued before the first store so that one L2 round trip covers them.
    auto sp_src = [&](int i) -> int {          // parameter-row offset of small-parameter slot i (-1: padding)
      if (i < 2 * SP_BR) {
        const int b = i >= SP_BR, r = i - b * SP_BR;
        if (r < SP_BF0) {
          const int seg = r >> 6, j = r & 63;
          const int o = seg == 0 ? L.bd[b] : seg == 1 ? L.bv[b] : seg == 2 ? L.bo[b] : seg == 3 ? L.g1[b] : L.be1[b];
          return o + j;
        }
        if (r < SP_BF3) return (r - SP_BF0) < L.ff ? L.bf0[b] + (r - SP_BF0) : -1;
        const int r2 = r - SP_BF3, seg = r2 >> 6, j = r2 & 63;
        const int o = seg == 0 ? L.bf3[b] : seg == 1 ? L.g2[b] : seg == 2 ? L.be2[b] : seg == 3 ? L.g3[b] : L.be3[b];
        return o + j;
      }
      const int r = i - 2 * SP_BR;
      if (r < 64) return L.b1 + r;
      if (r < 96) return L.b2 + (r - 64);
      if (r < 128) return L.w3 + (r - 96);
      return r == 128 ? L.b3 : -1;
    };
    auto load_small = [&]() {
      float vsp[3], vwd0, vwd1[2], vf3[2], vw3;
      int isp[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int i = etid + k * kEpi;
        isp[k] = i < SP_N ? sp_src(i) : -2;
        vsp[k] = isp[k] >= 0 ? __ldcg(w + isp[k]) : 0.f;
      }
      {  // dense V: [64 x 8]
        const int nn = etid >> 3, k = etid & 7;
        vwd0 = k < L.kin[0] ? __ldcg(w + L.wd[0] + nn * L.kin[0] + k) : 0.f;
      }
#pragma unroll
      for (int t = 0; t < 2; ++t) {  // dense L: [64 x 16]
        const int i = etid + t * kEpi, nn = i >> 4, k = i & 15;
        vwd1[t] = k < L.kin[1] ? __ldcg(w + L.wd[1] + nn * L.kin[1] + k) : 0.f;
      }
#pragma unroll
      for (int b = 0; b < 2; ++b) {  // ffn.3: [64 x 8]
        const int nn = etid >> 3, k = etid & 7;
        vf3[b] = k < L.ff ? __ldcg(w + L.wf3[b] + nn * L.ff + k) : 0.f;
      }
      {  // output: [16 x 32], row 0 only
        const int nn = etid >> 5, k = etid & 31;
        vw3 = nn == 0 ? __ldcg(w + L.w3 + k) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int i = etid + k * kEpi;
        if (i < SP_N) { sp[i] = vsp[k]; sg[i] = 0.f; }
      }
      uint8_t* sm = smem + kOffSm;
      {
        const int nn = etid >> 3, k = etid & 7;
        *reinterpret_cast<float*>(sm + kSmWd0 + (nn >> 3) * 256 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = vwd0;
        *reinterpret_cast<float*>(sm + kSmWf30 + (nn >> 3) * 256 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = vf3[0];
        *reinterpret_cast<float*>(sm + kSmWf31 + (nn >> 3) * 256 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = vf3[1];
      }
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const int i = etid + t * kEpi, nn = i >> 4, k = i & 15;
        *reinterpret_cast<float*>(sm + kSmWd1 + (nn >> 3) * 512 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = vwd1[t];
      }
      {
        const int nn = etid >> 5, k = etid & 31;
        *reinterpret_cast<float*>(sm + kSmW3 + (nn >> 3) * 1024 + (k >> 2) * 128 + (nn & 7) * 16 + (k & 3) * 4) = vw3;
      }
    };

    load_small();
    fence_async();
    bar_named(13, kEpi);
    if (lane == 0) mbar_arrive(&bars[B_STEP]);

    for (int e = 0; e < P.epochs; ++e) {
      const int32_t* idx_e = P.idx + (int64_t)c * P.idx_stride_c + (int64_t)e * P.idx_stride_e;
      for (int st = 0; st < steps_per_epoch; ++st, ++gstep) {
        const int b = min(P.batch, n - st * P.batch);
        if (b == 1) continue;                              // client.py:86-87
        rng.step = (uint32_t)gstep;
        const float inv_b = 1.0f / (float)b;
        const bool valid = row < b;
        const uint32_t accS = kColAcc + 64 * s, aS = kColA + 64 * s;     // this branch's accumulator / A-operand columns
        TSTAMP();
        // ---------------------------------------------------------------- E0: gather the batch
        if (ch2 == 0) {
          float x[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) x[j] = 0.f;
          float y = 0.f;
          if (valid) {
            const float* rec = P.data + (int64_t)idx_e[(int64_t)st * P.batch + row] * P.rec;
            if (s == 0) { for (int j = 0; j < L.kin[0]; ++j) x[j] = rec[j]; y = rec[L.kin[0] + L.kin[1]]; }
            else for (int j = 0; j < L.kin[1]; ++j) x[j] = rec[L.kin[0] + j];
          }
          if (s == 0) { ylab[row] = y; tmem_st8(tq + aS, x); } else tmem_st16(tq + aS, x);
          st16t(blk(S0 + SL_X), row0, lane, 0, x);
        }
        signal_fwd(s);
        TSTAMP();
        // ---------------------------------------------------------------- forward, this warp group's branch
        {                               // E1: h1 = gelu(x Wd^T + bd)
          wait_acc(s);
          float a[32], g[32];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            ld_acc16(accS + c0, a + 16 * hf);
#pragma unroll
            for (int j = 0; j < 16; ++j) gelu_both(a[16 * hf + j] + sp[SB + SP_BD + c0 + j], a[16 * hf + j], g[16 * hf + j]);
            tmem_st16(tq + aS + c0, a + 16 * hf);
          }
          signal_fwd(s);
          tmem_st16(tH1, a);
          tmem_st16(tH1 + 16, a + 16);
          el_st32(EL_GP1, g);
          st32t(blk(S0 + SL_H1 + ch2), row0, lane, a);
        }
        TSTAMP();
        {                               // E2: vd = headdrop(h1 Wv^T + bv)
          mkh = 0;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf)
            mkh |= (drop1(rng, L.rng_head[s], row, 2 * ch2 + hf, P.p1) != 0.f ? 1u : 0u) << hf;
          wait_acc(s);
          float a[32];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            const float ds = !rng.on ? 1.f : ((mkh >> hf) & 1u) ? inv1 : 0.f;
            ld_acc16(accS + c0, a + 16 * hf);
#pragma unroll
            for (int j = 0; j < 16; ++j) a[16 * hf + j] = (a[16 * hf + j] + sp[SB + SP_BV + c0 + j]) * ds;
            tmem_st16(tq + aS + c0, a + 16 * hf);
          }
          signal_fwd(s);
          st32t(blk(S0 + SL_VD + ch2), row0, lane, a);
        }
        TSTAMP();
        {                               // E3: n1 = LN1(h1 + drop(vd Wo^T + bo))
          TS2(0);
          float r[32];
          tmem_ld16_async(tH1, r);
          tmem_ld16_async(tH1 + 16, r + 16);
          tmem_ld_wait();
          mk1 = keep32(L.rng_ln1[s], 32 * ch2, P.p1);
          TS2(1);
          wait_acc(s);
          TS2(2);
          float sum = 0.f, sq = 0.f;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            float a[16];
            ld_acc16(accS + c0, a);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float ds = !rng.on ? 1.f : ((mk1 >> (16 * hf + j)) & 1u) ? inv1 : 0.f;
              const float x = r[16 * hf + j] + (a[j] + sp[SB + SP_BO + c0 + j]) * ds;
              r[16 * hf + j] = x; sum += x; sq += x * x;
            }
          }
          TS2(3);
          float ts, tq2;
          ln_stats2(sum, sq, ts, tq2);
          TS2(4);
          const float mean = ts * (1.f / 64.f);
          float var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          const float rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
          rstd_r[0] = rstd;
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = (r[j] - mean) * rstd;                        // xhat
          {
            float n1[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) n1[j] = r[j] * sp[SB + SP_G1 + 32 * ch2 + j] + sp[SB + SP_BE1 + 32 * ch2 + j];
            tmem_st16(tq + aS + 32 * ch2, n1);
            tmem_st16(tq + aS + 32 * ch2 + 16, n1 + 16);
            TS2(5);
            signal_fwd(s);
            TS2(6);
            tmem_st16(tN1, n1);
            tmem_st16(tN1 + 16, n1 + 16);
            st32t(blk(S0 + SL_N1 + ch2), row0, lane, n1);
          }
          el_st32(EL_XH1, r);
        }
        TSTAMP();
        {                               // E4: gf = drop(gelu(n1 Wf0^T + bf0))   (ff <= 8 columns)
          wait_acc(s);
          if (ch2 == 0) {
            float a[16], g[16];
            ld_acc16(accS, a);
            mkf = keep16(L.rng_f1[s], 0, P.p1) & 0xffu;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (j < 8 && j < L.ff) {
                const float ds = !rng.on ? 1.f : ((mkf >> j) & 1u) ? inv1 : 0.f;
                float gl, gp;
                gelu_both(a[j] + sp[SB + SP_BF0 + j], gl, gp);
                g[j] = gp * ds; a[j] = gl * ds;
              } else { g[j] = 0.f; a[j] = 0.f; }
            }
            tmem_st8(tq + aS, a);
            el_st16(EL_GPF, g);
            st16t(blk(S0 + SL_GF), row0, lane, 0, a);
          }
          signal_fwd(s);
        }
        TSTAMP();
        {                               // E5: n3 = LN3(LN2(n1 + drop(gf Wf3^T + bf3)))
          float r[32];
          tmem_ld16_async(tN1, r);
          tmem_ld16_async(tN1 + 16, r + 16);
          tmem_ld_wait();
          mk2 = keep32(L.rng_ln2[s], 32 * ch2, P.p1);
          wait_acc(s);
          float sum = 0.f, sq = 0.f;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            float a[16];
            ld_acc16(accS + c0, a);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float ds = !rng.on ? 1.f : ((mk2 >> (16 * hf + j)) & 1u) ? inv1 : 0.f;
              const float x = r[16 * hf + j] + (a[j] + sp[SB + SP_BF3 + c0 + j]) * ds;
              r[16 * hf + j] = x; sum += x; sq += x * x;
            }
          }
          float ts, tq2;
          ln_stats2(sum, sq, ts, tq2);
          float mean = ts * (1.f / 64.f);
          float var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          float rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
          rstd_r[1] = rstd;
          sum = 0.f; sq = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = (r[j] - mean) * rstd;                        // xhat2
          el_st32(EL_XH2, r);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float x = r[j] * sp[SB + SP_G2 + 32 * ch2 + j] + sp[SB + SP_BE2 + 32 * ch2 + j];
            r[j] = x; sum += x; sq += x * x;
          }
          ln_stats2(sum, sq, ts, tq2);
          mean = ts * (1.f / 64.f);
          var = tq2 * (1.f / 64.f) - mean * mean;
          var = var > 0.f ? var : 0.f;
          rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
          rstd_r[2] = rstd;
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = (r[j] - mean) * rstd;                        // xhat3
          {
            float n3[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) n3[j] = r[j] * sp[SB + SP_G3 + 32 * ch2 + j] + sp[SB + SP_BE3 + 32 * ch2 + j];
            tmem_st16(tq + aS + 32 * ch2, n3);
            tmem_st16(tq + aS + 32 * ch2 + 16, n3 + 16);
            signal_fwd(s);
            st32t(blk(SL_CAT + 2 * s + ch2), row0, lane, n3);
          }
          el_st32(EL_XH3, r);
        }
        TSTAMP();
        // ---------------------------------------------------------------- forward, head (all 16 warps, 16 columns each)
        {                               // E6: hd = drop(gelu(cat W1^T + b1))
          mk3 = keep16(L.rng_fc1, col0, P.p3);
          wait_acc(2);
          float a[16], g[16];
          ld_acc16(kColAcc + col0, a);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float ds = !rng.on ? 1.f : ((mk3 >> j) & 1u) ? inv3 : 0.f;
            float gl, gp;
            gelu_both(a[j] + sp[SP_B1 + col0 + j], gl, gp);
            g[j] = gp * ds; a[j] = gl * ds;
          }
          tmem_st16(tq + kColA + col0, a);
          signal_fwd(2);
          el_st16(EL_GPA, g);
          st16t(blk(SL_HD + (col0 >> 5)), row0, lane, col0 & 31, a);
        }
        TSTAMP();
        float h2[16], gp2[16];
        {                               // E7: h2 = gelu(hd W2^T + b2)   (32 columns: chunks 0, 1)
          wait_acc(2);
          if (chunk < 2) {
            ld_acc16(kColAcc + col0, h2);
#pragma unroll
            for (int j = 0; j < 16; ++j) gelu_both(h2[j] + sp[SP_B2 + col0 + j], h2[j], gp2[j]);
            tmem_st16(tq + kColA + col0, h2);
            st16t(blk(SL_H2), row0, lane, col0, h2);
          }
          signal_fwd(2);
        }
        TSTAMP();
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
          bar_named(9 + q, 128);
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
          // every stash store of the forward is fenced here, once, for the TMA reads of the backward
          tmem_st_wait();
          fence_async();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[B_A + 2]);
        }
        TSTAMP();
        // ---------------------------------------------------------------- backward, head
        {                               // B7: da = (da2 W2) gelu'(a) mask
          float g[16];
          el_ld16(EL_GPA, g);
          wait_acc(2);
          float a[16];
          ld_acc16(kColAcc + col0, a);
#pragma unroll
          for (int j = 0; j < 16; ++j) a[j] *= g[j];
          colsum16(a, sg + SP_B1 + col0, lane);
          tmem_st16(tq + kColA + col0, a);
          st_dymn(dymn[1], row, col0, a);
          wait_dw(2);                                         // dW2 (reads dymn[V]) finished before it may be lapped
          signal_bwd(2);
          drain16(kColDw, gw + L.w2, 32, 64, col0, 64);
        }
        TSTAMP();
        {                               // B6: dcat -> LN3, LN2 backward (own branch) -> do2
          float g[32], xh[32];
          el_ld32(EL_XH3, xh);
          wait_acc(2);
          ld_acc16(accS + 32 * ch2, g);
          ld_acc16(accS + 32 * ch2 + 16, g + 16);
          float sa = 0.f, sb = 0.f;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            float t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = g[16 * hf + j] * xh[16 * hf + j];
            colsum16(t, sg + SB + SP_G3 + c0, lane);
            colsum16(g + 16 * hf, sg + SB + SP_BE3 + c0, lane);
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[16 * hf + j] *= sp[SB + SP_G3 + c0 + j]; sa += g[16 * hf + j]; sb += g[16 * hf + j] * xh[16 * hf + j]; }
          }
          float ta, tb;
          ln_stats2(sa, sb, ta, tb);
          float rs = rstd_r[2], m1 = ta * (1.f / 64.f), m2 = tb * (1.f / 64.f);
#pragma unroll
          for (int j = 0; j < 32; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);             // dL/dn2
          el_ld32(EL_XH2, xh);
          sa = 0.f; sb = 0.f;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            float t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = g[16 * hf + j] * xh[16 * hf + j];
            colsum16(t, sg + SB + SP_G2 + c0, lane);
            colsum16(g + 16 * hf, sg + SB + SP_BE2 + c0, lane);
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[16 * hf + j] *= sp[SB + SP_G2 + c0 + j]; sa += g[16 * hf + j]; sb += g[16 * hf + j] * xh[16 * hf + j]; }
          }
          ln_stats2(sa, sb, ta, tb);
          rs = rstd_r[1]; m1 = ta * (1.f / 64.f); m2 = tb * (1.f / 64.f);
#pragma unroll
          for (int j = 0; j < 32; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);             // dL/dr2
          el_st32(EL_RES, g);
#pragma unroll
          for (int j = 0; j < 32; ++j) g[j] *= !rng.on ? 1.f : ((mk2 >> j) & 1u) ? inv1 : 0.f;   // dL/do2
          colsum16(g, sg + SB + SP_BF3 + 32 * ch2, lane);
          colsum16(g + 16, sg + SB + SP_BF3 + 32 * ch2 + 16, lane);
          {   // ffn.3 weights, MN-major, for dgf = do2 . Wf3 (N = in padded to 16, K = 64 outputs)
            uint8_t* wmn = smem + kOffW + (2 * s + 1) * kSlotB;
            for (int i = etid - 256 * s; i < 64 * 16; i += 256) {
              const int o = i >> 4, ii = i & 15;
              *reinterpret_cast<float*>(wmn + o * 128 + (((ii >> 3) ^ (o & 3)) << 5) + (ii & 7) * 4) =
                  ii < L.ff ? __ldcg(w + L.wf3[s] + o * L.ff + ii) : 0.f;
            }
          }
          if (s == 1) wait_dw(2);                             // dW1 (reads dymn[L]) finished
          tmem_st16(tq + aS + 32 * ch2, g);
          tmem_st16(tq + aS + 32 * ch2 + 16, g + 16);
          st_dymn(dymn[s], row, 32 * ch2, g);
          st_dymn(dymn[s], row, 32 * ch2 + 16, g + 16);
          if (s == 1) {                                       // dW1: 64 x 128 tile, 64 columns per warp
#pragma unroll
            for (int k = 0; k < 4; ++k) drain16(kColDw + 128, gw + L.w1, 64, 128, 64 * ch2 + 16 * k, 128);
          }
          signal_bwd(s);
          if (s == 0) wait_dw(2);                             // keep the parity of the head's dW barrier in step
          dwk = s == 0 ? 1 : 0;
        }
        TSTAMP();
        // ---------------------------------------------------------------- backward, this warp group's branch
        {                               // B5: df = (do2 Wf3) gelu'(f) mask
          wait_acc(s);
          float a[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) a[j] = 0.f;
          if (ch2 == 0) {
            float g[16];
            el_ld16(EL_GPF, g);
            ld_acc16(accS, a);
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = (j < L.ff) ? a[j] * g[j] : 0.f;
            colsum16(a, sg + SB + SP_BF0, lane);
          }
          wait_dw(s);                                           // dWf3 (reads dymn[s] = do2) finished
          if (ch2 == 0) {
            tmem_st8(tq + aS, a);
            st_dymn(dymn[s], row, 0, a);
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = 0.f;
            st_dymn(dymn[s], row, 16, a);
          }
          signal_bwd(s);
          {                                                     // drain dWf3: [64 out] x [ff in]
            float g[16];
            ld_acc16(dwcol(), g);
            const int r = 16 * q + lane;
            if (ch2 == 0 && lane < 16) for (int j = 0; j < L.ff; ++j) __stcg(gw + L.wf3[s] + r * L.ff + j, g[j]);
          }
          dwk ^= 1;
        }
        TSTAMP();
        {                               // B4: dr1 = LN1'(df Wf0 + dr2); do = dr1 mask
          float g[32], xh[32];
          el_ld32(EL_RES, g);
          el_ld32(EL_XH1, xh);
          wait_acc(s);
          float sa = 0.f, sb = 0.f;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int c0 = 32 * ch2 + 16 * hf;
            float t[16];
            ld_acc16(accS + c0, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[16 * hf + j] += t[j]; t[j] = g[16 * hf + j] * xh[16 * hf + j]; }
            colsum16(t, sg + SB + SP_G1 + c0, lane);
            colsum16(g + 16 * hf, sg + SB + SP_BE1 + c0, lane);
#pragma unroll
            for (int j = 0; j < 16; ++j) { g[16 * hf + j] *= sp[SB + SP_G1 + c0 + j]; sa += g[16 * hf + j]; sb += g[16 * hf + j] * xh[16 * hf + j]; }
          }
          float ta, tb;
          ln_stats2(sa, sb, ta, tb);
          const float rs = rstd_r[0], m1 = ta * (1.f / 64.f), m2 = tb * (1.f / 64.f);
#pragma unroll
          for (int j = 0; j < 32; ++j) g[j] = rs * (g[j] - m1 - xh[j] * m2);             // dL/dr1
          el_st32(EL_RES, g);
#pragma unroll
          for (int j = 0; j < 32; ++j) g[j] *= !rng.on ? 1.f : ((mk1 >> j) & 1u) ? inv1 : 0.f;   // dL/do
          colsum16(g, sg + SB + SP_BO + 32 * ch2, lane);
          colsum16(g + 16, sg + SB + SP_BO + 32 * ch2 + 16, lane);
          wait_dw(s);                                           // dWf0 finished
          tmem_st16(tq + aS + 32 * ch2, g);
          tmem_st16(tq + aS + 32 * ch2 + 16, g + 16);
          st_dymn(dymn[s], row, 32 * ch2, g);
          st_dymn(dymn[s], row, 32 * ch2 + 16, g + 16);
          signal_bwd(s);
          drain16(dwcol(), gw + L.wf0[s], L.ff, 64, 32 * ch2, 64);
          drain16(dwcol(), gw + L.wf0[s], L.ff, 64, 32 * ch2 + 16, 64);
          dwk ^= 1;
        }
        TSTAMP();
        {                               // B3: dv = (do Wo) headmask
          wait_acc(s);
          float g[32];
          ld_acc16(accS + 32 * ch2, g);
          ld_acc16(accS + 32 * ch2 + 16, g + 16);
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const float ds = !rng.on ? 1.f : ((mkh >> hf) & 1u) ? inv1 : 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) g[16 * hf + j] *= ds;
            colsum16(g + 16 * hf, sg + SB + SP_BV + 32 * ch2 + 16 * hf, lane);
          }
          wait_dw(s);                                           // dWo finished
          tmem_st16(tq + aS + 32 * ch2, g);
          tmem_st16(tq + aS + 32 * ch2 + 16, g + 16);
          st_dymn(dymn[s], row, 32 * ch2, g);
          st_dymn(dymn[s], row, 32 * ch2 + 16, g + 16);
          signal_bwd(s);
          drain16(dwcol(), gw + L.wo[s], 64, 64, 32 * ch2, 64);
          drain16(dwcol(), gw + L.wo[s], 64, 64, 32 * ch2 + 16, 64);
          dwk ^= 1;
        }
        TSTAMP();
        {                               // B2: da1 = (dv Wv + dr1) gelu'(a1)
          float g[32], gp[32];
          el_ld32(EL_RES, g);
          el_ld32(EL_GP1, gp);
          wait_acc(s);
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            float t[16];
            ld_acc16(accS + 32 * ch2 + 16 * hf, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) g[16 * hf + j] = (g[16 * hf + j] + t[j]) * gp[16 * hf + j];
            colsum16(g + 16 * hf, sg + SB + SP_BD + 32 * ch2 + 16 * hf, lane);
          }
          wait_dw(s);                                           // dWv finished
          st_dymn(dymn[s], row, 32 * ch2, g);
          st_dymn(dymn[s], row, 32 * ch2 + 16, g + 16);
          signal_bwd(s);
          drain16(dwcol(), gw + L.wv[s], 64, 64, 32 * ch2, 64);
          drain16(dwcol(), gw + L.wv[s], 64, 64, 32 * ch2 + 16, 64);
          dwk ^= 1;
        }
        TSTAMP();
        {                               // B1: drain dWd [64 x kin]
          wait_dw(s);
          float g[16];
          ld_acc16(dwcol(), g);
          const int r = 16 * q + lane, kin = L.kin[s];
          if (ch2 == 0 && lane < 16) for (int j = 0; j < kin; ++j) __stcg(gw + L.wd[s] + r * kin + j, g[j]);
          dwk ^= 1;
        }
        tc_fence_before();
        bar_named(13, kEpi);
        TSTAMP();
        // ---------------------------------------------------------------- small gradients -> gradient row
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const int i = etid + k * kEpi;
          if (i < SP_N) { const int o = sp_src(i); if (o >= 0) gw[o] = sg[i]; }
        }
        __threadfence_block();
        bar_named(13, kEpi);
        TSTAMP();
        // ---------------------------------------------------------------- optimiser step
        ++opt_steps;
        b1t *= (double)P.beta1; b2t *= (double)P.beta2;
        AflOpt o;
        o.kind = P.opt_kind; o.lr = P.lr; o.beta1 = P.beta1; o.beta2 = P.beta2; o.eps = P.eps;
        o.momentum = P.momentum; o.bc1 = (float)(1.0 - b1t); o.bc2 = (float)(1.0 - b2t);
        o.first_step = opt_steps == 1; o.clip_coef = 1.f;
        if (P.clip_proper) {
          float part = 0.f;
          for (int r = 0; r < P.n_ranges; ++r) part += afl_reduce_sqnorm(gw, 1, P.d_pad, P.range_lo[r], P.range_hi[r], etid, kEpi);
          part = warp_sum(part);
          if (lane == 0) red[ew] = part;
          bar_named(13, kEpi);
          float tot = 0.f;
          for (int i = 0; i < kEpi / 32; ++i) tot += red[i];
          const float coef = P.clip / (sqrtf(tot) + 1e-6f);
          o.clip_coef = coef < 1.f ? coef : 1.f;
          bar_named(13, kEpi);
        }
        TS2(8);
        if (CS > 1) {
          // The helpers take the whole pass.  Lane 0's fence is the one cumulative GPU-scope release for the gradient row
          // (the barrier above ordered every thread's stores before it); after the warp sync, lane r hands over to rank r.
          if (ew == 0) {
            if (lane == 0) __threadfence();
            __syncwarp();
            if (lane >= 1 && lane < CS) {
              if (P.clip_proper) remote_st_f32(map_to_rank(ccoef, (uint32_t)lane), o.clip_coef);
              remote_arrive(map_to_rank(&bars[B_GO], (uint32_t)lane));
            }
          }
          TS2(9);
          mbar_wait_cluster(&bars[B_DONE], (uint32_t)(opt_steps - 1) & 1u);
        } else {
          opt_share(o, etid, kEpi);
          __threadfence_block();
        }
        TS2(11);
        bar_named(13, kEpi);
        TSTAMP();
        load_small();
        fence_async();
        bar_named(13, kEpi);
        if (lane == 0) mbar_arrive(&bars[B_STEP]);
        TSTAMP();
      }
    }
    if (etid == 0) {
      atomicAdd(P.loss_out + 2 * c, loss_acc[0]);
      atomicAdd(P.loss_out + 2 * c + 1, loss_acc[1]);
      P.steps_out[c] = opt_steps;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);      // (helper CTAs returned above: only the main CTA owns TMEM)
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

static long long* g_tc_dbg = nullptr;
static int g_tc_last_cluster = 1;

extern "C" {

/* Profiling aid: device buffer of 64 int64 that receives clock64 stamps of the phases of one optimiser step. */
void afl_tc_set_debug(long long* dev_buf) { g_tc_dbg = dev_buf; }

int afl_tc_stash_floats_per_client(void) { return kNSlot * kSlotF; }

/* CTAs per client (cluster size) chosen by the last afl_tc_train_round call of this process. */
int afl_tc_last_cluster(void) { return g_tc_last_cluster; }

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
  P.loss_out = loss_out; P.steps_out = steps_out; P.dbg = g_tc_dbg;
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
  // CTAs per client: clusters of 8 (the portable limit) when every client's cluster can be resident at once, else one CTA
  // per client.  Measured on B200 (8 clients, us / optimiser step): 1 CTA 83.4, 2 CTAs 95.5, 4 CTAs 85.6, 8 CTAs 81.7 -- the
  // hand-over costs two fences and two DSMEM round trips, which only seven helpers' worth of parallel update math repays.
  // ATTACKFL_TC_CLUSTER=1|2|4|8 overrides (tests).
  int cs = 1;
  {
    int forced = 0;
    if (const char* e = getenv("ATTACKFL_TC_CLUSTER")) forced = atoi(e);
    for (int cand = 8; cand > 1; cand >>= 1) {
      if (forced ? cand != forced : cand != 8) continue;
      cudaLaunchConfig_t q = {};
      q.gridDim = dim3((unsigned)(n_clients * cand)); q.blockDim = dim3(kThreads); q.dynamicSmemBytes = kSmemBytes;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = (unsigned)cand; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
      q.attrs = qa; q.numAttrs = 1;
      int max_clusters = 0;
      if (cudaOccupancyMaxActiveClusters(&max_clusters, tc_train_kernel, &q) != cudaSuccess) { (void)cudaGetLastError(); continue; }
      if (max_clusters >= n_clients) { cs = cand; break; }
    }
  }
  g_tc_last_cluster = cs;
  P.cluster = cs;
  if (cs == 1) {
    tc_train_kernel<<<n_clients, kThreads, kSmemBytes, st>>>(maps, P);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n_clients * cs)); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmemBytes; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    AFL_CUDA(cudaLaunchKernelEx(&cfg, tc_train_kernel, maps, P));
  }
  afl_rt::count_launch();
  return afl_rt::check_last("afl_tc_train_round");
}

}  // extern "C"
