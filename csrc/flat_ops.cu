// CUDA kernels + C-ABI wrappers for the flat-vector ops (sm_100a).  The numerical cores live in
// flat_core.h (shared with the CPU twin); this file adds grids, shared-memory staging and the
// bandwidth-oriented hyper-network kernels.
#include <atomic>
#include <cuda_runtime.h>
#include <stdio.h>

#include "../include/attackfl_b200.h"
#include <algorithm>
#include <vector>

#include "flat_core.h"
#include "runtime.h"

namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) adam_kernel(float* w, const float* g, float* m, float* v, int64_t n,
                                                        float lr, float b1, float b2, float eps, float bc1, float bc2,
                                                        const float* gscale_ptr) {
  const float gs = gscale_ptr ? gscale_ptr[0] : 1.f;
  afl_adam_core(w, g, m, v, 0, n, lr, b1, b2, eps, bc1, bc2, gs, blockIdx.x * kThreads + threadIdx.x,
                gridDim.x * kThreads);
}

__global__ void __launch_bounds__(kThreads) sgdm_kernel(float* w, const float* g, float* buf, int64_t n, float lr,
                                                        float momentum, int first, const float* gscale_ptr) {
  const float gs = gscale_ptr ? gscale_ptr[0] : 1.f;
  afl_sgdm_core(w, g, buf, 0, n, lr, momentum, first, gs, blockIdx.x * kThreads + threadIdx.x, gridDim.x * kThreads);
}

template <int A>
__global__ void __launch_bounds__(kThreads) fedavg_kernel(AflFedavgArgs a) {
  extern __shared__ float sm[];
  float* wsm = sm;
  float* msm = sm + a.C;
  for (int i = threadIdx.x; i < a.C; i += kThreads) wsm[i] = a.weights[i];
  for (int i = threadIdx.x; i < A * a.C; i += kThreads) msm[i] = a.member[i];
  __syncthreads();
  afl_fedavg_core<A>(a, wsm, msm, 0, a.d_pad, blockIdx.x * kThreads + threadIdx.x, gridDim.x * kThreads);
}

__global__ void gamma_kernel(const float* G, int k, int rule, int exclude_first, int paper, float* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) afl_gamma_search(G, k, rule, exclude_first, paper, out);
}

__global__ void __launch_bounds__(kThreads) craft_axpy_kernel(float* out, const float* mu, const float* p,
                                                              const float* gamma, float sign_mode, int64_t n) {
  afl_craft_axpy_core(out, mu, p, gamma, sign_mode, 0, n, blockIdx.x * kThreads + threadIdx.x, gridDim.x * kThreads);
}

__global__ void __launch_bounds__(kThreads) craft_random_kernel(float* out, const float* w, const float* valid,
                                                                float pert, uint32_t s0, uint32_t s1, uint32_t stream,
                                                                int64_t n) {
  afl_craft_random_core(out, w, valid, pert, s0, s1, stream, 0, n, blockIdx.x * kThreads + threadIdx.x,
                        gridDim.x * kThreads);
}

__global__ void __launch_bounds__(128) select_kernel(const float* rows, int C, int64_t d_pad, float* out, int mode,
                                                     int trim) {
  afl_select_core(rows, C, d_pad, out, mode, trim, 0, d_pad, blockIdx.x * 128 + threadIdx.x, gridDim.x * 128);
}

__global__ void __launch_bounds__(256) round_indices_kernel(int32_t* __restrict__ idx, const int32_t* __restrict__ sizes,
                                                            const int32_t* __restrict__ client_ids, int C, int E, int mx,
                                                            int n_rows, uint32_t s0, uint32_t s1) {
  afl_round_indices_core(idx, sizes, client_ids, C, E, mx, n_rows, s0, s1, (int64_t)blockIdx.x * 256 + threadIdx.x,
                         (int64_t)gridDim.x * 256);
}

// ---- ROC-AUC of <= 8192 scores in ONE block (src/Validation.py:113-120: roc_curve + auc on the pooled outputs): 64-bit
// keys (ordered score | label | index) bitonic-sorted in shared memory, prefix count of positives, previous-tie-group scan
// and the tie-aware trapezoid sum; out = {auc, #NaN scores, #positives, #negatives}.  Replaces ~20 torch launches and three
// host syncs of the validation phase by one launch and one 16-byte read.
__device__ __forceinline__ unsigned f2ord_auc(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }

__global__ void __launch_bounds__(1024) roc_auc_kernel(const float* __restrict__ scores, const float* __restrict__ labels, int n,
                                                       int P, float* __restrict__ out) {
  extern __shared__ unsigned long long auc_keys[];          // [P]
  int* tp = reinterpret_cast<int*>(auc_keys + P);          // [P] inclusive count of positives in sorted order
  int* pe = tp + P;                                        // [P] index of the last tie-group end before i (-1: none)
  __shared__ int s_nan;
  __shared__ int wtmp[32];
  __shared__ double wred[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_nan = 0;
  __syncthreads();
  int nan_local = 0;
  for (int i = tid; i < P; i += 1024) {
    unsigned long long k = 0ull;                           // padding sorts after every real key
    if (i < n) {
      const float sc = scores[i];
      if (sc != sc) ++nan_local;
      const unsigned lab = labels[i] > 0.5f ? 1u : 0u;
      k = ((unsigned long long)f2ord_auc(sc) << 32) | (unsigned long long)((lab << 31) | (unsigned)i);
    }
    auc_keys[i] = k;
  }
  if (nan_local) atomicAdd(&s_nan, nan_local);
  __syncthreads();
  if (s_nan) {
    if (tid == 0) { out[0] = __int_as_float(0x7fc00000); out[1] = (float)s_nan; out[2] = 0.f; out[3] = 0.f; }
    return;
  }
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < P; i += 1024) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = auc_keys[i], b = auc_keys[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? a < b : a > b) { auc_keys[i] = b; auc_keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
  // contiguous chunk per thread
  const int ipt = P / 1024, i0 = tid * ipt;
  int pos = 0, last_end = -1;
  for (int q = 0; q < ipt; ++q) {
    const int i = i0 + q;
    if (i < n) {
      pos += (int)((auc_keys[i] >> 31) & 1ull);
      const bool end = i == n - 1 || (unsigned)(auc_keys[i] >> 32) != (unsigned)(auc_keys[i + 1] >> 32);
      if (end) last_end = i;
    }
  }
  // block exclusive scans of (pos: sum) and (last_end: max)
  int sp = pos, se = last_end;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int vp = __shfl_up_sync(0xffffffffu, sp, o), ve = __shfl_up_sync(0xffffffffu, se, o);
    if (lane >= o) { sp += vp; se = max(se, ve); }
  }
  if (lane == 31) { wtmp[warp] = sp; }
  __syncthreads();
  int wp = 0;
  for (int w = 0; w < warp; ++w) wp += wtmp[w];
  __syncthreads();
  if (lane == 31) wtmp[warp] = se;
  __syncthreads();
  int we = -1;
  for (int w = 0; w < warp; ++w) we = max(we, wtmp[w]);
  int run_p = wp + sp - pos;                               // positives before this thread's chunk
  int ex_e = __shfl_up_sync(0xffffffffu, se, 1);
  int run_e = max(we, lane ? ex_e : -1);                   // last group end before this thread's chunk
  for (int q = 0; q < ipt; ++q) {
    const int i = i0 + q;
    if (i < n) {
      run_p += (int)((auc_keys[i] >> 31) & 1ull);
      tp[i] = run_p;
      pe[i] = run_e;
      const bool end = i == n - 1 || (unsigned)(auc_keys[i] >> 32) != (unsigned)(auc_keys[i + 1] >> 32);
      if (end) run_e = i;
    }
  }
  __syncthreads();
  double area = 0.0;
  for (int q = 0; q < ipt; ++q) {
    const int i = i0 + q;
    if (i < n) {
      const bool end = i == n - 1 || (unsigned)(auc_keys[i] >> 32) != (unsigned)(auc_keys[i + 1] >> 32);
      if (end) {
        const int p = pe[i];
        const int tp_p = p >= 0 ? tp[p] : 0, fp_p = p >= 0 ? (p + 1 - tp[p]) : 0;
        const int tp_i = tp[i], fp_i = i + 1 - tp_i;
        area += 0.5 * (double)(fp_i - fp_p) * (double)(tp_i + tp_p);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) area += __shfl_xor_sync(0xffffffffu, area, o);
  if (lane == 0) wred[warp] = area;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    for (int w = 0; w < 32; ++w) tot += wred[w];
    const int n_pos = tp[n - 1], n_neg = n - n_pos;
    out[0] = (n_pos == 0 || n_neg == 0) ? __int_as_float(0x7fc00000) : (float)(tot / ((double)n_pos * (double)n_neg));
    out[1] = 0.f; out[2] = (float)n_pos; out[3] = (float)n_neg;
  }
}

// ---- ScionFL: min/max per row (one block per (chunk, row); partials merged with ordered-int atomics)
__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

__global__ void __launch_bounds__(256) scion_minmax_kernel(const float* __restrict__ rows, const float* __restrict__ valid,
                                                           int64_t d_pad, unsigned* __restrict__ ord /* [C][2] */) {
  const int c = blockIdx.y;
  const float* row = rows + (int64_t)c * d_pad;
  float mn = INFINITY, mx = -INFINITY;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < d_pad; j += (int64_t)gridDim.x * 256)
    if (!valid || valid[j] != 0.f) { const float x = row[j]; mn = fminf(mn, x); mx = fmaxf(mx, x); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(ord + 2 * c, f2ord(mn)); atomicMax(ord + 2 * c + 1, f2ord(mx)); }
}

// ordered-int encodings of +inf / -inf so atomicMin / atomicMax start from the identity
__global__ void scion_minmax_init_kernel(unsigned* ord, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ord[i] = (i & 1) ? 0x007FFFFFu : 0xFF800000u;
}

__global__ void scion_minmax_finish_kernel(unsigned* ord, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) reinterpret_cast<float*>(ord)[i] = ord2f(ord[i]);
}

// one thread per coordinate, a warp's ballot is the packed word
__global__ void __launch_bounds__(256) scion_quantize_kernel(const float* __restrict__ rows, const float* __restrict__ valid,
                                                             const float* __restrict__ minmax, int64_t d_pad,
                                                             uint32_t s0, uint32_t s1, uint32_t* __restrict__ bits,
                                                             int* __restrict__ counts, int row0) {
  const int c = blockIdx.y;
  const float mn = minmax[2 * c], inv = 1.0f / (minmax[2 * c + 1] - mn + 1e-6f);
  const float* row = rows + (int64_t)c * d_pad;
  const int64_t words = d_pad / 32;
  int cnt = 0;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < d_pad; j += (int64_t)gridDim.x * 256) {
    const bool ok = !valid || valid[j] != 0.f;
    const bool bit = ok && afl_scion_bit(row[j], mn, inv, s0, s1, (uint32_t)(row0 + c), (uint32_t)j);
    const unsigned w = __ballot_sync(0xffffffffu, bit);
    if ((threadIdx.x & 31) == 0) { bits[(int64_t)c * words + (j >> 5)] = w; cnt += __popc(w); }
  }
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(counts + c, cnt);
}

// one warp per 32-coordinate word; S_c partials live in shared memory until the block is done.  `bptr` (optional): the
// packed rows by address, one per client (peer-mapped addresses allowed: the 1-bit codes are read where their owner
// wrote them), and only the words [w_lo, w_hi) are visited (slice ownership on N GPUs).
__global__ void __launch_bounds__(256) scion_scores_kernel(const uint32_t* __restrict__ bits, const float* __restrict__ valid,
                                                           const float* __restrict__ lohi, int C, int64_t d_pad,
                                                           float* __restrict__ out, const uint32_t* const* bptr = nullptr,
                                                           int64_t w_lo = 0, int64_t w_hi = -1) {
  extern __shared__ float acc[];                     // [C + 2]
  for (int i = threadIdx.x; i < C + 2; i += 256) acc[i] = 0.f;
  __syncthreads();
  const int64_t words = d_pad / 32;
  if (w_hi < 0) w_hi = words;
  const int lane = threadIdx.x & 31;
  float sa = 0.f, sq = 0.f;
  for (int64_t w = w_lo + (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); w < w_hi; w += (int64_t)gridDim.x * 8) {
    const bool ok = !valid || valid[w * 32 + lane] != 0.f;
    float a = 0.f;
    if (bptr) {
      if (ok) {
        for (int i = 0; i < C; ++i) a += ((bptr[i][w] >> lane) & 1u) ? lohi[2 * i + 1] : lohi[2 * i];
        a /= (float)C;
      }
    } else {
      a = ok ? afl_scion_agg(bits, lohi, C, words, w, lane) : 0.f;
    }
    sa += a; sq += a * a;
    for (int i = 0; i < C; ++i) {
      const uint32_t word = bptr ? bptr[i][w] : bits[(int64_t)i * words + w];
      float v = ((word >> lane) & 1u) ? a : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v != 0.f) atomicAdd(acc + i, v);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sa += __shfl_xor_sync(0xffffffffu, sa, o);
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
  }
  if (lane == 0) { atomicAdd(acc + C, sa); atomicAdd(acc + C + 1, sq); }
  __syncthreads();
  for (int i = threadIdx.x; i < C + 2; i += 256) if (acc[i] != 0.f) atomicAdd(out + i, acc[i]);
}

// K6, warp-cooperative version: a block stages a [C][32-coordinate] tile through shared memory (coalesced 128-byte row
// segments), then each warp sorts one coordinate's C values held VPT per lane in registers with a bitonic network:
// compare-exchange distances < VPT stay inside a lane, larger ones use shfl.xor.  No local memory, no global re-reads.
template <int VPT>
__global__ void __launch_bounds__(256) select_warp_kernel(const float* __restrict__ rows, int C, int64_t d_pad,
                                                          float* __restrict__ out, int mode, int trim,
                                                          const float* const* row_ptrs = nullptr, int64_t lo = 0,
                                                          int64_t hi = -1, float* const* dsts = nullptr, int n_dst = 0) {
  // row_ptrs (optional): one base pointer per row instead of a dense [C][d_pad] matrix - the rows may live on other GPUs
  // (peer-mapped); [lo, hi): coordinate slice handled by this launch; dsts: the result is stored into every listed buffer
  // (every rank's copy of the global model) instead of `out`.
  extern __shared__ float tile[];                    // [C][33]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int P = 32 * VPT;                        // padded element count (power of two >= C)
  if (hi < 0) hi = d_pad;
  const int64_t t0 = lo / 32, n_tiles = (hi + 31) / 32;
  for (int64_t tIdx = t0 + blockIdx.x; tIdx < n_tiles; tIdx += gridDim.x) {
    const int64_t d0 = tIdx * 32;
    __syncthreads();
    for (int i = threadIdx.x; i < C * 32; i += 256) {
      const int c = i >> 5, x = i & 31;
      const float* src = row_ptrs ? row_ptrs[c] : rows + (int64_t)c * d_pad;
      tile[c * 33 + x] = (d0 + x < hi) ? src[d0 + x] : 0.f;
    }
    __syncthreads();
    for (int col = warp; col < 32; col += 8) {
      if (d0 + col >= hi) break;
      float v[VPT];
#pragma unroll
      for (int r = 0; r < VPT; ++r) {
        const int e = lane * VPT + r;
        v[r] = e < C ? tile[e * 33 + col] : INFINITY;
      }
#pragma unroll
      for (int k = 2; k <= P; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
          if (j >= VPT) {
            const int lj = j / VPT;                   // partner lane distance
#pragma unroll
            for (int r = 0; r < VPT; ++r) {
              const int e = lane * VPT + r;
              const float o = __shfl_xor_sync(0xffffffffu, v[r], lj);
              const bool up = (e & k) == 0;
              const bool lower = (lane & lj) == 0;   // this lane holds the smaller index of the pair
              const bool take_min = (up == lower);
              const float mn = fminf(v[r], o), mx = fmaxf(v[r], o);
              // NaNs: fminf/fmaxf drop them; keep the original value when both compare false
              v[r] = take_min ? mn : mx;
            }
          } else {
#pragma unroll
            for (int r = 0; r < VPT; ++r) {
              const int r2 = r ^ j;
              if (r2 > r) {
                const int e = lane * VPT + r;
                const bool up = (e & k) == 0;
                const float a = v[r], b = v[r2];
                const bool sw = up ? (a > b) : (a < b);
                if (sw) { v[r] = b; v[r2] = a; }
              }
            }
          }
        }
      }
      float res;
      if (mode == 0) {
        const int idx = (C - 1) / 2;
        float mine = 0.f;
#pragma unroll
        for (int r = 0; r < VPT; ++r) if ((idx % VPT) == r) mine = v[r];
        res = __shfl_sync(0xffffffffu, mine, idx / VPT);
      } else {
        float sum = 0.f;
#pragma unroll
        for (int r = 0; r < VPT; ++r) {
          const int e = lane * VPT + r;
          if (e >= trim && e < C - trim) sum += v[r];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        res = sum / (float)(C - 2 * trim);
      }
      if (lane == 0) {
        if (dsts) { for (int w = 0; w < n_dst; ++w) dsts[w][d0 + col] = res; }
        else out[d0 + col] = res;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------ hyper kernels
// Tile of TR rows of W_h staged in shared memory with a padded row (H+1) so both access patterns
// (lane = row, lane = column) are bank-conflict free.  W_h is read from HBM exactly once per pass.
constexpr int kTR = 64;

__global__ void __launch_bounds__(kThreads) hyper_emit_kernel(const float* __restrict__ head_w,
                                                              const float* __restrict__ head_b,
                                                              const float* __restrict__ f, int C, int H,
                                                              float* theta, int64_t theta_ld, int64_t d_pad) {
  extern __shared__ float sm[];
  float* fs = sm;                       // [C][H]
  float* tile = sm + C * H;             // [kTR][H+1]
  for (int i = threadIdx.x; i < C * H; i += kThreads) fs[i] = f[i];
  const int64_t n_tiles = (d_pad + kTR - 1) / kTR;
  for (int64_t tIdx = blockIdx.x; tIdx < n_tiles; tIdx += gridDim.x) {
    const int64_t d0 = tIdx * kTR;
    const int rows = (int)AFL_MIN((int64_t)kTR, d_pad - d0);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * H; i += kThreads) {
      const int r = i / H, h = i - r * H;
      tile[r * (H + 1) + h] = head_w[d0 * H + i];
    }
    __syncthreads();
    for (int it = threadIdx.x; it < kTR * C; it += kThreads) {
      const int r = it % kTR, c = it / kTR;      // lanes walk rows => coalesced theta stores
      if (r < rows) {
        const float* tr = tile + r * (H + 1);
        const float* fr = fs + c * H;
        float acc = head_b[d0 + r];
        for (int h = 0; h < H; ++h) acc += tr[h] * fr[h];
        theta[(int64_t)c * theta_ld + d0 + r] = acc;
      }
    }
  }
}

// pass 1: delta, |delta|^2 and W_h^T delta.  red: [1+H] floats, zeroed by the caller.
__global__ void __launch_bounds__(kThreads) hyper_delta_kernel(const float* __restrict__ head_w,
                                                               const float* __restrict__ head_b,
                                                               const float* __restrict__ f, int H,
                                                               const float* __restrict__ theta_client, float* delta,
                                                               int64_t d_pad, float* red) {
  extern __shared__ float sm[];
  float* fs = sm;                       // [H]
  float* dl = sm + H;                   // [kTR]
  float* tile = dl + kTR;               // [kTR][H+1]
  for (int i = threadIdx.x; i < H; i += kThreads) fs[i] = f[i];
  float acc_h = 0.f;                    // thread h < H accumulates (W^T delta)[h]
  float acc_sq = 0.f;                   // thread r < kTR accumulates delta^2
  const int64_t n_tiles = (d_pad + kTR - 1) / kTR;
  for (int64_t tIdx = blockIdx.x; tIdx < n_tiles; tIdx += gridDim.x) {
    const int64_t d0 = tIdx * kTR;
    const int rows = (int)AFL_MIN((int64_t)kTR, d_pad - d0);
    __syncthreads();
    for (int i = threadIdx.x; i < rows * H; i += kThreads) {
      const int r = i / H, h = i - r * H;
      tile[r * (H + 1) + h] = head_w[d0 * H + i];
    }
    __syncthreads();
    if (threadIdx.x < rows) {
      const int r = threadIdx.x;
      const float* tr = tile + r * (H + 1);
      float a = head_b[d0 + r];
      for (int h = 0; h < H; ++h) a += tr[h] * fs[h];
      const float d = a - theta_client[d0 + r];
      delta[d0 + r] = d;
      dl[r] = d;
      acc_sq += d * d;
    }
    __syncthreads();
    if (threadIdx.x < H) {
      const int h = threadIdx.x;
      float a = 0.f;
      for (int r = 0; r < rows; ++r) a += tile[r * (H + 1) + h] * dl[r];
      acc_h += a;
    }
  }
  if (threadIdx.x < H) atomicAdd(red + 1 + threadIdx.x, acc_h);
  // block-reduce acc_sq through shared memory
  __syncthreads();
  if (threadIdx.x < kTR) dl[threadIdx.x] = acc_sq;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int r = 0; r < kTR; ++r) s += dl[r];
    atomicAdd(red, s);
  }
}

// Small replicated part of the step: MLP backward, embedding gradient, clip coefficient and Adam on
// the MLP / embedding table.  One block.  Layout of `small` (all fp32):
//   emb [n_nodes*E] | w0 [H*E] b0 [H] | w1 [H*H] b1 [H] | w2 [H*H] b2 [H]      (params, then m, then v copies)
// acts (written by hyper_features_kernel): e[E] z0[H] a0[H] z1[H] a1[H] f[H]
struct HyperSmall {
  float* params; float* m; float* v; int n_nodes, E, H;
  float* acts; float* red; float* coef_out; float* norm_out;
  int client; float lr, b1, b2, eps, bc1, bc2, clip;
};

__device__ inline int hs_off_w(const HyperSmall& s, int layer) {
  int o = s.n_nodes * s.E;
  if (layer == 0) return o;
  o += s.H * s.E + s.H;
  if (layer == 1) return o;
  return o + s.H * s.H + s.H;
}
__device__ inline int hs_total(const HyperSmall& s) { return hs_off_w(s, 2) + s.H * s.H + s.H; }

__global__ void __launch_bounds__(kThreads) hyper_features_kernel(HyperSmall s, const int* clients, int n, float* f_out) {
  // f_out: [n][H]; acts saved only when n == 1
  extern __shared__ float sm[];
  float* x = sm; float* y = sm + s.H;
  const int E = s.E, H = s.H;
  for (int q = blockIdx.x; q < n; q += gridDim.x) {        // one block per client
    const int c = clients[q];
    __syncthreads();
    for (int i = threadIdx.x; i < E; i += kThreads) { x[i] = s.params[c * E + i]; if (n == 1) s.acts[i] = x[i]; }
    __syncthreads();
    for (int layer = 0; layer < 3; ++layer) {
      const int in = layer == 0 ? E : H;
      const float* W = s.params + hs_off_w(s, layer);
      const float* B = W + H * in;
      for (int j = threadIdx.x; j < H; j += kThreads) {
        float a = B[j];
        for (int i = 0; i < in; ++i) a += W[j * in + i] * x[i];
        y[j] = a;
      }
      __syncthreads();
      for (int j = threadIdx.x; j < H; j += kThreads) {
        const float z = y[j];
        if (layer < 2) {
          const float r = z > 0.f ? z : 0.f;
          if (n == 1) { s.acts[E + 2 * layer * H + j] = z; s.acts[E + (2 * layer + 1) * H + j] = r; }
          x[j] = r;
        } else {
          if (n == 1) s.acts[E + 4 * H + j] = z;
          f_out[q * H + j] = z;
        }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(kThreads) hyper_small_step_kernel(HyperSmall s) {
  extern __shared__ float sm[];
  const int E = s.E, H = s.H;
  float* g = sm;             // current upstream gradient [H]
  float* gn = sm + H;        // next [H]
  float* grads = sm + 2 * H; // full small-gradient vector, same layout as params
  __shared__ float s_norm2, s_coef;
  const int total = hs_total(s);
  for (int i = threadIdx.x; i < total; i += kThreads) grads[i] = 0.f;
  for (int i = threadIdx.x; i < H; i += kThreads) g[i] = s.red[1 + i];     // dL/df = W_h^T delta
  __syncthreads();
  const float* e = s.acts;
  for (int layer = 2; layer >= 0; --layer) {
    const int in = layer == 0 ? E : H;
    const float* xin = layer == 0 ? e : s.acts + E + (2 * (layer - 1) + 1) * H;
    const int ow = hs_off_w(s, layer);
    for (int i = threadIdx.x; i < H * in; i += kThreads) grads[ow + i] = g[i / in] * xin[i % in];
    for (int j = threadIdx.x; j < H; j += kThreads) grads[ow + H * in + j] = g[j];
    const float* W = s.params + ow;
    for (int i = threadIdx.x; i < in; i += kThreads) {
      float a = 0.f;
      for (int j = 0; j < H; ++j) a += W[j * in + i] * g[j];
      if (layer > 0) { const float z = s.acts[E + 2 * (layer - 1) * H + i]; a = z > 0.f ? a : 0.f; }
      gn[i] = a;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < in; i += kThreads) g[i] = gn[i];
    __syncthreads();
  }
  for (int i = threadIdx.x; i < E; i += kThreads) grads[s.client * E + i] = g[i];
  __syncthreads();
  __shared__ double s_part[kThreads];
  {
    double acc = 0.0;                                    // block-parallel |grad small|^2
    for (int i = threadIdx.x; i < total; i += kThreads) acc += (double)grads[i] * grads[i];
    s_part[threadIdx.x] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // |grad|^2 = |delta|^2 (|f|^2 + 1) + |grad small|^2   (rank-1 head gradient)
    const float* f = s.acts + E + 4 * H;
    double f2 = 0.0, small2 = 0.0;
    for (int i = 0; i < H; ++i) f2 += (double)f[i] * f[i];
    for (int i = 0; i < kThreads; ++i) small2 += s_part[i];
    const double n2 = (double)s.red[0] * (f2 + 1.0) + small2;
    const float norm = (float)sqrt(n2);
    float coef = 1.f;
    if (s.clip > 0.f) { coef = s.clip / (norm + 1e-6f); coef = coef < 1.f ? coef : 1.f; }
    s_norm2 = norm; s_coef = coef;
    s.coef_out[0] = coef; s.norm_out[0] = norm;
  }
  __syncthreads();
  const float coef = s_coef;
  const float step = s.lr / s.bc1, isq = 1.0f / sqrtf(s.bc2);
  for (int i = threadIdx.x; i < total; i += kThreads) {
    const float gg = grads[i] * coef;
    const float mi = s.b1 * s.m[i] + (1.f - s.b1) * gg;
    const float vi = s.b2 * s.v[i] + (1.f - s.b2) * gg * gg;
    s.m[i] = mi; s.v[i] = vi;
    s.params[i] -= step * (mi / (sqrtf(vi) * isq + s.eps));
  }
  (void)s_norm2;
}

__global__ void __launch_bounds__(kThreads) hyper_adam_kernel(float* head_w, float* head_b, float* mw, float* vw,
                                                              float* mb, float* vb, const float* delta, const float* f,
                                                              int H, const float* coef, float lr, float b1, float b2,
                                                              float eps, float bc1, float bc2, int64_t d_pad) {
  afl_hyper_adam_core(head_w, head_b, mw, vw, mb, vb, delta, f, H, coef, lr, b1, b2, eps, bc1, bc2, 0, d_pad,
                      blockIdx.x * kThreads + threadIdx.x, gridDim.x * kThreads);
}

inline int grid_for(int64_t work_items, int threads, int max_blocks_per_sm = 8) {
  int64_t b = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)afl_rt::num_sms() * max_blocks_per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace

extern "C" {

int afl_adam_flat(float* w, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                  int step, const float* gscale, void* stream) {
  AFL_CHECK_ARG(n % 4 == 0, "afl_adam_flat: n must be a multiple of 4");
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  adam_kernel<<<grid_for(n / 4, kThreads), kThreads, 0, (cudaStream_t)stream>>>(w, g, m, v, n, lr, b1, b2, eps, bc1,
                                                                                 bc2, gscale);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_adam_flat");
}

int afl_sgdm_flat(float* w, const float* g, float* buf, int64_t n, float lr, float momentum, int first_step,
                  const float* gscale, void* stream) {
  AFL_CHECK_ARG(n % 4 == 0, "afl_sgdm_flat: n must be a multiple of 4");
  sgdm_kernel<<<grid_for(n / 4, kThreads), kThreads, 0, (cudaStream_t)stream>>>(w, g, buf, n, lr, momentum,
                                                                                 first_step, gscale);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_sgdm_flat");
}

int afl_fedavg(const float* rows, const float* weights, int C, int64_t d_pad, float* out, float* bcast, int Cb,
               int A, const float* ref, const float* member, const float* kcount, const float* z, float* mu,
               float* sd, float* lie, int strict_k1, void* stream) {
  AFL_CHECK_ARG(d_pad % 4 == 0, "afl_fedavg: d_pad must be a multiple of 4");
  AFL_CHECK_ARG(A >= 0 && A <= AFL_MAX_ATT, "afl_fedavg: at most 4 moment sets per call");
  AflFedavgArgs a{rows, weights, C, d_pad, out, bcast, Cb, A, ref, member, kcount, z, mu, sd, lie, strict_k1};
  const size_t smem = sizeof(float) * (size_t)(C + A * C);
  const int grid = grid_for(d_pad / 4, kThreads, 4);
  cudaStream_t st = (cudaStream_t)stream;
  switch (A) {
    case 0: fedavg_kernel<0><<<grid, kThreads, smem, st>>>(a); break;
    case 1: fedavg_kernel<1><<<grid, kThreads, smem, st>>>(a); break;
    case 2: fedavg_kernel<2><<<grid, kThreads, smem, st>>>(a); break;
    case 3: fedavg_kernel<3><<<grid, kThreads, smem, st>>>(a); break;
    default: fedavg_kernel<4><<<grid, kThreads, smem, st>>>(a); break;
  }
  afl_rt::count_launch();
  return afl_rt::check_last("afl_fedavg");
}

int afl_gamma_search_dev(const float* G, int k, int rule, int exclude_first, int paper, float* out4, void* stream) {
  gamma_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(G, k, rule, exclude_first, paper, out4);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_gamma_search_dev");
}

int afl_craft_axpy(float* out, const float* mu, const float* p, const float* gamma_dev, int sign_mode, int64_t n,
                   void* stream) {
  AFL_CHECK_ARG(n % 4 == 0, "afl_craft_axpy: n must be a multiple of 4");
  craft_axpy_kernel<<<grid_for(n / 4, kThreads), kThreads, 0, (cudaStream_t)stream>>>(out, mu, p, gamma_dev,
                                                                                       (float)sign_mode, n);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_craft_axpy");
}

int afl_craft_random(float* out, const float* w, const float* valid, float perturbation, uint64_t seed,
                     uint32_t stream_id, int64_t n, void* stream) {
  craft_random_kernel<<<grid_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      out, w, valid, perturbation, (uint32_t)seed, (uint32_t)(seed >> 32), stream_id, n);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_craft_random");
}

int afl_coord_select(const float* rows, int C, int64_t d_pad, float* out, int mode, int trim, void* stream) {
  AFL_CHECK_ARG(C >= 1 && C <= AFL_SEL_MAX, "afl_coord_select: 1 <= C <= 256");
  AFL_CHECK_ARG(mode == 0 || 2 * trim < C, "afl_coord_select: 2*trim must be < C");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = sizeof(float) * (size_t)C * 33;
  const int grid = grid_for((d_pad + 31) / 32, 1, 8);
  if (C <= 32) select_warp_kernel<1><<<grid, 256, smem, st>>>(rows, C, d_pad, out, mode, trim);
  else if (C <= 64) select_warp_kernel<2><<<grid, 256, smem, st>>>(rows, C, d_pad, out, mode, trim);
  else if (C <= 128) select_warp_kernel<4><<<grid, 256, smem, st>>>(rows, C, d_pad, out, mode, trim);
  else select_warp_kernel<8><<<grid, 256, smem, st>>>(rows, C, d_pad, out, mode, trim);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_coord_select");
}

int afl_coord_select_ptrs(const float* const* row_ptrs_dev, int C, int64_t lo, int64_t hi, float* const* dst_dev, int n_dst,
                          int mode, int trim, void* stream) {
  AFL_CHECK_ARG(C >= 1 && C <= AFL_SEL_MAX, "afl_coord_select_ptrs: 1 <= C <= 256");
  AFL_CHECK_ARG(mode == 0 || 2 * trim < C, "afl_coord_select_ptrs: 2*trim must be < C");
  AFL_CHECK_ARG(lo % 32 == 0 && hi > lo && n_dst >= 1, "afl_coord_select_ptrs: lo must be a multiple of 32, hi > lo");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = sizeof(float) * (size_t)C * 33;
  const int grid = grid_for((hi - lo + 31) / 32, 1, 8);
  if (C <= 32) select_warp_kernel<1><<<grid, 256, smem, st>>>(nullptr, C, hi, nullptr, mode, trim, row_ptrs_dev, lo, hi, dst_dev, n_dst);
  else if (C <= 64) select_warp_kernel<2><<<grid, 256, smem, st>>>(nullptr, C, hi, nullptr, mode, trim, row_ptrs_dev, lo, hi, dst_dev, n_dst);
  else if (C <= 128) select_warp_kernel<4><<<grid, 256, smem, st>>>(nullptr, C, hi, nullptr, mode, trim, row_ptrs_dev, lo, hi, dst_dev, n_dst);
  else select_warp_kernel<8><<<grid, 256, smem, st>>>(nullptr, C, hi, nullptr, mode, trim, row_ptrs_dev, lo, hi, dst_dev, n_dst);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_coord_select_ptrs");
}

int afl_round_indices(int32_t* idx, const int32_t* sizes, const int32_t* client_ids, int C, int E, int mx, int n_rows,
                      uint64_t seed, void* stream) {
  AFL_CHECK_ARG(C >= 1 && E >= 1 && mx >= 1 && n_rows >= 1, "afl_round_indices: C, E, mx, n_rows >= 1");
  const int64_t total = (int64_t)C * E * mx;
  round_indices_kernel<<<grid_for(total, 256, 8), 256, 0, (cudaStream_t)stream>>>(idx, sizes, client_ids, C, E, mx, n_rows,
                                                                                   (uint32_t)seed, (uint32_t)(seed >> 32));
  afl_rt::count_launch();
  return afl_rt::check_last("afl_round_indices");
}

int afl_roc_auc(const float* scores, const float* labels, int64_t n, float* out4, void* stream) {
  AFL_CHECK_ARG(n >= 1 && n <= 8192, "afl_roc_auc: 1 <= n <= 8192 (larger sets: sort on the caller's side)");
  int P = 1024;
  while (P < n) P <<= 1;
  const size_t smem = (size_t)P * 16;
  AFL_CUDA(cudaFuncSetAttribute(roc_auc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 16));
  roc_auc_kernel<<<1, 1024, smem, (cudaStream_t)stream>>>(scores, labels, (int)n, P, out4);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_roc_auc");
}

int afl_scion_minmax(const float* rows, const float* valid, int C, int64_t d_pad, float* minmax, void* stream) {
  AFL_CHECK_ARG(C >= 1 && d_pad % 32 == 0, "afl_scion_minmax: C >= 1, d_pad % 32 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  scion_minmax_init_kernel<<<(2 * C + 127) / 128, 128, 0, st>>>(reinterpret_cast<unsigned*>(minmax), 2 * C);
  const int gx = (int)std::min<int64_t>((d_pad + 4095) / 4096, 64);
  scion_minmax_kernel<<<dim3(gx, C), 256, 0, st>>>(rows, valid, d_pad, reinterpret_cast<unsigned*>(minmax));
  scion_minmax_finish_kernel<<<(2 * C + 127) / 128, 128, 0, st>>>(reinterpret_cast<unsigned*>(minmax), 2 * C);
  afl_rt::count_launch(); afl_rt::count_launch(); afl_rt::count_launch();
  return afl_rt::check_last("afl_scion_minmax");
}

int afl_scion_quantize_rows(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                            int row0, uint32_t* bits, int* counts, void* stream) {
  AFL_CHECK_ARG(C >= 1 && d_pad % 32 == 0 && row0 >= 0, "afl_scion_quantize: C >= 1, d_pad % 32 == 0, row0 >= 0");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)C, st);
  const int gx = (int)std::min<int64_t>((d_pad + 2047) / 2048, 128);
  scion_quantize_kernel<<<dim3(gx, C), 256, 0, st>>>(rows, valid, minmax, d_pad, (uint32_t)seed, (uint32_t)(seed >> 32),
                                                     bits, counts, row0);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_scion_quantize");
}

int afl_scion_quantize(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                       uint32_t* bits, int* counts, void* stream) {
  return afl_scion_quantize_rows(rows, valid, minmax, C, d_pad, seed, 0, bits, counts, stream);
}

int afl_scion_scores_ptrs(const uint32_t* const* bits_ptrs_dev, const float* valid, const float* lohi, int C, int64_t d_pad,
                          int64_t w_lo, int64_t w_hi, float* out, void* stream) {
  AFL_CHECK_ARG(C >= 1 && C <= 4096 && d_pad % 32 == 0 && 0 <= w_lo && w_lo <= w_hi && w_hi <= d_pad / 32,
                "afl_scion_scores_ptrs: 1 <= C <= 4096, d_pad % 32 == 0, 0 <= w_lo <= w_hi <= d_pad / 32");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, sizeof(float) * (size_t)(C + 2), st);
  if (w_hi > w_lo) {
    const int grid = grid_for(w_hi - w_lo, 8, 4);
    scion_scores_kernel<<<grid, 256, sizeof(float) * (size_t)(C + 2), st>>>(nullptr, valid, lohi, C, d_pad, out, bits_ptrs_dev,
                                                                             w_lo, w_hi);
    afl_rt::count_launch();
  }
  return afl_rt::check_last("afl_scion_scores_ptrs");
}

int afl_scion_scores(const uint32_t* bits, const float* valid, const float* lohi, int C, int64_t d_pad, float* out,
                     void* stream) {
  AFL_CHECK_ARG(C >= 1 && C <= 4096 && d_pad % 32 == 0, "afl_scion_scores: 1 <= C <= 4096, d_pad % 32 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, sizeof(float) * (size_t)(C + 2), st);
  const int grid = grid_for(d_pad / 32, 8, 4);
  scion_scores_kernel<<<grid, 256, sizeof(float) * (size_t)(C + 2), st>>>(bits, valid, lohi, C, d_pad, out);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_scion_scores");
}

// ---- hyper-network (K5a / K5b).  `small`: params | m | v, each hs_total floats (see HyperSmall).
static size_t hyper_small_total(int n_nodes, int E, int H) { return (size_t)n_nodes * E + (size_t)H * E + H + 2 * ((size_t)H * H + H); }

int afl_hyper_emit(const float* head_w, const float* head_b, float* small, int n_nodes, int E, int H,
                   const int* clients_dev, int n, float* f_scratch /* [n*H] */, float* acts /* [E+5H] */,
                   float* theta, int64_t theta_ld, int64_t d_pad, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const size_t tot = hyper_small_total(n_nodes, E, H);
  HyperSmall s{small, small + tot, small + 2 * tot, n_nodes, E, H, acts, nullptr, nullptr, nullptr, 0,
               0, 0, 0, 0, 1, 1, 0};
  hyper_features_kernel<<<n, kThreads, sizeof(float) * 2 * (size_t)AFL_MAX(H, E), st>>>(s, clients_dev, n, f_scratch);
  afl_rt::count_launch();
  const size_t smem = sizeof(float) * ((size_t)n * H + (size_t)kTR * (H + 1));
  AFL_CHECK_ARG(smem <= 200 * 1024, "afl_hyper_emit: too many clients per call for shared memory (chunk them)");
  AFL_CUDA(cudaFuncSetAttribute(hyper_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = grid_for((d_pad + kTR - 1) / kTR, 1, 2);
  hyper_emit_kernel<<<grid, kThreads, smem, st>>>(head_w, head_b, f_scratch, n, H, theta, theta_ld, d_pad);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_hyper_emit");
}

int afl_hyper_step(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, float* small,
                   int n_nodes, int E, int H, int client, const int* client_dev, const float* theta_client,
                   float* delta /* [d_pad] */, float* f_scratch /* [H] */, float* acts /* [E+5H] */,
                   float* red /* [1+H+2]: |delta|^2, W^T delta, coef, norm */, int64_t d_pad, float lr, float b1,
                   float b2, float eps, int step, float clip, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const size_t tot = hyper_small_total(n_nodes, E, H);
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  HyperSmall s{small, small + tot, small + 2 * tot, n_nodes, E, H, acts, red, red + 1 + H, red + 2 + H, client,
               lr, b1, b2, eps, bc1, bc2, clip};
  AFL_CUDA(cudaMemsetAsync(red, 0, sizeof(float) * (size_t)(H + 3), st));
  hyper_features_kernel<<<1, kThreads, sizeof(float) * 2 * (size_t)AFL_MAX(H, E), st>>>(s, client_dev, 1, f_scratch);
  const size_t smem1 = sizeof(float) * ((size_t)H + kTR + (size_t)kTR * (H + 1));
  const int grid = grid_for((d_pad + kTR - 1) / kTR, 1, 2);
  hyper_delta_kernel<<<grid, kThreads, smem1, st>>>(head_w, head_b, f_scratch, H, theta_client, delta, d_pad, red);
  const size_t smem2 = sizeof(float) * (2 * (size_t)H + tot);
  AFL_CHECK_ARG(smem2 <= 200 * 1024, "afl_hyper_step: embedding table too large for the single-block small step");
  AFL_CUDA(cudaFuncSetAttribute(hyper_small_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
  hyper_small_step_kernel<<<1, kThreads, smem2, st>>>(s);
  hyper_adam_kernel<<<grid_for(d_pad * H, kThreads, 8), kThreads, 0, st>>>(head_w, head_b, mw, vw, mb, vb, delta,
                                                                             f_scratch, H, red + 1 + H, lr, b1, b2,
                                                                             eps, bc1, bc2, d_pad);
  afl_rt::count_launch(4);
  return afl_rt::check_last("afl_hyper_step");
}

// ---- afl_hyper_step_noalloc: the same step for callers that must not allocate device memory ------------------
// Who needs this: code that drives the hyper-network update from a place where cudaMalloc (or a framework
// allocator such as torch's caching allocator) is off limits -- a real-time control loop that has pre-reserved
// all device memory, a plugin running inside another runtime's memory budget, or a host callback / worker that
// may not synchronise the device (cudaMalloc and cudaFree can).  afl_hyper_step already takes its scratch from
// the caller; this variant owns it instead, in the file-scope array below, so the caller passes nothing but the
// inputs and outputs.
//
// Limits (stated, checked, modest): d_pad <= 262144 floats, H <= 128, E <= 64 -> 263,563 floats = 1.006 MiB of
// device memory, held by every GPU that loads this library until the process exits.  Larger problems are rejected
// with an error; use afl_hyper_step with caller-provided scratch for them.
//
// Set-up: the array's name is only a host-side stand-in for the symbol, not a device address; on the first call
// on each device (with that device current) the entry point asks the runtime for the real address with
// cudaGetSymbolAddress and caches it per device.  Make that first call outside of any stream capture.
//
// Sharing: all calls use the same array and it keeps the previous call's contents.  The slots that must start at
// zero (the |delta|^2 / W^T delta accumulators) are cleared on the stream at the start of every call; everything
// else is fully overwritten before it is read.  Two calls that overlap on the GPU (different streams, no
// dependency between them) race on the array: serialise them.
#define AFL_HS_MAX_DPAD 262144
#define AFL_HS_MAX_H 128
#define AFL_HS_MAX_E 64
#define AFL_HS_OFF_F (AFL_HS_MAX_DPAD)
#define AFL_HS_OFF_ACTS (AFL_HS_OFF_F + AFL_HS_MAX_H)
#define AFL_HS_OFF_RED (AFL_HS_OFF_ACTS + AFL_HS_MAX_E + 5 * AFL_HS_MAX_H)
#define AFL_HS_FLOATS (AFL_HS_OFF_RED + AFL_HS_MAX_H + 3)

static __device__ __align__(256) float g_hyper_step_scratch[AFL_HS_FLOATS];

static std::atomic<float*> g_hs_addr[64];   // per device: device address of g_hyper_step_scratch (null = not resolved yet);
                                            // concurrent first calls all store the same value

static int hyper_step_scratch(float** out) {
  int dev = 0;
  AFL_CUDA(cudaGetDevice(&dev));
  AFL_CHECK_ARG(dev >= 0 && dev < 64, "afl_hyper_step_noalloc: device ordinal out of range");
  float* cached = g_hs_addr[dev].load(std::memory_order_acquire);
  if (cached == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_hyper_step_scratch);
    if (e != cudaSuccess || p == nullptr) {
      afl_rt::set_error("afl_hyper_step_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                        cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -7;
    }
    g_hs_addr[dev].store(p, std::memory_order_release);
    cached = p;
  }
  *out = cached;
  return 0;
}

int afl_hyper_step_noalloc(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, float* small,
                           int n_nodes, int E, int H, int client, const int* client_dev, const float* theta_client,
                           int64_t d_pad, float lr, float b1, float b2, float eps, int step, float clip,
                           float* coef_norm_out /* [2] device, nullable */, void* stream) {
  if (d_pad > AFL_HS_MAX_DPAD || H > AFL_HS_MAX_H || E > AFL_HS_MAX_E || d_pad < 1 || H < 1 || E < 1) {
    afl_rt::set_error("afl_hyper_step_noalloc: d_pad=%lld H=%d E=%d exceed the built-in scratch (d_pad <= %d, H <= %d, "
                      "E <= %d); call afl_hyper_step with caller-provided scratch instead",
                      (long long)d_pad, H, E, AFL_HS_MAX_DPAD, AFL_HS_MAX_H, AFL_HS_MAX_E);
    return -1;
  }
  float* base = nullptr;
  int rc = hyper_step_scratch(&base);
  if (rc) return rc;
  float* delta = base;
  float* f = base + AFL_HS_OFF_F;
  float* acts = base + AFL_HS_OFF_ACTS;
  float* red = base + AFL_HS_OFF_RED;
  rc = afl_hyper_step(head_w, head_b, mw, vw, mb, vb, small, n_nodes, E, H, client, client_dev, theta_client, delta, f,
                      acts, red, d_pad, lr, b1, b2, eps, step, clip, stream);
  if (rc) return rc;
  if (coef_norm_out)
    AFL_CUDA(cudaMemcpyAsync(coef_norm_out, red + 1 + H, 2 * sizeof(float), cudaMemcpyDeviceToDevice,
                             (cudaStream_t)stream));
  return 0;
}

}  // extern "C"
