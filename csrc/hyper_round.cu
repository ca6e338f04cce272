// Hyper-network round (K5a / K5b), persistent and row-sharded (sm_100a).
//
// What it replaces: the reference's sequential pFedHN update `Server.train_hyper` (/root/reference/server.py:637-678):
// per reporting client one hnet forward, `autograd.grad` with grad_outputs = theta - theta_client, `clip_grad_norm_`,
// `Adam.step` over ~100 x D parameters, then the re-emission of every client's weights (:674-678, :589-590;
// /root/reference/src/Model.py:285-304).  With the heads fused (theta = W_h f + b_h) the gradient of the
// head is rank-1 (delta f^T, delta) and the clip norm is closed-form.
//
// afl_hyper_round: ONE cooperative launch per rank walks all reporting clients in order.  The head W_h [D, H], b_h and
// their Adam moments are row-sharded: rank r owns rows [lo, hi) and nothing else of them.  Per client:
//   pass A  delta = W_h f + b_h - theta_client on the owned rows (theta_client is read where it lives: the pointer may be
//           peer-mapped memory of the client's owner rank), partial |delta|^2 and W_h^T delta;
//   reduce  grid barrier, then the H + 1 partial sums are exchanged between ranks through a symmetric buffer with
//           per-step flags (st.release.sys / ld.acquire.sys, bounded spins) and summed in rank order on every rank;
//   small   every block redoes the 3-layer MLP backward on ~20 K replicated parameters (cheaper than a broadcast), gets the
//           closed-form clip coefficient, and Adam-updates its share of the replicated small parameters (identical on
//           every rank: torch's Adam is dense, the whole embedding table moves every step);
//   pass B  fused Adam on the owned rows of W_h / b_h from delta, f and the coefficient;
//   grid barrier (the next client's features read the updated small parameters).
// afl_hyper_emit_rows: theta_j = W_h f_j + b_h for n clients on the owned rows, stored straight into each client's row
// (pointer table; peer-mapped pointers make it the scatter of the personalised models, server.py:589-590).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../include/attackfl_b200.h"
#include "flat_core.h"
#include "runtime.h"

namespace {

constexpr int kThreads = 256;          // emit / feature kernels
constexpr int kRT = 1024;              // round kernel: one block per SM, 32 warps streaming rows, one grid barrier per 148 blocks
constexpr int kTR = 64;
constexpr int kMaxWorld = 16;
constexpr long long kSpinTimeout = 4000000000LL;

struct HRArgs {
  float* head_w; float* head_b; float* mw; float* vw; float* mb; float* vb;   // shards: row 0 = global row lo
  int64_t rows, lo;
  float* small; float* small_alt; int n_nodes, E, H, tot;   // small: params | m | v; small_alt: [tot] second copy of params
  const int* order; const float* const* theta; int n;
  float* delta;                 // [rows] scratch
  float* red;                   // [2][H + 1] local partial sums, double-buffered by step parity (zeroed by the host)
  float* xchg[kMaxWorld];       // symmetric: [n_max][world][H + 1]
  uint32_t* sig[kMaxWorld];     // symmetric: [n_max][world]
  int world, rank, n_max; uint32_t epoch;
  uint32_t* bar;                // local: [0] grid-barrier counter (monotonic within a launch, zeroed by the host), [1] err
  int step0; float lr, b1, b2, eps, clip;
  float* norms;                 // [n] pre-clip gradient norms (nullable)
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ bool grid_sync(uint32_t* bar, uint32_t target) {
  __shared__ int s_ok;
  __syncthreads();
  if (threadIdx.x == 0) {
    int ok = 1;
    __threadfence();
    atomicAdd(bar, 1u);
    const long long t0 = clock64();
    while (ld_acquire_gpu(bar) < target) {
      if (clock64() - t0 > kSpinTimeout) { bar[1] = 1u; ok = 0; break; }
    }
    s_ok = ok;
  }
  __syncthreads();
  return s_ok != 0;
}

__device__ __forceinline__ int off_w(int n_nodes, int E, int H, int layer) {
  int o = n_nodes * E;
  if (layer == 0) return o;
  o += H * E + H;
  if (layer == 1) return o;
  return o + H * H + H;
}

// <W[d], f> with lane l holding W[d][l + 32 j] and f[l + 32 j]: used by pass A of the round kernel AND by the emit kernel, so
// that theta emitted for a client and theta recomputed in its step are bit-identical (a coordinate the client did not
// change then has delta == 0 exactly, as in the reference, instead of rounding noise that Adam would normalise to +-lr).
template <int J>
__device__ __forceinline__ float row_dot(const float* wv, const float* fl) {
  float v = 0.f;
#pragma unroll
  for (int j = 0; j < J; ++j) v += wv[j] * fl[j];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kRT) hyper_round_kernel(HRArgs a) {
  extern __shared__ float sm[];
  const int E = a.E, H = a.H, tid = threadIdx.x;
  float* e = sm;                    // [E]          embedding of the client
  float* z0 = e + E;                // [H] pre-activations / activations of the MLP
  float* a0 = z0 + H;
  float* z1 = a0 + H;
  float* a1 = z1 + H;
  float* f = a1 + H;                // [H] features
  float* g2 = f + H;                // [H] dL/df
  float* g1 = g2 + H;               // [H]
  float* g0 = g1 + H;               // [H]
  float* ge = g0 + H;               // [E]
  float* dl = ge + E;               // [kTR]
  float* sred = dl + kTR;           // [H + 1] block-local partial sums
  // MLP weights of the current step staged in shared memory (coalesced load once per step; the matvecs below would
  // otherwise walk ~20 K parameters in L2 with a 400-byte stride between threads): W0 [H][E] b0 | W1 [H][H+1] b1 | W2 [H][H+1] b2
  float* wsm = sred + (H + 1);
  const int so1 = H * E + H, so2 = so1 + H * (H + 1) + H;
  __shared__ float s_coef;
  float* Mm = a.small + a.tot;
  float* Vv = a.small + 2 * (int64_t)a.tot;
  uint32_t bar_n = 0;

  for (int k = 0; k < a.n; ++k) {
    const int c = a.order[k];
    const float* theta = a.theta[k] + a.lo;
    float* red = a.red + (k & 1) * (H + 1);
    const int t_adam = a.step0 + k + 1;
    // The replicated small parameters are double-buffered: a fast block may already write step k's update while a slow
    // block still reads the parameters of step k in its MLP backward (there is no barrier in between).
    const float* P = (k & 1) ? a.small_alt : a.small;
    float* Pn = (k & 1) ? a.small : a.small_alt;
    const float bc1 = 1.f - powf(a.b1, (float)t_adam), bc2 = 1.f - powf(a.b2, (float)t_adam);

    // ---- features (every block, redundantly): e -> z0/a0 -> z1/a1 -> f
    for (int i = tid; i < E; i += kRT) e[i] = __ldcg(P + c * E + i);
    for (int i = tid; i <= H; i += kRT) sred[i] = 0.f;
    {
      const float* P0 = P + off_w(a.n_nodes, E, H, 0);
      for (int i = tid; i < H * E + H; i += kRT) wsm[i] = __ldcg(P0 + i);
      const float* P1 = P + off_w(a.n_nodes, E, H, 1);
      const float* P2 = P + off_w(a.n_nodes, E, H, 2);
      for (int i = tid; i < H * H; i += kRT) {
        const int j = i / H, q = i - j * H;
        wsm[so1 + j * (H + 1) + q] = __ldcg(P1 + i);
        wsm[so2 + j * (H + 1) + q] = __ldcg(P2 + i);
      }
      for (int i = tid; i < H; i += kRT) {
        wsm[so1 + H * (H + 1) + i] = __ldcg(P1 + H * H + i);
        wsm[so2 + H * (H + 1) + i] = __ldcg(P2 + H * H + i);
      }
    }
    __syncthreads();
    for (int layer = 0; layer < 3; ++layer) {
      const int in = layer == 0 ? E : H, ldw = layer == 0 ? E : H + 1;
      const float* x = layer == 0 ? e : (layer == 1 ? a0 : a1);
      const float* W = wsm + (layer == 0 ? 0 : (layer == 1 ? so1 : so2));
      const float* B = W + H * ldw;
      for (int j = tid; j < H; j += kRT) {
        float acc = B[j];
        for (int i = 0; i < in; ++i) acc += W[j * ldw + i] * x[i];
        if (layer == 0) { z0[j] = acc; a0[j] = acc > 0.f ? acc : 0.f; }
        else if (layer == 1) { z1[j] = acc; a1[j] = acc > 0.f ? acc : 0.f; }
        else f[j] = acc;
      }
      __syncthreads();
    }

    // ---- pass A: delta on the owned rows, partial |delta|^2 and W^T delta.  One warp per row: lane l holds W[d][l + 32 j],
    // the dot product is a shuffle reduction and the same registers feed W^T delta, so W_h is read exactly once, fully
    // coalesced (a row is H contiguous floats), four rows in flight per warp.
    {
      const int lane = tid & 31, wib = tid >> 5, nwarp = kRT / 32;
      const int64_t gw = (int64_t)blockIdx.x * nwarp + wib, nw = (int64_t)gridDim.x * nwarp;
      constexpr int J = 4;                                  // H <= 128
      float fl[J], acc[J];
#pragma unroll
      for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; fl[j] = h < H ? f[h] : 0.f; acc[j] = 0.f; }
      float acc_sq = 0.f;
      for (int64_t d = gw; d < a.rows; d += 4 * nw) {
        float wv[4][J];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t dd = d + u * nw;
#pragma unroll
          for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; wv[u][j] = (dd < a.rows && h < H) ? a.head_w[dd * H + h] : 0.f; }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t dd = d + u * nw;
          if (dd >= a.rows) break;
          const float v = row_dot<J>(wv[u], fl);
          const float dlt = (v + a.head_b[dd]) - theta[dd];
          if (lane == 0) { a.delta[dd] = dlt; acc_sq += dlt * dlt; }
#pragma unroll
          for (int j = 0; j < J; ++j) acc[j] += wv[u][j] * dlt;
        }
      }
      __syncthreads();
      if (lane == 0) atomicAdd(sred, acc_sq);
#pragma unroll
      for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; if (h < H) atomicAdd(sred + 1 + h, acc[j]); }
      __syncthreads();
      if (tid <= H) atomicAdd(red + tid, sred[tid]);
    }
    if (!grid_sync(a.bar, ++bar_n * gridDim.x)) return;

    // ---- cross-rank sum of the H + 1 scalars (identical order on every rank)
    if (a.world > 1) {
      if (blockIdx.x == 0) {
        if (tid <= H) {
          const float v = __ldcg(red + tid);
          for (int p = 0; p < a.world; ++p) a.xchg[p][((int64_t)k * a.world + a.rank) * (H + 1) + tid] = v;
        }
        __threadfence_system();
        __syncthreads();
        if (tid < a.world) st_release_sys(a.sig[tid] + (int64_t)k * a.world + a.rank, a.epoch);
      }
      __shared__ int s_ok;
      if (tid == 0) {
        int ok = 1;
        const long long t0 = clock64();
        for (int p = 0; p < a.world && ok; ++p)
          while (ld_acquire_sys(a.sig[a.rank] + (int64_t)k * a.world + p) < a.epoch)
            if (clock64() - t0 > kSpinTimeout) { a.bar[1] = 1u; ok = 0; break; }
        s_ok = ok;
      }
      __syncthreads();
      if (!s_ok) return;
      if (tid <= H) {
        float v = 0.f;
        for (int p = 0; p < a.world; ++p) v += __ldcg(a.xchg[a.rank] + ((int64_t)k * a.world + p) * (H + 1) + tid);
        sred[tid] = v;
      }
    } else if (tid <= H) {
      sred[tid] = __ldcg(red + tid);
    }
    __syncthreads();
    // the other parity's accumulator is free again: clear it for the next client
    if (blockIdx.x == 0 && tid <= H) a.red[((k + 1) & 1) * (H + 1) + tid] = 0.f;

    // ---- MLP backward (every block): g2 = W_h^T delta, g1, g0, ge; closed-form norm and clip coefficient
    if (tid < H) g2[tid] = sred[1 + tid];
    __syncthreads();
    {
      const float* W2 = wsm + so2;
      for (int i = tid; i < H; i += kRT) {
        float v = 0.f;
        for (int j = 0; j < H; ++j) v += W2[j * (H + 1) + i] * g2[j];
        g1[i] = z1[i] > 0.f ? v : 0.f;
      }
      __syncthreads();
      const float* W1 = wsm + so1;
      for (int i = tid; i < H; i += kRT) {
        float v = 0.f;
        for (int j = 0; j < H; ++j) v += W1[j * (H + 1) + i] * g1[j];
        g0[i] = z0[i] > 0.f ? v : 0.f;
      }
      __syncthreads();
      const float* W0 = wsm;
      for (int i = tid; i < E; i += kRT) {
        float v = 0.f;
        for (int j = 0; j < H; ++j) v += W0[j * E + i] * g0[j];
        ge[i] = v;
      }
      __syncthreads();
    }
    {
      // |grad|^2 = |delta|^2 (|f|^2 + 1) + sum_l |g_l|^2 (|x_l|^2 + 1) + |g_e|^2     (every gradient is an outer product)
      double part = 0.0;
      if (tid == 0) {
        double f2 = 0, a12 = 0, a02 = 0, e2 = 0, g22 = 0, g12 = 0, g02 = 0, ge2 = 0;
        for (int i = 0; i < H; ++i) {
          f2 += (double)f[i] * f[i]; a12 += (double)a1[i] * a1[i]; a02 += (double)a0[i] * a0[i];
          g22 += (double)g2[i] * g2[i]; g12 += (double)g1[i] * g1[i]; g02 += (double)g0[i] * g0[i];
        }
        for (int i = 0; i < E; ++i) { e2 += (double)e[i] * e[i]; ge2 += (double)ge[i] * ge[i]; }
        part = (double)sred[0] * (f2 + 1.0) + g22 * (a12 + 1.0) + g12 * (a02 + 1.0) + g02 * (e2 + 1.0) + ge2;
        const float norm = (float)sqrt(part);
        float coef = 1.f;
        if (a.clip > 0.f) { coef = a.clip / (norm + 1e-6f); coef = coef < 1.f ? coef : 1.f; }
        s_coef = coef;
        if (blockIdx.x == 0 && a.norms) a.norms[k] = norm;
      }
      __syncthreads();
    }
    const float coef = s_coef;
    const float stepsz = a.lr / bc1, isq = 1.0f / sqrtf(bc2);

    // ---- Adam on the replicated small parameters: the blocks share the index range (dense: every entry moves)
    {
      const int o0 = off_w(a.n_nodes, E, H, 0), o1 = off_w(a.n_nodes, E, H, 1), o2 = off_w(a.n_nodes, E, H, 2);
      for (int i = blockIdx.x * kRT + tid; i < a.tot; i += gridDim.x * kRT) {
        float g = 0.f;
        if (i < o0) { const int r = i / E; g = r == c ? ge[i - r * E] : 0.f; }
        else if (i < o1) { const int q = i - o0; g = q < H * E ? g0[q / E] * e[q % E] : g0[q - H * E]; }
        else if (i < o2) { const int q = i - o1; g = q < H * H ? g1[q / H] * a0[q % H] : g1[q - H * H]; }
        else { const int q = i - o2; g = q < H * H ? g2[q / H] * a1[q % H] : g2[q - H * H]; }
        g *= coef;
        const float mi = a.b1 * Mm[i] + (1.f - a.b1) * g;
        const float vi = a.b2 * Vv[i] + (1.f - a.b2) * g * g;
        Mm[i] = mi; Vv[i] = vi;
        Pn[i] = __ldcg(P + i) - stepsz * (mi / (sqrtf(vi) * isq + a.eps));
      }
    }

    // ---- pass B: Adam on the owned rows of W_h / b_h (same row -> warp map as pass A: no cross-block dependency on W_h)
    {
      const int lane = tid & 31, wib = tid >> 5, nwarp = kRT / 32;
      const int64_t gw = (int64_t)blockIdx.x * nwarp + wib, nw = (int64_t)gridDim.x * nwarp;
      constexpr int J = 4;
      float fl[J];
#pragma unroll
      for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; fl[j] = h < H ? f[h] : 0.f; }
      for (int64_t d = gw; d < a.rows; d += 2 * nw) {
        // every load of the two rows (W, m, v: 24 independent 128-byte warp loads) is issued before the first store: the
        // three arrays may alias as far as the compiler knows, so loads written after a store would wait for it and the
        // pass would run at the latency of one row at a time (measured 1.75 TB/s for the whole kernel before)
        float wr[2][J], mr[2][J], vr[2][J], cdr[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int64_t dd = d + u * nw;
          cdr[u] = dd < a.rows ? coef * a.delta[dd] : 0.f;
#pragma unroll
          for (int j = 0; j < J; ++j) {
            const int h = lane + 32 * j;
            const bool ok = dd < a.rows && h < H;
            const int64_t it = dd * H + h;
            wr[u][j] = ok ? a.head_w[it] : 0.f;
            mr[u][j] = ok ? a.mw[it] : 0.f;
            vr[u][j] = ok ? a.vw[it] : 0.f;
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int64_t dd = d + u * nw;
          if (dd >= a.rows) break;
          const float cd = cdr[u];
#pragma unroll
          for (int j = 0; j < J; ++j) {
            const int h = lane + 32 * j;
            if (h < H) {
              const int64_t it = dd * H + h;
              const float g = cd * fl[j];
              const float mi = a.b1 * mr[u][j] + (1.f - a.b1) * g;
              const float vi = a.b2 * vr[u][j] + (1.f - a.b2) * g * g;
              a.mw[it] = mi; a.vw[it] = vi;
              a.head_w[it] = wr[u][j] - stepsz * (mi / (sqrtf(vi) * isq + a.eps));
            }
          }
          if (lane == 0) {
            const float mi = a.b1 * a.mb[dd] + (1.f - a.b1) * cd;
            const float vi = a.b2 * a.vb[dd] + (1.f - a.b2) * cd * cd;
            a.mb[dd] = mi; a.vb[dd] = vi;
            a.head_b[dd] -= stepsz * (mi / (sqrtf(vi) * isq + a.eps));
          }
        }
      }
    }
    if (!grid_sync(a.bar, ++bar_n * gridDim.x)) return;
  }
  if (a.n & 1)   // the current parameters ended in the second copy
    for (int i = blockIdx.x * kRT + tid; i < a.tot; i += gridDim.x * kRT) a.small[i] = __ldcg(a.small_alt + i);
}

// f [n][H] for n clients: one block per client (tiny)
__global__ void __launch_bounds__(kThreads) hyper_feat_kernel(const float* small, int n_nodes, int E, int H,
                                                              const int* clients, int n, float* f_out) {
  extern __shared__ float sm[];
  float* x = sm; float* y = sm + AFL_MAX(H, E);
  for (int q = blockIdx.x; q < n; q += gridDim.x) {
    const int c = clients[q];
    __syncthreads();
    for (int i = threadIdx.x; i < E; i += kThreads) x[i] = small[c * E + i];
    __syncthreads();
    for (int layer = 0; layer < 3; ++layer) {
      const int in = layer == 0 ? E : H;
      const float* W = small + off_w(n_nodes, E, H, layer);
      const float* B = W + H * in;
      for (int j = threadIdx.x; j < H; j += kThreads) {
        float acc = B[j];
        for (int i = 0; i < in; ++i) acc += W[j * in + i] * x[i];
        y[j] = acc;
      }
      __syncthreads();
      for (int j = threadIdx.x; j < H; j += kThreads) {
        if (layer < 2) x[j] = y[j] > 0.f ? y[j] : 0.f;
        else f_out[q * H + j] = y[j];
      }
      __syncthreads();
    }
  }
}

// theta_j[lo + d] = W_h[d] . f_j + b_h[d] for the owned rows d, written through the pointer table.  One warp per row, the
// row is loaded once and reused for all n clients; same arithmetic as pass A of the round kernel (row_dot).
__global__ void __launch_bounds__(kThreads) hyper_emit_rows_kernel(const float* head_w, const float* head_b, int64_t rows,
                                                                   int64_t lo, const float* f_all, int n, int H,
                                                                   float* const* theta) {
  extern __shared__ float sm[];
  float* fs = sm;                       // [n][H]
  for (int i = threadIdx.x; i < n * H; i += kThreads) fs[i] = f_all[i];
  __syncthreads();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nwarp = kThreads / 32;
  const int64_t gw = (int64_t)blockIdx.x * nwarp + wib, nw = (int64_t)gridDim.x * nwarp;
  constexpr int J = 4;
  for (int64_t d = gw; d < rows; d += nw) {
    float wv[J];
#pragma unroll
    for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; wv[j] = h < H ? head_w[d * H + h] : 0.f; }
    const float b = head_b[d];
    for (int c = 0; c < n; ++c) {
      float fl[J];
#pragma unroll
      for (int j = 0; j < J; ++j) { const int h = lane + 32 * j; fl[j] = h < H ? fs[c * H + h] : 0.f; }
      const float v = row_dot<J>(wv, fl);
      if (lane == 0) theta[c][lo + d] = v + b;
    }
  }
}

}  // namespace

extern "C" {

/* Floats of the per-rank symmetric exchange buffer / words of the flag buffer for up to n_max clients per round. */
int64_t afl_hyper_round_xchg_floats(int n_max, int world, int H) { return (int64_t)n_max * world * (H + 1); }
int64_t afl_hyper_round_sig_words(int n_max, int world) { return (int64_t)n_max * world; }

int afl_hyper_round(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, int64_t rows, int64_t lo,
                    float* small, float* small_alt /* [tot] */, int n_nodes, int E, int H, const int* order_dev,
                    const float* const* theta_dev, int n, float* delta, float* red /* [2][H+1] */, float* const* xchg, uint32_t* const* sig, int world, int rank,
                    int n_max, uint32_t epoch, uint32_t* bar /* [2] */, int step0, float lr, float b1, float b2, float eps,
                    float clip, float* norms, void* stream) {
  AFL_CHECK_ARG(world >= 1 && world <= kMaxWorld, "afl_hyper_round: world size must be 1..16");
  AFL_CHECK_ARG(n >= 0 && n <= n_max, "afl_hyper_round: more clients than the exchange buffer holds");
  AFL_CHECK_ARG(H <= 128 && H >= 1 && E >= 1, "afl_hyper_round: hidden width must be <= 128");
  if (n == 0) return 0;
  HRArgs a;
  a.head_w = head_w; a.head_b = head_b; a.mw = mw; a.vw = vw; a.mb = mb; a.vb = vb; a.rows = rows; a.lo = lo;
  a.small = small; a.small_alt = small_alt; a.n_nodes = n_nodes; a.E = E; a.H = H;
  a.tot = n_nodes * E + H * E + H + 2 * (H * H + H);
  a.order = order_dev; a.theta = theta_dev; a.n = n; a.delta = delta; a.red = red;
  for (int r = 0; r < kMaxWorld; ++r) {
    a.xchg[r] = (world > 1 && r < world) ? xchg[r] : nullptr;
    a.sig[r] = (world > 1 && r < world) ? sig[r] : nullptr;
  }
  a.world = world; a.rank = rank; a.n_max = n_max; a.epoch = epoch; a.bar = bar; a.step0 = step0;
  a.lr = lr; a.b1 = b1; a.b2 = b2; a.eps = eps; a.clip = clip; a.norms = norms;
  cudaStream_t st = (cudaStream_t)stream;
  AFL_CUDA(cudaMemsetAsync(red, 0, sizeof(float) * 2 * (size_t)(H + 1), st));
  AFL_CUDA(cudaMemsetAsync(bar, 0, sizeof(uint32_t) * 2, st));
  const size_t smem = sizeof(float) * (size_t)(2 * E + 9 * H + kTR + (H + 1) + H * E + H + 2 * (H * (H + 1) + H));
  AFL_CUDA(cudaFuncSetAttribute(hyper_round_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hyper_round_kernel, kRT, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  if (per_sm > 2) per_sm = 2;
  const int64_t tiles = (rows + 127) / 128;               // at least 4 rows per warp before a block is worth launching
  int64_t cap = (int64_t)per_sm * afl_rt::num_sms();
  if (const char* ev = getenv("ATTACKFL_FUSED_MAX_BLOCKS")) { const int64_t m = atoi(ev); if (m >= 1 && m < cap) cap = m; }   // tests: emulated ranks on one GPU
  int grid = (int)(tiles < cap ? tiles : cap);
  if (grid < 1) grid = 1;
  void* args[] = {&a};
  AFL_CUDA(cudaLaunchCooperativeKernel((const void*)hyper_round_kernel, dim3(grid), dim3(kRT), args, smem, st));
  afl_rt::count_launch();
  return afl_rt::check_last("afl_hyper_round");
}

int afl_hyper_emit_rows(const float* head_w, const float* head_b, int64_t rows, int64_t lo, const float* small, int n_nodes,
                        int E, int H, const int* clients_dev, int n, float* f_scratch /* [n*H] */,
                        float* const* theta_dev, void* stream) {
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  hyper_feat_kernel<<<n, kThreads, sizeof(float) * 2 * (size_t)AFL_MAX(H, E), st>>>(small, n_nodes, E, H, clients_dev, n, f_scratch);
  const size_t smem = sizeof(float) * (size_t)n * H;
  AFL_CHECK_ARG(smem <= 200 * 1024 && H <= 128, "afl_hyper_emit_rows: too many clients per call for shared memory (chunk them)");
  AFL_CUDA(cudaFuncSetAttribute(hyper_emit_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t tiles = (rows + 7) / 8;
  int64_t cap = 4LL * afl_rt::num_sms();
  int grid = (int)(tiles < cap ? tiles : cap);
  if (grid < 1) grid = 1;
  hyper_emit_rows_kernel<<<grid, kThreads, smem, st>>>(head_w, head_b, rows, lo, f_scratch, n, H, theta_dev);
  afl_rt::count_launch(2);
  return afl_rt::check_last("afl_hyper_emit_rows");
}

}  // extern "C"
