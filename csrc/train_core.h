// Op-program interpreter for multi-client local training (forward + backward + fused optimiser).
//
// What it replaces: the reference's per-client Python loop `train_ICU` (/root/reference/client.py:75-111):
// ~525 optimiser steps per client per round, each ~150 ATen/cuDNN/cuBLAS launches, 3 H2D copies
// and 2 host syncs.  Here a whole round of a client is ONE persistent kernel: a model is a short
// list of ops (built in attackfl_b200/models/programs.py) executed by G thread blocks per client,
// each owning tiles of S samples whose activations live in shared memory; parameters, Adam
// state and per-block partial gradients are rows of the flat fp32 arena (L2-resident).
//
// Ops (forward semantics; backward is derived op by op, run in reverse order):
//   LOAD     gather S records of the device-resident data set by index (vitals | labs | label)
//   LINEAR   out = drop_p(act(in0 . W^T + b))         act in {none, relu, gelu}   (T5, T3/T4 pieces)
//   LN       out = LayerNorm(in0 + drop_p(in1)) * g + b   (residual + dropout fused)   (T3/T4)
//   GRU      T=1 bidirectional-GRU cell on precomputed W_ih x + b_ih      (T1)
//   CONV     Conv1d(k=3, pad=1) + ReLU on [C][L] rows                                   (T2)
//   POOL     AdaptiveAvgPool1d(4) (+ dropout)                                           (T2)
//   DROPOUT  grouped Bernoulli dropout (group 16 = attention-weight dropout at T=1)     (T6)
//   BCE      sigmoid + BCELoss, emits dL/dlogit, loss sum and a NaN flag                (T7)
// The optimiser phase (Adam re-created per round / SGD-momentum, T8/T9) reduces the G partial
// gradient rows and updates the owned 1/G slice of the parameter row.
//
// All functions are (tid, nthreads)-parametrised and contain no warp intrinsics, so the same
// source is the CPU twin (tid=0, nthreads=1) used by the non-GPU tests.
#pragma once
#include "common.h"

enum AflOpKind { AFL_OP_LOAD = 1, AFL_OP_LINEAR = 2, AFL_OP_LN = 3, AFL_OP_GRU = 4, AFL_OP_CONV = 5,
                 AFL_OP_POOL = 6, AFL_OP_DROPOUT = 7, AFL_OP_BCE = 8 };
enum AflAct { AFL_ACT_NONE = 0, AFL_ACT_RELU = 1, AFL_ACT_GELU = 2 };
enum AflFlags { AFL_F_ACC0 = 1, AFL_F_ACC1 = 2, AFL_F_NEED_GIN0 = 4, AFL_F_MASK_INPUT = 8 };

#define AFL_OP_WORDS 32
#define AFL_LN_PARTS 8
#define AFL_LN_EPS 1e-5f
#define AFL_LN_ROW (2 * AFL_LN_PARTS + 1)   // per-sample scratch row of LN: 2P partial sums, then rstd
#define AFL_PART_ACT 1     // backward parts: activation gradients (dX / dx) ...
#define AFL_PART_PARAM 2   // ... and parameter gradients (dW, db, dgamma, dbeta, db_hh)

// One instruction; 32 x int32 so Python can pack programs as a plain int32 array.
struct AflOp {
  int32_t kind, act;
  int32_t in0, in0_ld;      // pool offsets are in floats; ld = row (sample) stride
  int32_t in1, in1_ld;
  int32_t out, out_ld;
  int32_t aux, aux2;        // saved tensors for backward (pre-activation / xhat / partials+rstd)
  int32_t w, b, w2;         // parameter offsets inside the flat row (-1 = none)
  int32_t N, K, L;          // out width, in width, conv length / dropout group / pool length
  float p;                  // dropout probability
  int32_t rng;              // per-op RNG stream id
  int32_t g_out, g_in0, g_in1;   // gradient buffers (pool offsets)
  int32_t flags;
  int32_t in_cs, out_cs;    // conv/pool channel strides
  int32_t grp_start, grp_size;   // concurrent-op groups (see afl_run_forward)
  int32_t pad[6];
};

struct AflTile {
  float* pool;              // per-block activation pool (shared memory on the GPU)
  const float* w;           // this client's parameter row
  float* gw;                // this block's partial-gradient row
  const float* data;        // [n_rows, rec] packed records
  const int32_t* idx;       // this step's sample indices for the tile (may be null => consecutive)
  float* probs;             // inference output (global), or null
  float* loss_acc;          // [2]: loss sum, nan count  (per block, in pool tail or global)
  int64_t row0;             // inference: first record index of this tile
  int rec;                  // floats per record (24 for ICU)
  int S;                    // samples per tile
  int n_valid;              // valid samples in this tile
  int accumulate;           // partial grads: += instead of =
  int train;                // 1: dropout active
  float inv_b;              // 1 / (batch size of this optimiser step)
  uint32_t seed_lo, seed_hi, client, step;
  int sample_base;          // index of the tile's first sample inside the step's batch
};

// ----------------------------------------------------------------------------------------- helpers
AFL_DEV float afl_act_fwd(int act, float z) {
  if (act == AFL_ACT_RELU) return z > 0.f ? z : 0.f;
  if (act == AFL_ACT_GELU) return afl_gelu(z);
  return z;
}

AFL_DEV float afl_drop_scale(const AflTile& t, const AflOp& op, int s, uint32_t elem) {
  // returns 0 (dropped) or 1/(1-p); 1 when dropout is inactive
  if (!(t.train && op.p > 0.f)) return 1.f;
  bool keep = afl_keep(t.seed_lo, t.seed_hi, t.client, t.step, (uint32_t)(t.sample_base + s),
                       (uint32_t)op.rng, elem, op.p);
  return keep ? 1.f / (1.f - op.p) : 0.f;
}

AFL_DEV void afl_store(float* dst, float v, int acc) { *dst = acc ? *dst + v : v; }

// Scales for elements elem4..elem4+3 (elem4 % 4 == 0) from ONE Philox call.
AFL_DEV float4 afl_drop_scale4(const AflTile& t, const AflOp& op, int s, uint32_t elem4) {
  float4 r; r.x = r.y = r.z = r.w = 1.f;
  if (!(t.train && op.p > 0.f)) return r;
  const AflU4 u = afl_philox(elem4 >> 2, ((uint32_t)op.rng & 0xFFFFu) | (t.client << 16), t.step,
                             (uint32_t)(t.sample_base + s), t.seed_lo, t.seed_hi);
  const float inv = 1.f / (1.f - op.p), sc = 1.0f / 16777216.0f;
  r.x = ((float)(u.x >> 8) * sc >= op.p) ? inv : 0.f;
  r.y = ((float)(u.y >> 8) * sc >= op.p) ? inv : 0.f;
  r.z = ((float)(u.z >> 8) * sc >= op.p) ? inv : 0.f;
  r.w = ((float)(u.w >> 8) * sc >= op.p) ? inv : 0.f;
  return r;
}

AFL_DEV void afl_store4(float* dst, float4 v, int acc) {
  float4* d = reinterpret_cast<float4*>(dst);
  if (acc) { const float4 o = *d; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
  *d = v;
}

// ----------------------------------------------------------------------------------------- LOAD
AFL_DEV void afl_load_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int nv = op.N, nl = op.K;       // vitals / labs widths
  const int per = nv + nl + 1;
  for (int it = tid; it < t.S * per; it += nt) {
    int s = it / per, f = it - s * per;
    float v = 0.f;
    if (s < t.n_valid) {
      int64_t r = t.idx ? (int64_t)t.idx[s] : t.row0 + s;
      v = t.data[r * t.rec + f];
      if ((op.flags & AFL_F_MASK_INPUT) && f < nv + nl && v == -2.0f) v = 0.f;
    }
    if (f < nv) t.pool[op.out + s * op.out_ld + f] = v;
    else if (f < nv + nl) t.pool[op.aux + s * op.aux2 + (f - nv)] = v;
    else t.pool[op.in1 + s] = v;
  }
}

// ----------------------------------------------------------------------------------------- LINEAR
#define AFL_TS 4   // samples per thread in the forward / dX mappings

// Weight row chunk of 4*Q floats is loaded into registers first (Q independent 128-bit loads in flight => one
// L2 latency per chunk instead of one per 4 k), then swept over the thread's AFL_TS samples from shared memory.
template <int Q, int TS>
AFL_DEV void afl_lin_chunk(const float* wr, const float* x0, int ldx, int ns, float* acc) {
  float4 wv[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) wv[q] = *reinterpret_cast<const float4*>(wr + 4 * q);
#pragma unroll
  for (int j = 0; j < TS; ++j) {
    if (j < ns) {
      float a = acc[j];
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const float4 xv = *reinterpret_cast<const float4*>(x0 + j * ldx + 4 * q);
        a += wv[q].x * xv.x + wv[q].y * xv.y + wv[q].z * xv.z + wv[q].w * xv.w;
      }
      acc[j] = a;
    }
  }
}

template <int TS>
AFL_DEV void afl_linear_fwd_t(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, K = op.K, S = t.S;
  const int SG = (S + TS - 1) / TS;
  const bool vec = (K & 3) == 0 && (op.in0_ld & 3) == 0 && (op.in0 & 3) == 0 && (op.w & 3) == 0;
  for (int it = tid; it < N * SG; it += nt) {
    const int n = it % N, sg = it / N;
    const int s0 = sg * TS;
    const int ns = AFL_MIN(TS, S - s0);
    const float* wr = t.w + op.w + (int64_t)n * K;
    const float* x0 = t.pool + op.in0 + s0 * op.in0_ld;
    const float bias = op.b >= 0 ? t.w[op.b + n] : 0.f;      // issued early: overlaps the weight loads
    float acc[TS];
#pragma unroll
    for (int j = 0; j < TS; ++j) acc[j] = 0.f;
    if (vec) {
      int k = 0;
      for (; k + 64 <= K; k += 64) afl_lin_chunk<16, TS>(wr + k, x0 + k, op.in0_ld, ns, acc);
      for (; k + 16 <= K; k += 16) afl_lin_chunk<4, TS>(wr + k, x0 + k, op.in0_ld, ns, acc);
      for (; k + 4 <= K; k += 4) afl_lin_chunk<1, TS>(wr + k, x0 + k, op.in0_ld, ns, acc);
    } else {
      float wv[8];
      for (int k0 = 0; k0 < K; k0 += 8) {
        const int kn = AFL_MIN(8, K - k0);
#pragma unroll
        for (int q = 0; q < 8; ++q) wv[q] = q < kn ? wr[k0 + q] : 0.f;
#pragma unroll
        for (int j = 0; j < TS; ++j)
          if (j < ns)
#pragma unroll
            for (int q = 0; q < 8; ++q)
              if (q < kn) acc[j] += wv[q] * x0[j * op.in0_ld + k0 + q];
      }
    }
#pragma unroll
    for (int j = 0; j < TS; ++j) {
      const int s = s0 + j;
      if (j >= ns) break;
      const float z = acc[j] + bias;
      if (op.aux >= 0) t.pool[op.aux + s * N + n] = z;
      float y = afl_act_fwd(op.act, z);
      y *= afl_drop_scale(t, op, s, (uint32_t)n);
      t.pool[op.out + s * op.out_ld + n] = y;
    }
  }
}

AFL_DEV void afl_linear_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  // small tiles: 2 samples per thread so that every thread of the block has an item
  if (t.S >= 16) afl_linear_fwd_t<4>(op, t, tid, nt);
  else if (t.S >= 2) afl_linear_fwd_t<2>(op, t, tid, nt);
  else afl_linear_fwd_t<1>(op, t, tid, nt);
}

// Phase A: turn dL/dout into dL/dz in place (dropout mask, activation derivative).
AFL_DEV void afl_linear_bwd_a(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S;
  const bool drop = t.train && op.p > 0.f;
  if (op.act == AFL_ACT_NONE && !drop) return;
  for (int it = tid; it < N * S; it += nt) {
    const int s = it / N, n = it - s * N;
    float g = t.pool[op.g_out + s * op.out_ld + n];
    if (drop) g *= afl_drop_scale(t, op, s, (uint32_t)n);
    if (op.act == AFL_ACT_RELU) {
      if (!(t.pool[op.out + s * op.out_ld + n] > 0.f)) g = 0.f;
    } else if (op.act == AFL_ACT_GELU) {
      g *= afl_gelu_grad(t.pool[op.aux + s * N + n]);
    }
    t.pool[op.g_out + s * op.out_ld + n] = g;
  }
}

// Phase B: dX, dW, db from dz (three independent item ranges, one flat loop).
template <int TS>
AFL_DEV void afl_linear_bwd_b_t(const AflOp& op, const AflTile& t, int tid, int nt, int parts) {
  const int N = op.N, K = op.K, S = t.S;
  const float* dz = t.pool + op.g_out;        // [S][.], row stride = out_ld (same layout as out)
  const int ldz = op.out_ld;
  const float* x = t.pool + op.in0;
  const int ldx = op.in0_ld;
  const int SG = (S + TS - 1) / TS;
  const bool need_dx = (op.flags & AFL_F_NEED_GIN0) != 0 && (parts & AFL_PART_ACT);
  const int n_dx = need_dx ? K * SG : 0;
  const bool nvec = (N & 3) == 0 && (op.g_out & 3) == 0 && (ldz & 3) == 0;
  const bool kvec = (K & 3) == 0 && (op.in0 & 3) == 0 && (ldx & 3) == 0 && (op.w & 3) == 0;
  const bool tile44 = nvec && kvec;                       // 4(n) x 4(k) register tile for dW
  const int KQ = K >> 2;
  const bool do_par = (parts & AFL_PART_PARAM) != 0;
  const int n_dw = !do_par ? 0 : tile44 ? (N >> 2) * KQ : (nvec ? (N >> 2) * K : N * K);
  const int n_db = (do_par && op.b >= 0) ? N : 0;
  const int total = n_dx + n_dw + n_db;
  for (int it = tid; it < total; it += nt) {
    if (it < n_dx) {
      // dX[s][k] = sum_n dz[s][n] W[n][k]; lanes walk k => coalesced W reads; 16 W values preloaded per pass
      const int k = it % K, sg = it / K, s0 = sg * TS;
      const int ns = AFL_MIN(TS, S - s0);
      const float* wc = t.w + op.w + k;
      float acc[TS];
#pragma unroll
      for (int j = 0; j < TS; ++j) acc[j] = 0.f;
      int n = 0;
      if (nvec) {
        for (; n + 16 <= N; n += 16) {
          float wv[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) wv[q] = wc[(int64_t)(n + q) * K];
#pragma unroll
          for (int j = 0; j < TS; ++j) {
            if (j < ns) {
              float a = acc[j];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 d = *reinterpret_cast<const float4*>(dz + (s0 + j) * ldz + n + 4 * q);
                a += d.x * wv[4 * q] + d.y * wv[4 * q + 1] + d.z * wv[4 * q + 2] + d.w * wv[4 * q + 3];
              }
              acc[j] = a;
            }
          }
        }
      }
      for (; n < N; ++n) {
        const float wv = wc[(int64_t)n * K];
#pragma unroll
        for (int j = 0; j < TS; ++j)
          if (j < ns) acc[j] += dz[(s0 + j) * ldz + n] * wv;
      }
#pragma unroll
      for (int j = 0; j < TS; ++j)
        if (j < ns) afl_store(t.pool + op.g_in0 + (s0 + j) * ldx + k, acc[j], op.flags & AFL_F_ACC0);
    } else if (it < n_dx + n_dw) {
      const int j2 = it - n_dx;
      if (tile44) {
        // dW[n0..n0+3][k0..k0+3] = sum_s dz[s][n0..] (x) x[s][k0..]: 2 x 128-bit shared loads per 16 FMAs
        const int kq = j2 % KQ, nq = j2 / KQ;
        float4 a0, a1, a2, a3;
        a0.x = a0.y = a0.z = a0.w = 0.f; a1 = a0; a2 = a0; a3 = a0;
        for (int s = 0; s < S; ++s) {
          const float4 d = *reinterpret_cast<const float4*>(dz + s * ldz + 4 * nq);
          const float4 xv = *reinterpret_cast<const float4*>(x + s * ldx + 4 * kq);
          a0.x += d.x * xv.x; a0.y += d.x * xv.y; a0.z += d.x * xv.z; a0.w += d.x * xv.w;
          a1.x += d.y * xv.x; a1.y += d.y * xv.y; a1.z += d.y * xv.z; a1.w += d.y * xv.w;
          a2.x += d.z * xv.x; a2.y += d.z * xv.y; a2.z += d.z * xv.z; a2.w += d.z * xv.w;
          a3.x += d.w * xv.x; a3.y += d.w * xv.y; a3.z += d.w * xv.z; a3.w += d.w * xv.w;
        }
        float* g = t.gw + op.w + (int64_t)(4 * nq) * K + 4 * kq;
        afl_store4(g, a0, t.accumulate);
        afl_store4(g + K, a1, t.accumulate);
        afl_store4(g + 2 * (int64_t)K, a2, t.accumulate);
        afl_store4(g + 3 * (int64_t)K, a3, t.accumulate);
      } else if (nvec) {
        const int k = j2 % K, nq = j2 / K;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int s = 0; s < S; ++s) {
          const float xv = x[s * ldx + k];
          const float4 d = *reinterpret_cast<const float4*>(dz + s * ldz + nq * 4);
          a0 += d.x * xv; a1 += d.y * xv; a2 += d.z * xv; a3 += d.w * xv;
        }
        float* g = t.gw + op.w + (int64_t)(nq * 4) * K + k;
        afl_store(g, a0, t.accumulate);
        afl_store(g + K, a1, t.accumulate);
        afl_store(g + 2 * (int64_t)K, a2, t.accumulate);
        afl_store(g + 3 * (int64_t)K, a3, t.accumulate);
      } else {
        const int k = j2 % K, n = j2 / K;
        float a0 = 0.f;
        for (int s = 0; s < S; ++s) a0 += dz[s * ldz + n] * x[s * ldx + k];
        afl_store(t.gw + op.w + (int64_t)n * K + k, a0, t.accumulate);
      }
    } else {
      const int n = it - n_dx - n_dw;
      float a = 0.f;
      for (int s = 0; s < S; ++s) a += dz[s * ldz + n];
      afl_store(t.gw + op.b + n, a, t.accumulate);
    }
  }
}

AFL_DEV void afl_linear_bwd_b(const AflOp& op, const AflTile& t, int tid, int nt,
                              int parts = AFL_PART_ACT | AFL_PART_PARAM) {
  if (t.S >= 16) afl_linear_bwd_b_t<4>(op, t, tid, nt, parts);
  else if (t.S >= 2) afl_linear_bwd_b_t<2>(op, t, tid, nt, parts);
  else afl_linear_bwd_b_t<1>(op, t, tid, nt, parts);
}

// ----------------------------------------------------------------------------------------- LAYERNORM
// aux  : xhat [S][N];  aux2 : partials [S][2*P] followed by rstd [S]
AFL_DEV void afl_ln_fwd_a(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S, P = AFL_LN_PARTS;
  const int chunk = (N + P - 1) / P;
  const bool v4 = (chunk & 3) == 0 && ((op.in0 | op.in0_ld | op.out | op.out_ld) & 3) == 0 &&
                  (op.in1 < 0 || ((op.in1 | op.in1_ld) & 3) == 0);
  for (int it = tid; it < S * P; it += nt) {
    const int s = it / P, part = it - s * P;
    const int n0 = part * chunk, n1 = AFL_MIN(N, n0 + chunk);
    float sum = 0.f, sq = 0.f;
    if (v4 && n1 - n0 == chunk) {
      for (int n = n0; n < n1; n += 4) {
        float4 x = *reinterpret_cast<const float4*>(t.pool + op.in0 + s * op.in0_ld + n);
        if (op.in1 >= 0) {
          const float4 r = *reinterpret_cast<const float4*>(t.pool + op.in1 + s * op.in1_ld + n);
          const float4 d = afl_drop_scale4(t, op, s, (uint32_t)n);
          x.x += r.x * d.x; x.y += r.y * d.y; x.z += r.z * d.z; x.w += r.w * d.w;
        }
        *reinterpret_cast<float4*>(t.pool + op.out + s * op.out_ld + n) = x;
        sum += (x.x + x.y) + (x.z + x.w);
        sq += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
      }
    } else {
      for (int n = n0; n < n1; ++n) {
        float x = t.pool[op.in0 + s * op.in0_ld + n];
        if (op.in1 >= 0) x += t.pool[op.in1 + s * op.in1_ld + n] * afl_drop_scale(t, op, s, (uint32_t)n);
        t.pool[op.out + s * op.out_ld + n] = x;
        sum += x; sq += x * x;
      }
    }
    t.pool[op.aux2 + s * AFL_LN_ROW + 2 * part] = sum;
    t.pool[op.aux2 + s * AFL_LN_ROW + 2 * part + 1] = sq;
  }
}

AFL_DEV void afl_ln_fwd_b(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S, P = AFL_LN_PARTS;
  const bool v4 = (N & 3) == 0 && ((op.out | op.out_ld | op.aux | op.w | op.b) & 3) == 0;
  const int W = v4 ? 4 : 1;
  const int NQ = N / W;
  for (int it = tid; it < S * NQ; it += nt) {
    const int s = it / NQ, n = (it - s * NQ) * W;
    float sum = 0.f, sq = 0.f;
#pragma unroll
    for (int q = 0; q < P; ++q) {
      sum += t.pool[op.aux2 + s * AFL_LN_ROW + 2 * q];
      sq += t.pool[op.aux2 + s * AFL_LN_ROW + 2 * q + 1];
    }
    const float mean = sum / (float)N;
    float var = sq / (float)N - mean * mean;
    var = var > 0.f ? var : 0.f;
    const float rstd = 1.0f / sqrtf(var + AFL_LN_EPS);
    if (v4) {
      float4 x = *reinterpret_cast<const float4*>(t.pool + op.out + s * op.out_ld + n);
      const float4 ga = *reinterpret_cast<const float4*>(t.w + op.w + n);
      const float4 be = *reinterpret_cast<const float4*>(t.w + op.b + n);
      float4 xh; xh.x = (x.x - mean) * rstd; xh.y = (x.y - mean) * rstd; xh.z = (x.z - mean) * rstd; xh.w = (x.w - mean) * rstd;
      *reinterpret_cast<float4*>(t.pool + op.aux + s * N + n) = xh;
      x.x = xh.x * ga.x + be.x; x.y = xh.y * ga.y + be.y; x.z = xh.z * ga.z + be.z; x.w = xh.w * ga.w + be.w;
      *reinterpret_cast<float4*>(t.pool + op.out + s * op.out_ld + n) = x;
    } else {
      const float xh = (t.pool[op.out + s * op.out_ld + n] - mean) * rstd;
      t.pool[op.aux + s * N + n] = xh;
      t.pool[op.out + s * op.out_ld + n] = xh * t.w[op.w + n] + t.w[op.b + n];
    }
    if (n == 0) t.pool[op.aux2 + s * AFL_LN_ROW + 2 * P] = rstd;
  }
}

// backward phase A: per-row partial sums of dxhat and dxhat*xhat, plus dgamma / dbeta
AFL_DEV void afl_ln_bwd_a(const AflOp& op, const AflTile& t, int tid, int nt, int parts = AFL_PART_ACT | AFL_PART_PARAM) {
  const int N = op.N, S = t.S, P = AFL_LN_PARTS;
  const int chunk = (N + P - 1) / P;
  const int n_part = (parts & AFL_PART_ACT) ? S * P : 0;
  const int n_par = (parts & AFL_PART_PARAM) ? N : 0;
  for (int it = tid; it < n_part + n_par; it += nt) {
    if (it < n_part) {
      const int s = it / P, part = it - s * P;
      const int n0 = part * chunk, n1 = AFL_MIN(N, n0 + chunk);
      float a = 0.f, b = 0.f;
      for (int n = n0; n < n1; ++n) {
        const float dxh = t.pool[op.g_out + s * op.out_ld + n] * t.w[op.w + n];
        a += dxh; b += dxh * t.pool[op.aux + s * N + n];
      }
      t.pool[op.aux2 + s * AFL_LN_ROW + 2 * part] = a;
      t.pool[op.aux2 + s * AFL_LN_ROW + 2 * part + 1] = b;
    } else {
      const int n = it - n_part;
      float dg = 0.f, db = 0.f;
      for (int s = 0; s < S; ++s) {
        const float g = t.pool[op.g_out + s * op.out_ld + n];
        dg += g * t.pool[op.aux + s * N + n]; db += g;
      }
      afl_store(t.gw + op.w + n, dg, t.accumulate);
      afl_store(t.gw + op.b + n, db, t.accumulate);
    }
  }
}

AFL_DEV void afl_ln_bwd_b(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S, P = AFL_LN_PARTS;
  const bool v4 = (N & 3) == 0 && ((op.g_out | op.out_ld | op.aux | op.g_in0 | op.in0_ld | op.w) & 3) == 0 &&
                  (op.in1 < 0 || ((op.g_in1 | op.in1_ld) & 3) == 0);
  const int W = v4 ? 4 : 1;
  const int NQ = N / W;
  for (int it = tid; it < S * NQ; it += nt) {
    const int s = it / NQ, n = (it - s * NQ) * W;
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int q = 0; q < P; ++q) {
      a += t.pool[op.aux2 + s * AFL_LN_ROW + 2 * q];
      b += t.pool[op.aux2 + s * AFL_LN_ROW + 2 * q + 1];
    }
    const float rstd = t.pool[op.aux2 + s * AFL_LN_ROW + 2 * P];
    const float m1 = a / (float)N, m2 = b / (float)N;
    if (v4) {
      const float4 g = *reinterpret_cast<const float4*>(t.pool + op.g_out + s * op.out_ld + n);
      const float4 ga = *reinterpret_cast<const float4*>(t.w + op.w + n);
      const float4 xh = *reinterpret_cast<const float4*>(t.pool + op.aux + s * N + n);
      float4 dx;
      dx.x = rstd * (g.x * ga.x - m1 - xh.x * m2); dx.y = rstd * (g.y * ga.y - m1 - xh.y * m2);
      dx.z = rstd * (g.z * ga.z - m1 - xh.z * m2); dx.w = rstd * (g.w * ga.w - m1 - xh.w * m2);
      if (op.flags & AFL_F_NEED_GIN0) afl_store4(t.pool + op.g_in0 + s * op.in0_ld + n, dx, op.flags & AFL_F_ACC0);
      if (op.in1 >= 0) {
        const float4 d = afl_drop_scale4(t, op, s, (uint32_t)n);
        float4 r; r.x = dx.x * d.x; r.y = dx.y * d.y; r.z = dx.z * d.z; r.w = dx.w * d.w;
        afl_store4(t.pool + op.g_in1 + s * op.in1_ld + n, r, op.flags & AFL_F_ACC1);
      }
    } else {
      const float dxh = t.pool[op.g_out + s * op.out_ld + n] * t.w[op.w + n];
      const float dx = rstd * (dxh - m1 - t.pool[op.aux + s * N + n] * m2);
      if (op.flags & AFL_F_NEED_GIN0) afl_store(t.pool + op.g_in0 + s * op.in0_ld + n, dx, op.flags & AFL_F_ACC0);
      if (op.in1 >= 0)
        afl_store(t.pool + op.g_in1 + s * op.in1_ld + n, dx * afl_drop_scale(t, op, s, (uint32_t)n), op.flags & AFL_F_ACC1);
    }
  }
}

// ----------------------------------------------------------------------------------------- GRU (T=1, h0=0)
// in0: [S][3H] = W_ih x + b_ih (gate order r,z,n); w2: b_hh [3H]; out: [S][H] at column offset (ld = out_ld).
AFL_DEV void afl_gru_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int H = op.N, S = t.S;
  const float* bhh = t.w + op.w2;
  for (int it = tid; it < S * H; it += nt) {
    const int s = it / H, j = it - s * H;
    float* g = t.pool + op.in0 + s * op.in0_ld;
    const float r = afl_sigmoid(g[j] + bhh[j]);
    const float z = afl_sigmoid(g[H + j] + bhh[H + j]);
    const float n = tanhf(g[2 * H + j] + r * bhh[2 * H + j]);
    g[j] = r; g[H + j] = z; g[2 * H + j] = n;
    t.pool[op.out + s * op.out_ld + j] = (1.f - z) * n;
  }
}

AFL_DEV void afl_gru_bwd_a(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int H = op.N, S = t.S;
  const float* bhh = t.w + op.w2;
  for (int it = tid; it < S * H; it += nt) {
    const int s = it / H, j = it - s * H;
    const float* g = t.pool + op.in0 + s * op.in0_ld;
    const float r = g[j], z = g[H + j], n = g[2 * H + j];
    const float dh = t.pool[op.g_out + s * op.out_ld + j];
    const float dpn = dh * (1.f - z) * (1.f - n * n);
    const float dpr = dpn * bhh[2 * H + j] * r * (1.f - r);
    const float dpz = -dh * n * z * (1.f - z);
    float* gi = t.pool + op.g_in0 + s * op.in0_ld;
    gi[j] = dpr; gi[H + j] = dpz; gi[2 * H + j] = dpn;
  }
}

AFL_DEV void afl_gru_bwd_b(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int H = op.N, S = t.S;
  for (int j = tid; j < 3 * H; j += nt) {
    float a = 0.f;
    for (int s = 0; s < S; ++s) {
      float d = t.pool[op.g_in0 + s * op.in0_ld + j];
      if (j >= 2 * H) d *= t.pool[op.in0 + s * op.in0_ld + (j - 2 * H)];   // d b_hn = dpn * r
      a += d;
    }
    afl_store(t.gw + op.w2 + j, a, t.accumulate);
  }
}

// ----------------------------------------------------------------------------------------- CONV1D k=3 pad=1 + ReLU
// in0: [S][Cin][.] (sample stride in0_ld, channel stride in_cs); out likewise; W [Cout][Cin][3]; N=Cout, K=Cin, L=len
template <int L>
AFL_DEV void afl_conv_fwd_t(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int Co = op.N, Ci = op.K, S = t.S;
  constexpr int CB = 8;                      // input channels per weight chunk: 24 floats = 6 x 128-bit loads in flight
  const bool vec = (Ci % CB) == 0 && (op.w & 3) == 0;
  for (int it = tid; it < S * Co; it += nt) {
    const int co = it % Co, s = it / Co;
    const float* wr = t.w + op.w + (int64_t)co * Ci * 3;
    const float* xs = t.pool + op.in0 + s * op.in0_ld;
    float acc[L];
    const float bias = t.w[op.b + co];
#pragma unroll
    for (int l = 0; l < L; ++l) acc[l] = bias;
    for (int c0 = 0; c0 < Ci; c0 += CB) {
      float wv[3 * CB];
      const int cn = AFL_MIN(CB, Ci - c0);
      if (vec) {
#pragma unroll
        for (int q = 0; q < 3 * CB / 4; ++q) {
          const float4 v = *reinterpret_cast<const float4*>(wr + c0 * 3 + 4 * q);
          wv[4 * q] = v.x; wv[4 * q + 1] = v.y; wv[4 * q + 2] = v.z; wv[4 * q + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int q = 0; q < 3 * CB; ++q) wv[q] = q < 3 * cn ? wr[c0 * 3 + q] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < CB; ++j) {
        if (j < cn) {
          const float w0 = wv[3 * j], w1 = wv[3 * j + 1], w2 = wv[3 * j + 2];
          const float* xr = xs + (c0 + j) * op.in_cs;
          float xm = 0.f, xc = xr[0];
#pragma unroll
          for (int l = 0; l < L; ++l) {
            const float xp = (l + 1 < L) ? xr[l + 1] : 0.f;
            acc[l] += w0 * xm + w1 * xc + w2 * xp;
            xm = xc; xc = xp;
          }
        }
      }
    }
    float* yr = t.pool + op.out + s * op.out_ld + co * op.out_cs;
#pragma unroll
    for (int l = 0; l < L; ++l) yr[l] = acc[l] > 0.f ? acc[l] : 0.f;
  }
}

// Phase A: dz = dy * (y > 0) in place.
AFL_DEV void afl_conv_bwd_a(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int Co = op.N, S = t.S, L = op.L;
  for (int it = tid; it < S * Co * L; it += nt) {
    const int l = it % L, r = it / L, co = r % Co, s = r / Co;
    const int o = s * op.out_ld + co * op.out_cs + l;
    if (!(t.pool[op.out + o] > 0.f)) t.pool[op.g_out + o] = 0.f;
  }
}

template <int L>
AFL_DEV void afl_conv_bwd_b_t(const AflOp& op, const AflTile& t, int tid, int nt, int parts) {
  const int Co = op.N, Ci = op.K, S = t.S;
  const bool need_dx = (op.flags & AFL_F_NEED_GIN0) != 0 && (parts & AFL_PART_ACT);
  const bool do_par = (parts & AFL_PART_PARAM) != 0;
  const int n_dx = need_dx ? S * Ci : 0;
  const int n_dw = do_par ? ((Co + 3) / 4) * Ci : 0;
  const int total = n_dx + n_dw + (do_par ? Co : 0);
  for (int it = tid; it < total; it += nt) {
    if (it < n_dx) {
      // dx[s][ci][j] = sum_co sum_k w[co][ci][k] dz[s][co][j-k+1]
      const int ci = it % Ci, s = it / Ci;
      const float* dzs = t.pool + op.g_out + s * op.out_ld;
      float acc[L];
#pragma unroll
      for (int l = 0; l < L; ++l) acc[l] = 0.f;
      constexpr int CB = 8;                  // output channels per chunk: 24 scalar weight loads in flight
      for (int co0 = 0; co0 < Co; co0 += CB) {
        float wv[3 * CB];
        const int cn = AFL_MIN(CB, Co - co0);
#pragma unroll
        for (int j = 0; j < CB; ++j) {
          const float* wr = t.w + op.w + ((int64_t)(co0 + (j < cn ? j : 0)) * Ci + ci) * 3;
          wv[3 * j] = wr[0]; wv[3 * j + 1] = wr[1]; wv[3 * j + 2] = wr[2];
        }
#pragma unroll
        for (int j = 0; j < CB; ++j) {
          if (j < cn) {
            const float w0 = wv[3 * j], w1 = wv[3 * j + 1], w2 = wv[3 * j + 2];
            const float* dr = dzs + (co0 + j) * op.out_cs;
            float dm = 0.f, dc = dr[0];
#pragma unroll
            for (int l = 0; l < L; ++l) {
              const float dp = (l + 1 < L) ? dr[l + 1] : 0.f;
              acc[l] += w0 * dp + w1 * dc + w2 * dm;
              dm = dc; dc = dp;
            }
          }
        }
      }
      float* gx = t.pool + op.g_in0 + s * op.in0_ld + ci * op.in_cs;
#pragma unroll
      for (int l = 0; l < L; ++l) afl_store(gx + l, acc[l], op.flags & AFL_F_ACC0);
    } else if (it < n_dx + n_dw) {
      // dW[co0..co0+3][ci][k] = sum_s sum_l dz[s][co][l] x[s][ci][l+k-1]: one x row feeds 4 output channels
      const int j2 = it - n_dx;
      const int ci = j2 % Ci, cq = j2 / Ci;
      const int co0 = cq * 4;
      float a[4][3];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q][0] = 0.f; a[q][1] = 0.f; a[q][2] = 0.f; }
      for (int s = 0; s < S; ++s) {
        const float* xr = t.pool + op.in0 + s * op.in0_ld + ci * op.in_cs;
        const float* d0 = t.pool + op.g_out + s * op.out_ld + co0 * op.out_cs;
        float xm = 0.f, xc = xr[0];
#pragma unroll
        for (int l = 0; l < L; ++l) {
          const float xp = (l + 1 < L) ? xr[l + 1] : 0.f;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (co0 + q < Co) {
              const float d = d0[q * op.out_cs + l];
              a[q][0] += d * xm; a[q][1] += d * xc; a[q][2] += d * xp;
            }
          }
          xm = xc; xc = xp;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (co0 + q < Co) {
          float* g = t.gw + op.w + ((int64_t)(co0 + q) * Ci + ci) * 3;
          afl_store(g, a[q][0], t.accumulate); afl_store(g + 1, a[q][1], t.accumulate); afl_store(g + 2, a[q][2], t.accumulate);
        }
      }
    } else {
      const int co = it - n_dx - n_dw;
      float a = 0.f;
      for (int s = 0; s < S; ++s) {
        const float* dr = t.pool + op.g_out + s * op.out_ld + co * op.out_cs;
#pragma unroll
        for (int l = 0; l < L; ++l) a += dr[l];
      }
      afl_store(t.gw + op.b + co, a, t.accumulate);
    }
  }
}

// ----------------------------------------------------------------------------------------- POOL (adaptive avg -> 4) + dropout
AFL_DEV void afl_pool_bounds(int L, int j, int& a, int& b) { a = (j * L) / 4; b = ((j + 1) * L + 3) / 4; }

AFL_DEV void afl_pool_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int C = op.N, S = t.S, L = op.L;
  for (int it = tid; it < S * C; it += nt) {
    const int c = it % C, s = it / C;
    const float* xr = t.pool + op.in0 + s * op.in0_ld + c * op.in_cs;
    const float4 d = afl_drop_scale4(t, op, s, (uint32_t)(c * 4));      // one Philox call for the 4 windows
    const float ds[4] = {d.x, d.y, d.z, d.w};
    float* y = t.pool + op.out + s * op.out_ld + c * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int a, b; afl_pool_bounds(L, j, a, b);
      float sum = 0.f;
      for (int l = a; l < b; ++l) sum += xr[l];
      y[j] = sum / (float)(b - a) * ds[j];
    }
  }
}

AFL_DEV void afl_pool_bwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int C = op.N, S = t.S, L = op.L;
  for (int it = tid; it < S * C; it += nt) {
    const int c = it % C, s = it / C;
    const float4 d = afl_drop_scale4(t, op, s, (uint32_t)(c * 4));
    const float* go = t.pool + op.g_out + s * op.out_ld + c * 4;
    float gj[4] = {go[0] * d.x, go[1] * d.y, go[2] * d.z, go[3] * d.w};
    float* gi = t.pool + op.g_in0 + s * op.in0_ld + c * op.in_cs;
    for (int l = 0; l < L; ++l) gi[l] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int a, b; afl_pool_bounds(L, j, a, b);
      const float v = gj[j] / (float)(b - a);
      for (int l = a; l < b; ++l) gi[l] += v;
    }
  }
}

// ----------------------------------------------------------------------------------------- DROPOUT (grouped)
AFL_DEV void afl_dropout_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S, G = op.L > 0 ? op.L : 1;
  for (int it = tid; it < S * N; it += nt) {
    const int s = it / N, n = it - s * N;
    t.pool[op.out + s * op.out_ld + n] = t.pool[op.in0 + s * op.in0_ld + n] * afl_drop_scale(t, op, s, (uint32_t)(n / G));
  }
}

AFL_DEV void afl_dropout_bwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  const int N = op.N, S = t.S, G = op.L > 0 ? op.L : 1;
  for (int it = tid; it < S * N; it += nt) {
    const int s = it / N, n = it - s * N;
    const float g = t.pool[op.g_out + s * op.out_ld + n] * afl_drop_scale(t, op, s, (uint32_t)(n / G));
    afl_store(t.pool + op.g_in0 + s * op.in0_ld + n, g, op.flags & AFL_F_ACC0);
  }
}

// ----------------------------------------------------------------------------------------- BCE (sigmoid + BCELoss)
// in0: logits [S] (ld 1); in1: labels [S]; g_in0: dL/dlogit; loss_acc[0] += sum loss, loss_acc[1] += #NaN
AFL_DEV void afl_bce(const AflOp& op, const AflTile& t, int tid, int nt) {
  for (int s = tid; s < t.S; s += nt) {
    const float p = afl_sigmoid(t.pool[op.in0 + s * op.in0_ld]);
    const float y = t.pool[op.in1 + s];
    const bool valid = s < t.n_valid;
    if (t.probs) {
      if (valid) t.probs[t.row0 + s] = p;
      continue;
    }
    float g = 0.f;
    if (valid) {
      float lp = logf(p), lq = logf(1.f - p);
      lp = lp < -100.f ? -100.f : lp;
      lq = lq < -100.f ? -100.f : lq;
      const float loss = -(y * lp + (1.f - y) * lq);
      g = (p - y) * t.inv_b;
      // one lane per sample; S is small, serialise through slot s to stay warp-intrinsic free
      t.pool[op.out + s] = loss;
    } else {
      t.pool[op.out + s] = 0.f;
    }
    t.pool[op.g_in0 + s * op.in0_ld] = g;
  }
}

AFL_DEV void afl_bce_reduce(const AflOp& op, const AflTile& t, int tid, int nt) {
  if (t.probs || tid != 0) return;
  float sum = 0.f, bad = 0.f;
  for (int s = 0; s < t.n_valid; ++s) {
    const float l = t.pool[op.out + s];
    if (l != l) bad += 1.f; else sum += l;
  }
  t.loss_acc[0] += sum;
  t.loss_acc[1] += bad;
}

// ----------------------------------------------------------------------------------------- dispatch
AFL_DEV void afl_conv_fwd(const AflOp& op, const AflTile& t, int tid, int nt) {
  if (op.L == 7) afl_conv_fwd_t<7>(op, t, tid, nt);
  else if (op.L == 16) afl_conv_fwd_t<16>(op, t, tid, nt);
}
AFL_DEV void afl_conv_bwd_b(const AflOp& op, const AflTile& t, int tid, int nt, int parts = AFL_PART_ACT | AFL_PART_PARAM) {
  if (op.L == 7) afl_conv_bwd_b_t<7>(op, t, tid, nt, parts);
  else if (op.L == 16) afl_conv_bwd_b_t<16>(op, t, tid, nt, parts);
}

// Ops of one *group* are mutually independent (e.g. the vitals and labs branches); the block's threads are split
// evenly between them so both run in the same barrier interval.  op.grp_size is stored on the first op of a group,
// op.grp_start on every op.  All ops of a group have the same kind (same phase structure).
AFL_DEV void afl_fwd_phases(int kind, const AflOp& op, const AflTile& t, int tid, int nt, bool active) {
  switch (kind) {
    case AFL_OP_LOAD: if (active) afl_load_fwd(op, t, tid, nt); break;
    case AFL_OP_LINEAR: if (active) afl_linear_fwd(op, t, tid, nt); break;
    case AFL_OP_LN: if (active) afl_ln_fwd_a(op, t, tid, nt); AFL_SYNC(); if (active) afl_ln_fwd_b(op, t, tid, nt); break;
    case AFL_OP_GRU: if (active) afl_gru_fwd(op, t, tid, nt); break;
    case AFL_OP_CONV: if (active) afl_conv_fwd(op, t, tid, nt); break;
    case AFL_OP_POOL: if (active) afl_pool_fwd(op, t, tid, nt); break;
    case AFL_OP_DROPOUT: if (active) afl_dropout_fwd(op, t, tid, nt); break;
    case AFL_OP_BCE: if (active) afl_bce(op, t, tid, nt); AFL_SYNC(); if (active) afl_bce_reduce(op, t, tid, nt); break;
    default: break;
  }
}

AFL_DEV void afl_bwd_phases(int kind, const AflOp& op, const AflTile& t, int tid, int nt, bool active) {
  switch (kind) {
    case AFL_OP_LINEAR: if (active) afl_linear_bwd_a(op, t, tid, nt); AFL_SYNC(); if (active) afl_linear_bwd_b(op, t, tid, nt); break;
    case AFL_OP_LN: if (active) afl_ln_bwd_a(op, t, tid, nt); AFL_SYNC(); if (active) afl_ln_bwd_b(op, t, tid, nt); break;
    case AFL_OP_GRU: if (active) afl_gru_bwd_a(op, t, tid, nt); AFL_SYNC(); if (active) afl_gru_bwd_b(op, t, tid, nt); break;
    case AFL_OP_CONV: if (active) afl_conv_bwd_a(op, t, tid, nt); AFL_SYNC(); if (active) afl_conv_bwd_b(op, t, tid, nt); break;
    case AFL_OP_POOL: if (active) afl_pool_bwd(op, t, tid, nt); break;
    case AFL_OP_DROPOUT: if (active) afl_dropout_bwd(op, t, tid, nt); break;
    default: break;   // LOAD / BCE have no backward work (BCE wrote dL/dlogit in forward)
  }
}

// While a group computes, its threads pull the parameters of the group that runs next (forward: following group,
// backward: preceding group) from L2 into L1, 128 bytes per thread-iteration, so that the next phase does not start
// with a cold weight fetch (~600-800 cycles per dependent chunk).  L1 is only invalidated at the per-step barriers.
// Measured on B200 (profiles/train_kernel_experiments.md): this made the step SLOWER (87 -> 102 us), the prefetch
// instructions cost more issue slots than the latency they hide; it is compiled out unless AFL_ENABLE_PREFETCH is set.
AFL_DEV void afl_prefetch_params(const AflOp* ops, int first, int count, const AflTile& t, int tid, int nt) {
#if defined(__CUDA_ARCH__) && defined(AFL_ENABLE_PREFETCH)
  for (int j = 0; j < count; ++j) {
    const AflOp& op = ops[first + j];
    int n_w = 0;
    if (op.kind == AFL_OP_LINEAR) n_w = op.N * op.K;
    else if (op.kind == AFL_OP_CONV) n_w = op.N * op.K * 3;
    else if (op.kind == AFL_OP_LN) n_w = op.N;
    if (op.w >= 0 && n_w > 0)
      for (int l = tid; l * 32 < n_w; l += nt) AFL_PREFETCH_L1(t.w + op.w + l * 32);
    if (op.b >= 0 && tid < (op.N + 31) / 32) AFL_PREFETCH_L1(t.w + op.b + tid * 32);
    if (op.w2 >= 0 && tid < 4) AFL_PREFETCH_L1(t.w + op.w2 + tid * 32);
  }
#else
  (void)ops; (void)first; (void)count; (void)t; (void)tid; (void)nt;
#endif
}

AFL_DEV void afl_run_forward(const AflOp* ops, int n_ops, const AflTile& t, int tid, int nt) {
  for (int i = 0; i < n_ops;) {
    const int g = ops[i].grp_size > 1 ? ops[i].grp_size : 1;
    if (i + g < n_ops) {
      const int ng = ops[i + g].grp_size > 1 ? ops[i + g].grp_size : 1;
      afl_prefetch_params(ops, i + g, ng, t, tid, nt);
    }
    const int per = nt / g > 0 ? nt / g : 1;                 // threads per member (nt = 1 on the CPU twin)
    if (nt >= g) {
      const int sub = tid / per;
      const bool active = sub < g;
      afl_fwd_phases(ops[i].kind, ops[i + (active ? sub : 0)], t, tid - sub * per, per, active);
    } else {
      for (int j = 0; j < g; ++j) afl_fwd_phases(ops[i].kind, ops[i + j], t, tid, nt, true);
    }
    AFL_SYNC();
    i += g;
  }
}

AFL_DEV void afl_run_backward(const AflOp* ops, int n_ops, const AflTile& t, int tid, int nt) {
  for (int i = n_ops - 1; i >= 0;) {
    const int st = ops[i].grp_start;
    const int g = ops[st].grp_size > 1 ? ops[st].grp_size : 1;
    if (st > 0) {
      const int pst = ops[st - 1].grp_start;
      const int pg = ops[pst].grp_size > 1 ? ops[pst].grp_size : 1;
      afl_prefetch_params(ops, pst, pg, t, tid, nt);
    }
    const int per = nt / g > 0 ? nt / g : 1;
    if (nt >= g) {
      const int sub = tid / per;
      const bool active = sub < g;
      afl_bwd_phases(ops[st].kind, ops[st + (active ? sub : 0)], t, tid - sub * per, per, active);
    } else {
      for (int j = 0; j < g; ++j) afl_bwd_phases(ops[st].kind, ops[st + j], t, tid, nt, true);
    }
    AFL_SYNC();
    i = st - 1;
  }
}

// ----------------------------------------------------------------------------------------- chain mode
// Every forward op and every activation-gradient op works on one sample at a time, so a *team* (the warps assigned to
// one sample) can carry that sample through the whole layer chain with team-local synchronisation only: one warp per
// branch while the program is in a 2-member group (vitals | labs), the whole team for the single ops (LOAD, head, BCE).
// Only parameter gradients need all samples; they are computed afterwards in one barrier-free block-wide pass
// (afl_param_grads).  Block mode needs ~45 block-wide barrier intervals per optimiser step; chain mode needs 2.
// ops_s: per-sample copy of the program with every pool offset shifted to that sample (afl_shift_op), so the unchanged
// op functions run on a 1-sample tile.
AFL_DEV void afl_shift_op(const AflOp& src, int s, AflOp& o) {
  o = src;
  if (src.in0 >= 0) o.in0 += s * src.in0_ld;
  if (src.in1 >= 0) o.in1 += s * src.in1_ld;
  if (src.out >= 0) o.out += s * src.out_ld;
  o.g_out += s * src.out_ld;
  o.g_in0 += s * src.in0_ld;
  o.g_in1 += s * src.in1_ld;
  if (src.kind == AFL_OP_LOAD) { if (src.aux >= 0) o.aux += s * src.aux2; }
  else {
    if (src.aux >= 0) o.aux += s * src.N;
    if (src.kind == AFL_OP_LN && src.aux2 >= 0) o.aux2 += s * AFL_LN_ROW;
  }
}

// team barrier: one warp => __syncwarp; several warps => named barrier `bar_id` (1..15) over `team` threads
AFL_DEV void afl_team_sync(int team, int bar_id) {
#if defined(__CUDA_ARCH__)
  if (team <= 32) __syncwarp();
  else asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(team) : "memory");
#else
  (void)team; (void)bar_id;
#endif
}

AFL_DEV void afl_chain_fwd_op(const AflOp& op, const AflTile& t, int tid, int nt, int bar_id) {
  switch (op.kind) {
    case AFL_OP_LOAD: afl_load_fwd(op, t, tid, nt); break;
    case AFL_OP_LINEAR: afl_linear_fwd(op, t, tid, nt); break;
    case AFL_OP_LN: afl_ln_fwd_a(op, t, tid, nt); afl_team_sync(nt, bar_id); afl_ln_fwd_b(op, t, tid, nt); break;
    case AFL_OP_GRU: afl_gru_fwd(op, t, tid, nt); break;
    case AFL_OP_CONV: afl_conv_fwd(op, t, tid, nt); break;
    case AFL_OP_POOL: afl_pool_fwd(op, t, tid, nt); break;
    case AFL_OP_DROPOUT: afl_dropout_fwd(op, t, tid, nt); break;
    case AFL_OP_BCE: afl_bce(op, t, tid, nt); break;          // loss summed later by afl_param_grads
    default: break;
  }
}

AFL_DEV void afl_chain_bwd_op(const AflOp& op, const AflTile& t, int tid, int nt, int bar_id) {
  switch (op.kind) {
    case AFL_OP_LINEAR: afl_linear_bwd_a(op, t, tid, nt); afl_team_sync(nt, bar_id);
                        afl_linear_bwd_b(op, t, tid, nt, AFL_PART_ACT); break;
    case AFL_OP_LN: afl_ln_bwd_a(op, t, tid, nt, AFL_PART_ACT); afl_team_sync(nt, bar_id); afl_ln_bwd_b(op, t, tid, nt); break;
    case AFL_OP_GRU: afl_gru_bwd_a(op, t, tid, nt); break;
    case AFL_OP_CONV: afl_conv_bwd_a(op, t, tid, nt); afl_team_sync(nt, bar_id);
                      afl_conv_bwd_b(op, t, tid, nt, AFL_PART_ACT); break;
    case AFL_OP_POOL: afl_pool_bwd(op, t, tid, nt); break;
    case AFL_OP_DROPOUT: afl_dropout_bwd(op, t, tid, nt); break;
    default: break;
  }
}

// One sample's forward + backward chain, executed by a team of `tw` warps (tw = group size of the branch groups).
// ops: this sample's shifted program; ts: 1-sample tile; wteam: warp index inside the team; lane: lane in the warp.
AFL_DEV void afl_chain_sample(const AflOp* ops, int n_ops, const AflTile& ts, int wteam, int tw, int lane, int lanes,
                              int bar_id) {
  const int team = tw * lanes;
  // ---- forward
  for (int i = 0; i < n_ops;) {
    const int g = ops[i].grp_size > 1 ? ops[i].grp_size : 1;
    if (g > 1 && tw >= g) {
      if (wteam < g) afl_chain_fwd_op(ops[i + wteam], ts, lane, lanes, bar_id);
      const int ng = (i + g < n_ops && ops[i + g].grp_size > 1) ? 1 : 0;
      if (ng) afl_team_sync(lanes, bar_id); else afl_team_sync(team, bar_id);     // join before a single op
    } else {
      for (int j = 0; j < g; ++j) {
        afl_chain_fwd_op(ops[i + j], ts, wteam * lanes + lane, team, bar_id);
        afl_team_sync(team, bar_id);
      }
    }
    i += g;
  }
  // ---- backward (activation gradients only)
  for (int i = n_ops - 1; i >= 0;) {
    const int st = ops[i].grp_start;
    const int g = ops[st].grp_size > 1 ? ops[st].grp_size : 1;
    if (g > 1 && tw >= g) {
      if (wteam < g) afl_chain_bwd_op(ops[st + wteam], ts, lane, lanes, bar_id);
      const int pg = (st > 0 && ops[ops[st - 1].grp_start].grp_size > 1) ? 1 : 0;
      if (pg) afl_team_sync(lanes, bar_id); else afl_team_sync(team, bar_id);
    } else {
      for (int j = g - 1; j >= 0; --j) {
        afl_chain_bwd_op(ops[st + j], ts, wteam * lanes + lane, team, bar_id);
        afl_team_sync(team, bar_id);
      }
    }
    i = st - 1;
  }
}

// Deferred parameter gradients over all S samples of the tile: no barriers, every op's items are independent.
AFL_DEV void afl_param_grads(const AflOp* ops, int n_ops, const AflTile& t, int tid, int nt) {
  for (int i = 0; i < n_ops; ++i) {
    const AflOp& op = ops[i];
    switch (op.kind) {
      case AFL_OP_LINEAR: afl_linear_bwd_b(op, t, tid, nt, AFL_PART_PARAM); break;
      case AFL_OP_LN: afl_ln_bwd_a(op, t, tid, nt, AFL_PART_PARAM); break;
      case AFL_OP_GRU: afl_gru_bwd_b(op, t, tid, nt); break;
      case AFL_OP_CONV: afl_conv_bwd_b(op, t, tid, nt, AFL_PART_PARAM); break;
      case AFL_OP_BCE: afl_bce_reduce(op, t, tid, nt); break;
      default: break;
    }
  }
}

// Chain-mode tile step.  ops_s: [S][n_ops] shifted programs; the block's warps are split into S teams of `tw` warps.
// On the CPU twin (nt = 1) the samples and team members simply run one after the other.
AFL_DEV void afl_run_tile_chain(const AflOp* ops, const AflOp* ops_s, int n_ops, const AflTile& t, int tid, int nt) {
  AflTile ts = t;
  ts.S = 1;
#if defined(__CUDA_ARCH__)
  const int warps = nt >> 5, tw = warps / t.S;          // caller guarantees warps == tw * S, tw >= 1
  const int w = tid >> 5, lane = tid & 31;
  const int s = w / tw, wteam = w - s * tw;
  ts.n_valid = s < t.n_valid ? 1 : 0;
  ts.sample_base = t.sample_base + s;
  ts.idx = t.idx ? t.idx + s : nullptr;
  ts.row0 = t.row0 + s;
  afl_chain_sample(ops_s + (size_t)s * n_ops, n_ops, ts, wteam, tw, lane, 32, 1 + s);
  __syncthreads();
#else
  for (int s = 0; s < t.S; ++s) {
    ts.n_valid = s < t.n_valid ? 1 : 0;
    ts.sample_base = t.sample_base + s;
    ts.idx = t.idx ? t.idx + s : nullptr;
    ts.row0 = t.row0 + s;
    afl_chain_sample(ops_s + (size_t)s * n_ops, n_ops, ts, 0, 1, 0, 1, 0);
  }
#endif
  afl_param_grads(ops, n_ops, t, tid, nt);
  AFL_SYNC();
}

// ----------------------------------------------------------------------------------------- optimiser phase
struct AflOpt {
  int kind;            // 0 = Adam, 1 = SGD-momentum
  float lr, beta1, beta2, eps, momentum;
  float bc1, bc2;      // Adam bias corrections 1-beta^t for this step
  int first_step;      // SGD: momentum buffer initialised with the gradient (torch semantics)
  float clip_coef;     // gradient scale (1 = none)
};

// One 128-bit chunk of the update (stepsz = lr / bc1, isq = 1 / sqrt(bc2)); vv is only touched by Adam.
AFL_DEV void afl_opt_apply4(float4& wv, float4& mv, float4& vv, const float4& g, const AflOpt& o, float stepsz, float isq) {
  if (o.kind == 0) {
#define AFL_UPD(c) { const float gg = g.c * o.clip_coef; mv.c = o.beta1 * mv.c + (1.f - o.beta1) * gg; \
      vv.c = o.beta2 * vv.c + (1.f - o.beta2) * gg * gg; wv.c -= stepsz * (mv.c / (sqrtf(vv.c) * isq + o.eps)); }
    AFL_UPD(x) AFL_UPD(y) AFL_UPD(z) AFL_UPD(w)
#undef AFL_UPD
  } else {
#define AFL_UPD(c) { const float gg = g.c * o.clip_coef; mv.c = o.first_step ? gg : o.momentum * mv.c + gg; wv.c -= o.lr * mv.c; }
    AFL_UPD(x) AFL_UPD(y) AFL_UPD(z) AFL_UPD(w)
#undef AFL_UPD
  }
}

// Reduce G partial rows over [lo, hi) and apply the update.  gpart: [G][d_pad].
AFL_DEV void afl_reduce_update(float* w, float* m, float* v, const float* gpart, int G, int64_t d_pad,
                               int64_t lo, int64_t hi, const AflOpt& o, int tid, int nt) {
  const float stepsz = o.lr / o.bc1, isq = 1.0f / sqrtf(o.bc2);
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 g; g.x = g.y = g.z = g.w = 0.f;
    int q = 0;
    for (; q + 4 <= G; q += 4) {          // 4 independent 128-bit loads in flight
      const float4 p0 = *reinterpret_cast<const float4*>(gpart + (int64_t)q * d_pad + i);
      const float4 p1 = *reinterpret_cast<const float4*>(gpart + (int64_t)(q + 1) * d_pad + i);
      const float4 p2 = *reinterpret_cast<const float4*>(gpart + (int64_t)(q + 2) * d_pad + i);
      const float4 p3 = *reinterpret_cast<const float4*>(gpart + (int64_t)(q + 3) * d_pad + i);
      g.x += (p0.x + p1.x) + (p2.x + p3.x); g.y += (p0.y + p1.y) + (p2.y + p3.y);
      g.z += (p0.z + p1.z) + (p2.z + p3.z); g.w += (p0.w + p1.w) + (p2.w + p3.w);
    }
    for (; q < G; ++q) {
      const float4 p0 = *reinterpret_cast<const float4*>(gpart + (int64_t)q * d_pad + i);
      g.x += p0.x; g.y += p0.y; g.z += p0.z; g.w += p0.w;
    }
    float4 mv = *reinterpret_cast<const float4*>(m + i);
    float4 wv = *reinterpret_cast<const float4*>(w + i);
    float4 vv; vv.x = vv.y = vv.z = vv.w = 0.f;
    if (o.kind == 0) vv = *reinterpret_cast<const float4*>(v + i);
    afl_opt_apply4(wv, mv, vv, g, o, stepsz, isq);
    if (o.kind == 0) *reinterpret_cast<float4*>(v + i) = vv;
    *reinterpret_cast<float4*>(m + i) = mv;
    *reinterpret_cast<float4*>(w + i) = wv;
  }
}

// Sum of squares of the reduced gradient over [lo, hi) (for clip-mode "proper"); returns thread partial.
AFL_DEV float afl_reduce_sqnorm(const float* gpart, int G, int64_t d_pad, int64_t lo, int64_t hi, int tid, int nt) {
  float acc = 0.f;
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 g; g.x = g.y = g.z = g.w = 0.f;
    for (int q = 0; q < G; ++q) {
      const float4 p0 = *reinterpret_cast<const float4*>(gpart + (int64_t)q * d_pad + i);
      g.x += p0.x; g.y += p0.y; g.z += p0.z; g.w += p0.w;
    }
    acc += (g.x * g.x + g.y * g.y) + (g.z * g.z + g.w * g.w);
  }
  return acc;
}
