// Flat-vector kernels shared by the CUDA build and the CPU twin.
//
//   K1/K2  afl_fedavg_core      sample-weighted reduce (+ per-attacker shifted moments, LIE rows) over a
//                               coordinate range; the caller owns the range (whole vector on one GPU, the
//                               rank's slice in the multi-GPU fused kernel, csrc/comm.cu)
//                               reference: server.py:751-775 (FedAvg), src/Utils.py:83-98,207-214 (mean/std, LIE)
//   K4     afl_gamma_search     the reference's gamma bisection on Gram scalars (src/Utils.py:115-130)
//          afl_craft_core       mu - gamma*p   /  w + N(0,1)*perturbation (src/Utils.py:52-57)
//   K6     afl_select_core      per-coordinate sort: lower median / trimmed mean (src/Utils.py:267-302,344-357)
//   T8     afl_adam_core / afl_sgdm_core   flat optimisers over [rows, d_pad] (client.py:78,107)
//   K5a/b  afl_hyper_*          fused hyper-network head: emit, delta / W^T delta, rank-1 Adam (server.py:637-678)
#pragma once
#include "common.h"

AFL_HD float4 afl_f4(float v) { float4 r; r.x = v; r.y = v; r.z = v; r.w = v; return r; }
AFL_HD float4 afl_ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
AFL_HD void afl_st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
#define AFL_F4_OP(r, expr_x, expr_y, expr_z, expr_w) do { (r).x = (expr_x); (r).y = (expr_y); (r).z = (expr_z); (r).w = (expr_w); } while (0)

// ----------------------------------------------------------------------------------------- T8 optimisers
AFL_DEV void afl_adam_core(float* w, const float* g, float* m, float* v, int64_t lo, int64_t hi,
                           float lr, float b1, float b2, float eps, float bc1, float bc2, float gscale,
                           int tid, int nt) {
  const float step = lr / bc1, isq = 1.0f / sqrtf(bc2);
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 gv = afl_ld4(g + i), mv = afl_ld4(m + i), vv = afl_ld4(v + i), wv = afl_ld4(w + i);
#define AFL_ADAM1(c) { const float gg = gv.c * gscale; mv.c = b1 * mv.c + (1.f - b1) * gg; \
      vv.c = b2 * vv.c + (1.f - b2) * gg * gg; wv.c -= step * (mv.c / (sqrtf(vv.c) * isq + eps)); }
    AFL_ADAM1(x) AFL_ADAM1(y) AFL_ADAM1(z) AFL_ADAM1(w)
#undef AFL_ADAM1
    afl_st4(m + i, mv); afl_st4(v + i, vv); afl_st4(w + i, wv);
  }
}

AFL_DEV void afl_sgdm_core(float* w, const float* g, float* buf, int64_t lo, int64_t hi, float lr,
                           float momentum, int first_step, float gscale, int tid, int nt) {
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 gv = afl_ld4(g + i), bv = afl_ld4(buf + i), wv = afl_ld4(w + i);
#define AFL_SGD1(c) { const float gg = gv.c * gscale; bv.c = first_step ? gg : momentum * bv.c + gg; wv.c -= lr * bv.c; }
    AFL_SGD1(x) AFL_SGD1(y) AFL_SGD1(z) AFL_SGD1(w)
#undef AFL_SGD1
    afl_st4(buf + i, bv); afl_st4(w + i, wv);
  }
}

// ----------------------------------------------------------------------------------------- K1 / K2
#define AFL_MAX_ATT 4      // attackers' moment sets per pass (the wrapper loops passes)

struct AflFedavgArgs {
  const float* rows;       // [C][d_pad]
  const float* weights;    // [C]  n_i / sum n  (or any aggregation weights)
  int C;
  int64_t d_pad;
  float* out;              // [d_pad] weighted sum (nullable)
  float* bcast;            // [Cb][d_pad]: the result fanned out to every row (nullable)
  int Cb;
  // K2: per-attacker moments over subsets of the rows
  int A;                   // 0..AFL_MAX_ATT
  const float* ref;        // [d_pad] shift (cancellation-safe sums), nullable => rows[0]
  const float* member;     // [A][C] 0/1
  const float* kcount;     // [A] subset sizes (float)
  const float* z;          // [A] LIE scaling factors
  float* mu;               // [A][d_pad] (nullable)
  float* sd;               // [A][d_pad] (nullable)
  float* lie;              // [A][d_pad]  mu + z*sd (nullable)
  int strict_k1;           // 1: sd = NaN when k < 2 (reference), 0: sd = 0
};

// wsm: weights in fast memory [C]; msm: member flags [A][C]
template <int A>
AFL_DEV void afl_fedavg_core(const AflFedavgArgs& a, const float* wsm, const float* msm,
                             int64_t lo, int64_t hi, int tid, int nt) {
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 acc = afl_f4(0.f);
    float4 s1[A > 0 ? A : 1], s2[A > 0 ? A : 1];
    float4 rf = afl_f4(0.f);
    if (A > 0) {
      rf = a.ref ? afl_ld4(a.ref + i) : afl_ld4(a.rows + i);
#pragma unroll
      for (int q = 0; q < A; ++q) { s1[q] = afl_f4(0.f); s2[q] = afl_f4(0.f); }
    }
    for (int c = 0; c < a.C; ++c) {
      const float4 x = afl_ld4(a.rows + (int64_t)c * a.d_pad + i);
      const float wc = wsm[c];
      acc.x += wc * x.x; acc.y += wc * x.y; acc.z += wc * x.z; acc.w += wc * x.w;
      if (A > 0) {
        const float dx = x.x - rf.x, dy = x.y - rf.y, dz = x.z - rf.z, dw = x.w - rf.w;
#pragma unroll
        for (int q = 0; q < A; ++q) {
          const float mq = msm[q * a.C + c];
          s1[q].x += mq * dx; s1[q].y += mq * dy; s1[q].z += mq * dz; s1[q].w += mq * dw;
          s2[q].x += mq * dx * dx; s2[q].y += mq * dy * dy; s2[q].z += mq * dz * dz; s2[q].w += mq * dw * dw;
        }
      }
    }
    if (a.out) afl_st4(a.out + i, acc);
    if (a.bcast)
      for (int c = 0; c < a.Cb; ++c) afl_st4(a.bcast + (int64_t)c * a.d_pad + i, acc);
    if (A > 0) {
#pragma unroll
      for (int q = 0; q < A; ++q) {
        const float k = a.kcount[q];
        float4 mu, sd;
#define AFL_MOM1(c) { mu.c = rf.c + s1[q].c / k; \
          if (k < 2.f) sd.c = a.strict_k1 ? NAN : 0.f; \
          else { float var = (s2[q].c - s1[q].c * s1[q].c / k) / (k - 1.f); sd.c = sqrtf(var > 0.f ? var : 0.f); } }
        AFL_MOM1(x) AFL_MOM1(y) AFL_MOM1(z) AFL_MOM1(w)
#undef AFL_MOM1
        if (a.mu) afl_st4(a.mu + (int64_t)q * a.d_pad + i, mu);
        if (a.sd) afl_st4(a.sd + (int64_t)q * a.d_pad + i, sd);
        if (a.lie) {
          const float zz = a.z[q];
          float4 l; AFL_F4_OP(l, mu.x + zz * sd.x, mu.y + zz * sd.y, mu.z + zz * sd.z, mu.w + zz * sd.w);
          afl_st4(a.lie + (int64_t)q * a.d_pad + i, l);
        }
      }
    }
  }
}

// ----------------------------------------------------------------------------------------- K4
// Gram layout (R = k + 2): rows 0..k-1 genuine (centred), row k = mu - c, row k+1 = p (sigma or sign mu).
// rule: 0 = max (Min-Max / Opt-Fang), 1 = sum of squares (Min-Sum).  out[0] = gamma to use,
// out[1] = gamma_succ, out[2] = threshold, out[3] = number of probes.
AFL_HD void afl_gamma_search(const float* G, int k, int rule, int exclude_first, int paper, float* out) {
  const int R = k + 2;
  if (k <= 1) { out[0] = 0.f; out[1] = 0.f; out[2] = 0.f; out[3] = 0.f; return; }
  double thr = 0.0;
  for (int i = 0; i < k; ++i) {
    double rowsum = 0.0;
    for (int j = 0; j < k; ++j) {
      if (i == j) continue;
      double d2 = (double)G[i * R + i] + (double)G[j * R + j] - 2.0 * (double)G[i * R + j];
      d2 = d2 > 0.0 ? d2 : 0.0;
      if (rule == 0) { double d = sqrt(d2); thr = d > thr ? d : thr; }
      else rowsum += d2;
    }
    if (rule == 1) thr = rowsum > thr ? rowsum : thr;
  }
  const double gmm = G[k * R + k], gpp = G[(k + 1) * R + (k + 1)], gmp = G[k * R + (k + 1)];
  double gamma = 50.0, step = 50.0, succ = 0.0, last = 50.0;
  int probes = 0;
  while (fabs(succ - gamma) > 1.0 && probes < 64) {
    last = gamma;
    double worst = 0.0, total = 0.0;
    for (int j = exclude_first ? 1 : 0; j < k; ++j) {
      double d2 = gmm + gamma * gamma * gpp + (double)G[j * R + j] - 2.0 * gamma * gmp
                  - 2.0 * (double)G[k * R + j] + 2.0 * gamma * (double)G[(k + 1) * R + j];
      d2 = d2 > 0.0 ? d2 : 0.0;
      total += d2;
      double d = sqrt(d2); worst = d > worst ? d : worst;
    }
    const bool ok = rule == 0 ? (worst < thr) : (total < thr);
    if (ok) { succ = gamma; gamma += step / 2; } else { gamma -= step / 2; }
    step /= 2; ++probes;
  }
  out[0] = (float)(paper ? succ : last); out[1] = (float)succ; out[2] = (float)thr; out[3] = (float)probes;
}

// out = mu - gamma*p  (gamma read from device memory so no host sync is needed)
AFL_DEV void afl_craft_axpy_core(float* out, const float* mu, const float* p, const float* gamma_ptr, float sign_mode,
                                 int64_t lo, int64_t hi, int tid, int nt) {
  const float gamma = gamma_ptr[0];
  for (int64_t i = lo + 4 * (int64_t)tid; i < hi; i += 4 * (int64_t)nt) {
    float4 m = afl_ld4(mu + i), d;
    if (sign_mode != 0.f) {   // Opt-Fang: direction = sign(mu)
      AFL_F4_OP(d, (float)((m.x > 0.f) - (m.x < 0.f)), (float)((m.y > 0.f) - (m.y < 0.f)),
                (float)((m.z > 0.f) - (m.z < 0.f)), (float)((m.w > 0.f) - (m.w < 0.f)));
    } else {
      d = afl_ld4(p + i);
    }
    float4 r; AFL_F4_OP(r, m.x - gamma * d.x, m.y - gamma * d.y, m.z - gamma * d.z, m.w - gamma * d.w);
    afl_st4(out + i, r);
  }
}

// out = w + N(0,1)*perturbation on valid coordinates (mask: 1 valid / 0 padding), Philox + Box-Muller
AFL_DEV void afl_craft_random_core(float* out, const float* w, const float* valid, float pert, uint32_t seed_lo,
                                   uint32_t seed_hi, uint32_t stream, int64_t lo, int64_t hi, int tid, int nt) {
  for (int64_t i = lo + tid; i < hi; i += nt) {
    const float n = afl_normal(seed_lo, seed_hi, stream, (uint32_t)i);
    out[i] = (w[i] + n * pert) * (valid ? valid[i] : 1.f);
  }
}

// ----------------------------------------------------------------------------------------- per-round sample indices
// A client's round draws `size` of the `n_rows` training records without replacement (random.sample, RpcClient.py:166) and
// walks them in a fresh random order every epoch (DataLoader(shuffle=True), :169).  Instead of sorting random keys, both
// are keyed pseudo-random PERMUTATIONS evaluated per element: an unbalanced Feistel network on ceil(log2 n) bits with
// cycle walking is a bijection of [0, n), so idx[e][i] = P_subset(P_epoch_e(i)) needs no memory, no sort and is the same
// function on the CPU twin and on the GPU.
AFL_HD uint32_t afl_mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

AFL_HD uint32_t afl_prp(uint32_t i, uint32_t n, uint32_t k0, uint32_t k1) {
  if (n <= 1) return 0;
  int bits = 2;
  while (bits < 32 && (1u << bits) < n) ++bits;
  const int lb = bits / 2, rb = bits - lb;
  const uint32_t lm = (1u << lb) - 1u, rm = (1u << rb) - 1u;
  uint32_t x = i;
  do {
    uint32_t L = x >> rb, R = x & rm;
    for (uint32_t r = 0; r < 6; ++r) {
      if (r & 1u) R = (R ^ afl_mix32(L * 0x9E3779B9u + k1 + r * 0x632BE5ABu)) & rm;
      else        L = (L ^ afl_mix32(R * 0x85EBCA6Bu + k0 + r * 0xC2B2AE35u)) & lm;
    }
    x = (L << rb) | R;
  } while (x >= n);
  return x;
}

// record index of position i (< size) of epoch `epoch` of client `client`
AFL_HD int32_t afl_round_index(uint32_t seed_lo, uint32_t seed_hi, uint32_t client, uint32_t epoch, uint32_t i,
                               uint32_t size, uint32_t n_rows) {
  const uint32_t kc = afl_mix32(seed_lo ^ afl_mix32(client + 0x9E3779B9u)), kd = afl_mix32(seed_hi + 0x7F4A7C15u * (client + 1u));
  const uint32_t ke = afl_mix32(kc ^ afl_mix32(epoch + 0x51ED270Bu));
  const uint32_t pos = afl_prp(i, size, ke, kd ^ epoch);               // this epoch's order
  return (int32_t)afl_prp(pos, n_rows, kc, kd);                       // the round's subset (first `size` images of the PRP)
}

AFL_DEV void afl_round_indices_core(int32_t* idx, const int32_t* sizes, const int32_t* client_ids, int C, int E, int mx,
                                    int n_rows, uint32_t seed_lo, uint32_t seed_hi, int64_t tid, int64_t nt) {
  const int64_t total = (int64_t)C * E * mx;
  for (int64_t t = tid; t < total; t += nt) {
    const int i = (int)(t % mx), e = (int)((t / mx) % E), c = (int)(t / ((int64_t)mx * E));
    const int size = sizes[c] < n_rows ? sizes[c] : n_rows;
    idx[t] = i < size ? afl_round_index(seed_lo, seed_hi, (uint32_t)client_ids[c], (uint32_t)e, (uint32_t)i, (uint32_t)size,
                                        (uint32_t)n_rows)
                      : -1;
  }
}

// ----------------------------------------------------------------------------------------- ScionFL (S7)
// 1-bit stochastic quantisation of a row: sigma_j ~ Bernoulli((x_j - min) / (max - min + 1e-6)), one Philox stream
// per (seed, row).  The dequantised rows  lo_i + sigma_ij (hi_i - lo_i)  are never materialised: everything the filter
// needs (bit counts, <deq_i, agg>, |agg|^2) is computed from the packed bits.
AFL_HD bool afl_scion_bit(float x, float mn, float inv_range, uint32_t seed_lo, uint32_t seed_hi, uint32_t row,
                          uint32_t j) {
  AflU4 r = afl_philox(j >> 2, row, 0x5C10u, 0u, seed_lo, seed_hi);
  const uint32_t v = (j & 3u) == 0 ? r.x : (j & 3u) == 1 ? r.y : (j & 3u) == 2 ? r.z : r.w;
  return (float)(v >> 8) * (1.0f / 16777216.0f) < (x - mn) * inv_range;
}

// agg_j = mean_i (lo_i + sigma_ij * (hi_i - lo_i)) for the 32 coordinates of word `w`, lane `b`
AFL_HD float afl_scion_agg(const uint32_t* bits, const float* lohi, int C, int64_t words, int64_t w, int b) {
  float a = 0.f;
  for (int i = 0; i < C; ++i) {
    const float lo = lohi[2 * i], hi = lohi[2 * i + 1];
    a += ((bits[(int64_t)i * words + w] >> b) & 1u) ? hi : lo;
  }
  return a / (float)C;
}

// ----------------------------------------------------------------------------------------- K6
#define AFL_SEL_MAX 256
// mode 0: lower median; mode 1: trimmed mean dropping `trim` values at each end.
AFL_DEV void afl_select_core(const float* rows, int C, int64_t d_pad, float* out, int mode, int trim,
                             int64_t lo, int64_t hi, int tid, int nt) {
  int P = 1;
  while (P < C) P <<= 1;
  for (int64_t i = lo + tid; i < hi; i += nt) {
    float v[AFL_SEL_MAX];
    for (int c = 0; c < P; ++c) v[c] = c < C ? rows[(int64_t)c * d_pad + i] : INFINITY;
    for (int k = 2; k <= P; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1)
        for (int c = 0; c < P; ++c) {
          const int l = c ^ j;
          if (l > c) {
            const bool up = (c & k) == 0;
            const float a = v[c], b = v[l];
            const bool sw = up ? (a > b) : (a < b);     // NaN compares false: stays in place
            if (sw) { v[c] = b; v[l] = a; }
          }
        }
    if (mode == 0) {
      out[i] = v[(C - 1) / 2];
    } else {
      float s = 0.f;
      for (int c = trim; c < C - trim; ++c) s += v[c];
      out[i] = s / (float)(C - 2 * trim);
    }
  }
}

// ----------------------------------------------------------------------------------------- K5a / K5b
// theta[c][d] = W_h[d,:] . f[c,:] + b_h[d]   for d in [lo,hi), c in [0,C);  f: [C][H] in fast memory
AFL_DEV void afl_hyper_emit_core(const float* head_w, const float* head_b, const float* f, int C, int H,
                                 float* theta, int64_t theta_ld, int64_t lo, int64_t hi, int tid, int nt) {
  for (int64_t it = lo * C + tid; it < hi * C; it += nt) {
    const int64_t d = it / C; const int c = (int)(it - d * C);
    const float* wr = head_w + d * H;
    const float* fr = f + c * H;
    float acc = head_b[d];
    for (int h = 0; h < H; ++h) acc += wr[h] * fr[h];
    theta[(int64_t)c * theta_ld + d] = acc;
  }
}

// Pass 1 of a sequential hyper step for one client: delta[d] = W_h[d,:].f + b_h[d] - theta_client[d];
// thread-partials of |delta|^2 -> part[0], of W_h^T delta -> part[1..H].  (caller reduces the partials)
AFL_DEV void afl_hyper_delta_core(const float* head_w, const float* head_b, const float* f, int H,
                                  const float* theta_client, float* delta, int64_t lo, int64_t hi,
                                  float* part /* [1+H] thread-private */, int tid, int nt) {
  for (int64_t d = lo + tid; d < hi; d += nt) {
    const float* wr = head_w + d * H;
    float acc = head_b[d];
    for (int h = 0; h < H; ++h) acc += wr[h] * f[h];
    const float dl = acc - theta_client[d];
    delta[d] = dl;
    part[0] += dl * dl;
    for (int h = 0; h < H; ++h) part[1 + h] += wr[h] * dl;
  }
}

// Pass 2: Adam on W_h (rank-1 gradient coef*delta[d]*f[h]) and b_h (coef*delta[d]) over rows [lo,hi).
AFL_DEV void afl_hyper_adam_core(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb,
                                 const float* delta, const float* f, int H, const float* coef_ptr,
                                 float lr, float b1, float b2, float eps, float bc1, float bc2,
                                 int64_t lo, int64_t hi, int tid, int nt) {
  const float coef = coef_ptr[0];
  const float step = lr / bc1, isq = 1.0f / sqrtf(bc2);
  for (int64_t it = lo * H + tid; it < hi * H; it += nt) {
    const int64_t d = it / H; const int h = (int)(it - d * H);
    const float g = coef * delta[d] * f[h];
    const float mi = b1 * mw[it] + (1.f - b1) * g;
    const float vi = b2 * vw[it] + (1.f - b2) * g * g;
    mw[it] = mi; vw[it] = vi;
    head_w[it] -= step * (mi / (sqrtf(vi) * isq + eps));
  }
  for (int64_t d = lo + tid; d < hi; d += nt) {
    const float g = coef * delta[d];
    const float mi = b1 * mb[d] + (1.f - b1) * g;
    const float vi = b2 * vb[d] + (1.f - b2) * g * g;
    mb[d] = mi; vb[d] = vi;
    head_b[d] -= step * (mi / (sqrtf(vi) * isq + eps));
  }
}
