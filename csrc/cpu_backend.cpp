// CPU twin of the native library: the SAME kernel cores (flat_core.h, train_core.h) compiled with g++ and
// driven by plain loops (tid = 0, nthreads = 1; the G blocks of a client and the barrier between the
// gradient and optimiser phases become sequential loops).  Exports the same C ABI as the CUDA build, so
// Python picks the library by tensor device.  Used by `--device cpu` (BASELINE config 1) and by the
// non-GPU test-suite, which thereby exercises the kernel source itself.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../include/attackfl_b200.h"
#include "flat_core.h"
#include "train_core.h"

static thread_local char g_err[512] = "";
static void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
#define CHECK_ARG(cond, msg) do { if (!(cond)) { set_error("%s", msg); return -1; } } while (0)

static float* alloc_pool(size_t floats) {
  void* p = nullptr;
  if (posix_memalign(&p, 64, sizeof(float) * (floats + 16)) != 0) return nullptr;
  memset(p, 0, sizeof(float) * (floats + 16));
  return (float*)p;
}

extern "C" {

const char* afl_last_error(void) { return g_err; }
long long afl_launch_count(void) { return 0; }
void afl_reset_launch_count(void) {}
int afl_is_cuda_build(void) { return 0; }
int afl_num_sms(void) { return 0; }

int afl_adam_flat(float* w, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                  int step, const float* gscale, void*) {
  CHECK_ARG(n % 4 == 0, "afl_adam_flat: n must be a multiple of 4");
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  afl_adam_core(w, g, m, v, 0, n, lr, b1, b2, eps, bc1, bc2, gscale ? gscale[0] : 1.f, 0, 1);
  return 0;
}

int afl_sgdm_flat(float* w, const float* g, float* buf, int64_t n, float lr, float momentum, int first_step,
                  const float* gscale, void*) {
  CHECK_ARG(n % 4 == 0, "afl_sgdm_flat: n must be a multiple of 4");
  afl_sgdm_core(w, g, buf, 0, n, lr, momentum, first_step, gscale ? gscale[0] : 1.f, 0, 1);
  return 0;
}

int afl_fedavg(const float* rows, const float* weights, int C, int64_t d_pad, float* out, float* bcast, int Cb,
               int A, const float* ref, const float* member, const float* kcount, const float* z, float* mu,
               float* sd, float* lie, int strict_k1, void*) {
  CHECK_ARG(d_pad % 4 == 0, "afl_fedavg: d_pad must be a multiple of 4");
  CHECK_ARG(A >= 0 && A <= AFL_MAX_ATT, "afl_fedavg: at most 4 moment sets per call");
  AflFedavgArgs a{rows, weights, C, d_pad, out, bcast, Cb, A, ref, member, kcount, z, mu, sd, lie, strict_k1};
  switch (A) {
    case 0: afl_fedavg_core<0>(a, weights, member, 0, d_pad, 0, 1); break;
    case 1: afl_fedavg_core<1>(a, weights, member, 0, d_pad, 0, 1); break;
    case 2: afl_fedavg_core<2>(a, weights, member, 0, d_pad, 0, 1); break;
    case 3: afl_fedavg_core<3>(a, weights, member, 0, d_pad, 0, 1); break;
    default: afl_fedavg_core<4>(a, weights, member, 0, d_pad, 0, 1); break;
  }
  return 0;
}

int afl_gamma_search_dev(const float* G, int k, int rule, int exclude_first, int paper, float* out4, void*) {
  afl_gamma_search(G, k, rule, exclude_first, paper, out4);
  return 0;
}

int afl_craft_axpy(float* out, const float* mu, const float* p, const float* gamma_dev, int sign_mode, int64_t n, void*) {
  CHECK_ARG(n % 4 == 0, "afl_craft_axpy: n must be a multiple of 4");
  afl_craft_axpy_core(out, mu, p, gamma_dev, (float)sign_mode, 0, n, 0, 1);
  return 0;
}

int afl_craft_random(float* out, const float* w, const float* valid, float perturbation, uint64_t seed,
                     uint32_t stream_id, int64_t n, void*) {
  afl_craft_random_core(out, w, valid, perturbation, (uint32_t)seed, (uint32_t)(seed >> 32), stream_id, 0, n, 0, 1);
  return 0;
}

int afl_coord_select(const float* rows, int C, int64_t d_pad, float* out, int mode, int trim, void*) {
  CHECK_ARG(C >= 1 && C <= AFL_SEL_MAX, "afl_coord_select: 1 <= C <= 256");
  CHECK_ARG(mode == 0 || 2 * trim < C, "afl_coord_select: 2*trim must be < C");
  afl_select_core(rows, C, d_pad, out, mode, trim, 0, d_pad, 0, 1);
  return 0;
}

int afl_roc_auc(const float* scores, const float* labels, int64_t n, float* out4, void*) {
  CHECK_ARG(n >= 1 && n <= 8192, "afl_roc_auc: 1 <= n <= 8192 (larger sets: sort on the caller's side)");
  int n_nan = 0;
  for (int64_t i = 0; i < n; ++i) if (scores[i] != scores[i]) ++n_nan;
  if (n_nan) { out4[0] = NAN; out4[1] = (float)n_nan; out4[2] = 0.f; out4[3] = 0.f; return 0; }
  std::vector<int> order((size_t)n);
  for (int64_t i = 0; i < n; ++i) order[(size_t)i] = (int)i;
  std::sort(order.begin(), order.end(), [&](int a, int b) { return scores[a] > scores[b]; });
  double area = 0.0;
  int64_t tp = 0, fp = 0, tp_prev = 0, fp_prev = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (labels[order[(size_t)i]] > 0.5f) ++tp; else ++fp;
    if (i == n - 1 || scores[order[(size_t)i]] != scores[order[(size_t)(i + 1)]]) {      // end of a tie group
      area += 0.5 * (double)(fp - fp_prev) * (double)(tp + tp_prev);
      tp_prev = tp; fp_prev = fp;
    }
  }
  out4[0] = (tp == 0 || fp == 0) ? NAN : (float)(area / ((double)tp * (double)fp));
  out4[1] = 0.f; out4[2] = (float)tp; out4[3] = (float)fp;
  return 0;
}

int afl_round_indices(int32_t* idx, const int32_t* sizes, const int32_t* client_ids, int C, int E, int mx, int n_rows,
                      uint64_t seed, void*) {
  CHECK_ARG(C >= 1 && E >= 1 && mx >= 1 && n_rows >= 1, "afl_round_indices: C, E, mx, n_rows >= 1");
  afl_round_indices_core(idx, sizes, client_ids, C, E, mx, n_rows, (uint32_t)seed, (uint32_t)(seed >> 32), 0, 1);
  return 0;
}

int afl_scion_minmax(const float* rows, const float* valid, int C, int64_t d_pad, float* minmax, void*) {
  CHECK_ARG(C >= 1 && d_pad % 32 == 0, "afl_scion_minmax: C >= 1, d_pad % 32 == 0");
  for (int c = 0; c < C; ++c) {
    float mn = INFINITY, mx = -INFINITY;
    for (int64_t j = 0; j < d_pad; ++j)
      if (!valid || valid[j] != 0.f) { const float x = rows[(int64_t)c * d_pad + j]; mn = fminf(mn, x); mx = fmaxf(mx, x); }
    minmax[2 * c] = mn; minmax[2 * c + 1] = mx;
  }
  return 0;
}

int afl_scion_quantize_rows(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                            int row0, uint32_t* bits, int* counts, void*) {
  CHECK_ARG(C >= 1 && d_pad % 32 == 0 && row0 >= 0, "afl_scion_quantize: C >= 1, d_pad % 32 == 0, row0 >= 0");
  const int64_t words = d_pad / 32;
  for (int c = 0; c < C; ++c) {
    const float mn = minmax[2 * c], inv = 1.0f / (minmax[2 * c + 1] - mn + 1e-6f);
    int cnt = 0;
    for (int64_t w = 0; w < words; ++w) {
      uint32_t word = 0;
      for (int b = 0; b < 32; ++b) {
        const int64_t j = w * 32 + b;
        if ((!valid || valid[j] != 0.f) &&
            afl_scion_bit(rows[(int64_t)c * d_pad + j], mn, inv, (uint32_t)seed, (uint32_t)(seed >> 32), (uint32_t)(row0 + c), (uint32_t)j)) {
          word |= 1u << b; ++cnt;
        }
      }
      bits[(int64_t)c * words + w] = word;
    }
    counts[c] = cnt;
  }
  return 0;
}

int afl_scion_quantize(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                       uint32_t* bits, int* counts, void* stream) {
  return afl_scion_quantize_rows(rows, valid, minmax, C, d_pad, seed, 0, bits, counts, stream);
}

int afl_scion_scores_ptrs(const uint32_t* const* bits_ptrs, const float* valid, const float* lohi, int C, int64_t d_pad,
                          int64_t w_lo, int64_t w_hi, float* out, void*) {
  CHECK_ARG(C >= 1 && C <= 4096 && d_pad % 32 == 0 && 0 <= w_lo && w_lo <= w_hi && w_hi <= d_pad / 32,
            "afl_scion_scores_ptrs: 1 <= C <= 4096, d_pad % 32 == 0, 0 <= w_lo <= w_hi <= d_pad / 32");
  std::vector<double> acc((size_t)C + 2, 0.0);
  for (int64_t w = w_lo; w < w_hi; ++w)
    for (int b = 0; b < 32; ++b) {
      if (valid && valid[w * 32 + b] == 0.f) continue;
      float a = 0.f;
      for (int i = 0; i < C; ++i) a += ((bits_ptrs[i][w] >> b) & 1u) ? lohi[2 * i + 1] : lohi[2 * i];
      a /= (float)C;
      acc[C] += a; acc[C + 1] += (double)a * a;
      for (int i = 0; i < C; ++i) if ((bits_ptrs[i][w] >> b) & 1u) acc[i] += a;
    }
  for (int i = 0; i < C + 2; ++i) out[i] = (float)acc[i];
  return 0;
}

int afl_scion_scores(const uint32_t* bits, const float* valid, const float* lohi, int C, int64_t d_pad, float* out,
                     void* stream) {
  CHECK_ARG(C >= 1 && C <= 4096 && d_pad % 32 == 0, "afl_scion_scores: 1 <= C <= 4096, d_pad % 32 == 0");
  const int64_t words = d_pad / 32;
  std::vector<const uint32_t*> ptrs((size_t)C);
  for (int i = 0; i < C; ++i) ptrs[i] = bits + (int64_t)i * words;
  return afl_scion_scores_ptrs(ptrs.data(), valid, lohi, C, d_pad, 0, words, out, stream);
}

// ---------------------------------------------------------------------------------------- hyper (host loops)
static size_t small_total(int n_nodes, int E, int H) { return (size_t)n_nodes * E + (size_t)H * E + H + 2 * ((size_t)H * H + H); }
static size_t off_w(int n_nodes, int E, int H, int layer) {
  size_t o = (size_t)n_nodes * E;
  if (layer == 0) return o;
  o += (size_t)H * E + H;
  if (layer == 1) return o;
  return o + (size_t)H * H + H;
}

static void features(const float* params, int n_nodes, int E, int H, int c, float* acts /* E+5H */) {
  for (int i = 0; i < E; ++i) acts[i] = params[(size_t)c * E + i];
  const float* x = acts;
  for (int layer = 0; layer < 3; ++layer) {
    const int in = layer == 0 ? E : H;
    const float* W = params + off_w(n_nodes, E, H, layer);
    const float* B = W + (size_t)H * in;
    float* zbuf = acts + E + (layer < 2 ? 2 * layer * H : 4 * H);
    for (int j = 0; j < H; ++j) {
      float a = B[j];
      for (int i = 0; i < in; ++i) a += W[(size_t)j * in + i] * x[i];
      zbuf[j] = a;
    }
    if (layer < 2) {
      float* abuf = acts + E + (2 * layer + 1) * H;
      for (int j = 0; j < H; ++j) abuf[j] = zbuf[j] > 0.f ? zbuf[j] : 0.f;
      x = abuf;
    }
  }
}

int afl_hyper_emit(const float* head_w, const float* head_b, float* small, int n_nodes, int E, int H,
                   const int* clients, int n, float* f_scratch, float* acts, float* theta, int64_t theta_ld,
                   int64_t d_pad, void*) {
  std::vector<float> tmp((size_t)E + 5 * H);
  for (int q = 0; q < n; ++q) {
    float* a = (n == 1 && acts) ? acts : tmp.data();
    features(small, n_nodes, E, H, clients[q], a);
    memcpy(f_scratch + (size_t)q * H, a + E + 4 * H, sizeof(float) * H);
  }
  afl_hyper_emit_core(head_w, head_b, f_scratch, n, H, theta, theta_ld, 0, d_pad, 0, 1);
  return 0;
}

int afl_hyper_step(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, float* small,
                   int n_nodes, int E, int H, int client, const int*, const float* theta_client, float* delta,
                   float* f_scratch, float* acts, float* red, int64_t d_pad, float lr, float b1, float b2, float eps,
                   int step, float clip, void*) {
  const size_t tot = small_total(n_nodes, E, H);
  float* params = small; float* m = small + tot; float* v = small + 2 * tot;
  features(params, n_nodes, E, H, client, acts);
  const float* f = acts + E + 4 * H;
  memcpy(f_scratch, f, sizeof(float) * H);
  for (int i = 0; i < H + 3; ++i) red[i] = 0.f;
  afl_hyper_delta_core(head_w, head_b, f, H, theta_client, delta, 0, d_pad, red, 0, 1);
  std::vector<float> grads(tot, 0.f), g(red + 1, red + 1 + H), gn(H > E ? H : E);
  for (int layer = 2; layer >= 0; --layer) {
    const int in = layer == 0 ? E : H;
    const float* xin = layer == 0 ? acts : acts + E + (2 * (layer - 1) + 1) * H;
    const size_t ow = off_w(n_nodes, E, H, layer);
    for (int j = 0; j < H; ++j) {
      for (int i = 0; i < in; ++i) grads[ow + (size_t)j * in + i] = g[j] * xin[i];
      grads[ow + (size_t)H * in + j] = g[j];
    }
    const float* W = params + ow;
    for (int i = 0; i < in; ++i) {
      float a = 0.f;
      for (int j = 0; j < H; ++j) a += W[(size_t)j * in + i] * g[j];
      if (layer > 0 && !(acts[E + 2 * (layer - 1) * H + i] > 0.f)) a = 0.f;
      gn[i] = a;
    }
    for (int i = 0; i < in; ++i) g[i] = gn[i];
  }
  for (int i = 0; i < E; ++i) grads[(size_t)client * E + i] = g[i];
  double f2 = 0.0, small2 = 0.0;
  for (int i = 0; i < H; ++i) f2 += (double)f[i] * f[i];
  for (size_t i = 0; i < tot; ++i) small2 += (double)grads[i] * grads[i];
  const float norm = (float)sqrt((double)red[0] * (f2 + 1.0) + small2);
  float coef = 1.f;
  if (clip > 0.f) { coef = clip / (norm + 1e-6f); coef = coef < 1.f ? coef : 1.f; }
  red[1 + H] = coef; red[2 + H] = norm;
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  const float stepsz = lr / bc1, isq = 1.0f / sqrtf(bc2);
  for (size_t i = 0; i < tot; ++i) {
    const float gg = grads[i] * coef;
    m[i] = b1 * m[i] + (1.f - b1) * gg;
    v[i] = b2 * v[i] + (1.f - b2) * gg * gg;
    params[i] -= stepsz * (m[i] / (sqrtf(v[i]) * isq + eps));
  }
  afl_hyper_adam_core(head_w, head_b, mw, vw, mb, vb, delta, f_scratch, H, red + 1 + H, lr, b1, b2, eps, bc1, bc2,
                      0, d_pad, 0, 1);
  return 0;
}

// ---------------------------------------------------------------------------------------- training
int afl_train_capacity(int, int, int) { return 1 << 30; }

int afl_train_round(const int32_t* ops_i, int n_ops, int pool_floats, float* w_all, float* m_all, float* v_all,
                    float* gpart_all, int64_t d_pad, const float* data, int rec, const int32_t* idx,
                    int64_t idx_stride_c, int64_t idx_stride_e, const int32_t* n_data, const int32_t* client_ids,
                    int n_clients, int G, int M, int S, int batch, int epochs, int opt_kind, float lr, float beta1,
                    float beta2, float eps, float momentum, int clip_proper, float clip, uint64_t seed, int dropout,
                    float* loss_out, int32_t* steps_out, uint32_t*, float*, int chain, void*) {
  CHECK_ARG(sizeof(AflOp) == 4 * AFL_OP_WORDS, "AflOp layout mismatch");
  CHECK_ARG(G >= 1 && M >= 1 && S >= 1 && G * M * S >= batch, "afl_train_round: G*M*S must cover the batch");
  CHECK_ARG(d_pad % (4 * (int64_t)G) == 0, "afl_train_round: d_pad must be divisible by 4*G");
  const AflOp* ops = reinterpret_cast<const AflOp*>(ops_i);
  float* pool = alloc_pool((size_t)pool_floats);
  if (!pool) { set_error("out of memory"); return -4; }
  std::vector<AflOp> ops_s;
  if (chain) {
    ops_s.resize((size_t)S * n_ops);
    for (int i = 0; i < S * n_ops; ++i) afl_shift_op(ops[i % n_ops], i / n_ops, ops_s[i]);
  }
  for (int c = 0; c < n_clients; ++c) {
    float* w = w_all + (int64_t)c * d_pad;
    float* m = m_all + (int64_t)c * d_pad;
    float* v = v_all ? v_all + (int64_t)c * d_pad : nullptr;
    float* gp = gpart_all + (int64_t)c * G * d_pad;
    memset(gp, 0, sizeof(float) * (size_t)G * d_pad);
    const int n = n_data[c];
    const int steps_per_epoch = (n + batch - 1) / batch;
    const int64_t slice = d_pad / G;
    float loss_acc[2] = {0.f, 0.f};
    double b1t = 1.0, b2t = 1.0;
    int opt_steps = 0, gstep = 0;
    AflTile t;
    memset(&t, 0, sizeof(t));
    t.pool = pool; t.w = w; t.data = data; t.rec = rec; t.probs = nullptr; t.loss_acc = loss_acc;
    t.S = S; t.train = dropout; t.seed_lo = (uint32_t)seed; t.seed_hi = (uint32_t)(seed >> 32);
    t.client = (uint32_t)client_ids[c];
    for (int e = 0; e < epochs; ++e) {
      const int32_t* idx_e = idx + (int64_t)c * idx_stride_c + (int64_t)e * idx_stride_e;
      for (int s = 0; s < steps_per_epoch; ++s, ++gstep) {
        const int b = AFL_MIN(batch, n - s * batch);
        if (b == 1) continue;
        t.step = (uint32_t)gstep; t.inv_b = 1.0f / (float)b;
        for (int g = 0; g < G; ++g) {
          t.gw = gp + (int64_t)g * d_pad;
          for (int mt = 0; mt < M; ++mt) {
            const int base = (g * M + mt) * S;
            const int nv = AFL_MAX(0, AFL_MIN(S, b - base));
            if (nv == 0 && mt > 0) break;
            t.idx = idx_e + (int64_t)s * batch + base;
            t.n_valid = nv; t.sample_base = base; t.accumulate = mt > 0;
            if (chain) {
              afl_run_tile_chain(ops, ops_s.data(), n_ops, t, 0, 1);
            } else {
              afl_run_forward(ops, n_ops, t, 0, 1);
              afl_run_backward(ops, n_ops, t, 0, 1);
            }
          }
        }
        ++opt_steps;
        b1t *= (double)beta1; b2t *= (double)beta2;
        AflOpt o;
        o.kind = opt_kind; o.lr = lr; o.beta1 = beta1; o.beta2 = beta2; o.eps = eps; o.momentum = momentum;
        o.bc1 = (float)(1.0 - b1t); o.bc2 = (float)(1.0 - b2t); o.first_step = opt_steps == 1; o.clip_coef = 1.f;
        if (clip_proper) {
          float sq = 0.f;
          for (int g = 0; g < G; ++g) sq += afl_reduce_sqnorm(gp, G, d_pad, g * slice, (g + 1) * slice, 0, 1);
          const float coef = clip / (sqrtf(sq) + 1e-6f);
          o.clip_coef = coef < 1.f ? coef : 1.f;
        }
        for (int g = 0; g < G; ++g) afl_reduce_update(w, m, v, gp, G, d_pad, g * slice, (g + 1) * slice, o, 0, 1);
      }
    }
    loss_out[2 * c] = loss_acc[0];
    loss_out[2 * c + 1] = loss_acc[1];
    steps_out[c] = opt_steps;
  }
  free(pool);
  return 0;
}

int afl_infer(const int32_t* ops_i, int n_ops, int pool_floats, const float* w, int64_t d_pad, int n_models,
              const float* data, int rec, int64_t n_rows, float* probs, int S, void*) {
  const AflOp* ops = reinterpret_cast<const AflOp*>(ops_i);
  float* pool = alloc_pool((size_t)pool_floats);
  if (!pool) { set_error("out of memory"); return -4; }
  AflTile t;
  memset(&t, 0, sizeof(t));
  t.pool = pool; t.data = data; t.rec = rec; t.S = S;
  for (int mdl = 0; mdl < n_models; ++mdl) {
    t.w = w + (int64_t)mdl * d_pad;
    t.probs = probs + (int64_t)mdl * n_rows;
    for (int64_t r0 = 0; r0 < n_rows; r0 += S) {
      t.row0 = r0;
      t.n_valid = (int)AFL_MIN((int64_t)S, n_rows - r0);
      afl_run_forward(ops, n_ops, t, 0, 1);
    }
  }
  free(pool);
  return 0;
}

int afl_debug_tile(const int32_t* ops_i, int n_ops, int pool_floats, const float* w, float* gw, const float* data,
                   int rec, const int32_t* idx, int S, int n_valid, float inv_b, uint64_t seed, uint32_t client,
                   uint32_t step, int sample_base, int train, float* pool_out, float* loss_out2) {
  const AflOp* ops = reinterpret_cast<const AflOp*>(ops_i);
  float* pool = alloc_pool((size_t)pool_floats);
  if (!pool) { set_error("out of memory"); return -4; }
  float loss_acc[2] = {0.f, 0.f};
  AflTile t;
  memset(&t, 0, sizeof(t));
  t.pool = pool; t.w = w; t.gw = gw; t.data = data; t.rec = rec; t.idx = idx; t.probs = nullptr;
  t.loss_acc = loss_acc; t.S = S; t.n_valid = n_valid; t.accumulate = 0; t.train = train; t.inv_b = inv_b;
  t.seed_lo = (uint32_t)seed; t.seed_hi = (uint32_t)(seed >> 32); t.client = client; t.step = step;
  t.sample_base = sample_base;
  afl_run_forward(ops, n_ops, t, 0, 1);
  afl_run_backward(ops, n_ops, t, 0, 1);
  if (pool_out) memcpy(pool_out, pool, sizeof(float) * (size_t)pool_floats);
  if (loss_out2) { loss_out2[0] = loss_acc[0]; loss_out2[1] = loss_acc[1]; }
  free(pool);
  return 0;
}

int afl_dropout_keep(uint64_t seed, uint32_t client, uint32_t step, uint32_t sample, uint32_t rng, uint32_t elem,
                     float p) {
  return afl_keep((uint32_t)seed, (uint32_t)(seed >> 32), client, step, sample, rng, elem, p) ? 1 : 0;
}

}  // extern "C"
