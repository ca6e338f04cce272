// One Gram for all attackers (K2 + K3 + K4 batched, sm_100a).
//
// What it replaces: every Min-Max / Min-Sum / Opt-Fang attacker of the reference recomputes mean / std and all pairwise
// distances of its own sample of genuine models (/root/reference/src/Utils.py:83-98,108-113,142-147,176-185; the samples
// are drawn per attacker, server.py:596-600).  All samples come from the same n genuine updates, so ONE operand matrix
//     X = [ g_1 - c, ..., g_n - c,  p_1, ..., p_A ]          (p_a = sigma_a or sign(mu_a), R = n + A <= 256 rows)
// and ONE Gram G = X X^T serve every attacker: mu_a - c is the mean of k rows of X, so its Gram entries are means of
// entries of G (O(k^2) scalars), and the gamma bisection of each attacker runs on its own (k+2) x (k+2) sub-Gram.
//   afl_multi_moments  one pass over the n genuine rows of a coordinate slice (rows may be peer-mapped: read in place over
//                      NVLink): writes the centred rows of X, and per attacker mu, p (shifted sums, unbiased std);
//   (Gram)             afl_gram_tf32 on X (csrc/gram_tcgen05.cu); multi-GPU: the R x R partials are summed with K1-fused;
//   afl_multi_gamma    one block per attacker: sub-Gram assembly + the reference's bisection (afl_gamma_search);
//   afl_multi_craft    mu_a - gamma_a p_a for every attacker in one launch, stored through a pointer table (every rank's
//                      copy of the crafted row on multi-GPU runs).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "flat_core.h"
#include "runtime.h"

namespace {

constexpr int kThreads = 256;
constexpr int kGroup = 4;          // attackers per register group in the moments pass

__global__ void __launch_bounds__(kThreads) multi_moments_kernel(const float* const* row_ptrs, int n, const float* center,
                                                                 int64_t lo, int64_t hi, int64_t ld, const float* member,
                                                                 const float* kcount, const int* sign_mode, int A, float* X,
                                                                 float* mu_out, float* p_out) {
  extern __shared__ float msm[];     // [A][n] membership
  for (int i = threadIdx.x; i < A * n; i += kThreads) msm[i] = member[i];
  __syncthreads();
  const int64_t n4 = (hi - lo) / 4;
  for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < n4; q += (int64_t)gridDim.x * kThreads) {
    const int64_t i = lo + 4 * q, o = 4 * q;
    const float4 c = afl_ld4(center + i);
    for (int a0 = 0; a0 < A; a0 += kGroup) {
      float4 s1[kGroup], s2[kGroup];
#pragma unroll
      for (int g = 0; g < kGroup; ++g) { s1[g] = afl_f4(0.f); s2[g] = afl_f4(0.f); }
      for (int j = 0; j < n; ++j) {
        const float4 x = afl_ld4(row_ptrs[j] + i);          // first group: NVLink / HBM; later groups: L1 / L2 hits
        float4 d; AFL_F4_OP(d, x.x - c.x, x.y - c.y, x.z - c.z, x.w - c.w);
        if (a0 == 0) afl_st4(X + (int64_t)j * ld + o, d);
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
          if (a0 + g < A) {
            const float m = msm[(a0 + g) * n + j];
            s1[g].x += m * d.x; s1[g].y += m * d.y; s1[g].z += m * d.z; s1[g].w += m * d.w;
            s2[g].x += m * d.x * d.x; s2[g].y += m * d.y * d.y; s2[g].z += m * d.z * d.z; s2[g].w += m * d.w * d.w;
          }
        }
      }
#pragma unroll
      for (int g = 0; g < kGroup; ++g) {
        const int a = a0 + g;
        if (a >= A) break;
        const float kf = kcount[a];
        const int sgn = sign_mode[a];
        float4 mu, p;
#define AFL_MM1(e) { const float mc = s1[g].e / kf; const float var = kf > 1.f ? (s2[g].e - s1[g].e * s1[g].e / kf) / (kf - 1.f) : 0.f; \
        const float sd = sqrtf(var > 0.f ? var : 0.f); mu.e = c.e + mc; p.e = sgn ? (float)((mu.e > 0.f) - (mu.e < 0.f)) : sd; }
        AFL_MM1(x) AFL_MM1(y) AFL_MM1(z) AFL_MM1(w)
#undef AFL_MM1
        afl_st4(X + (int64_t)(n + a) * ld + o, p);
        afl_st4(mu_out + (int64_t)a * ld + o, mu);
        afl_st4(p_out + (int64_t)a * ld + o, p);
      }
    }
  }
}

// One block per attacker: sub-Gram [S_a..., mu_a, p_a] from the shared Gram, then the bisection.
__global__ void __launch_bounds__(128) multi_gamma_kernel(const float* G, int R, int n, const int* subsets, const int* kk,
                                                          int kmax, const int* rule, int paper, float* out) {
  extern __shared__ float gs[];      // [(k+2)^2]
  const int a = blockIdx.x, k = kk[a], Rs = k + 2;
  const int* S = subsets + (int64_t)a * kmax;
  const int pa = n + a;
  if (k <= 1) { if (threadIdx.x == 0) { out[4 * a] = out[4 * a + 1] = out[4 * a + 2] = out[4 * a + 3] = 0.f; } return; }
  const double inv = 1.0 / (double)k;
  __shared__ int s_sub[200];                                            // the sample, staged once (kmax + 2 <= 200)
  for (int i = threadIdx.x; i < k; i += blockDim.x) s_sub[i] = S[i];
  __syncthreads();
  for (int e = threadIdx.x; e < k * k; e += blockDim.x) {
    const int i = e / k, j = e - i * k;
    gs[i * Rs + j] = G[(int64_t)s_sub[i] * R + s_sub[j]];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    double sm = 0.0;
    for (int t = 0; t < k; ++t) sm += (double)gs[i * Rs + t];            // same terms, same order as the global gather
    const float gm = (float)(sm * inv);
    gs[i * Rs + k] = gm; gs[k * Rs + i] = gm;                          // <g_i, mu>
    const float gp = G[(int64_t)s_sub[i] * R + pa];
    gs[i * Rs + k + 1] = gp; gs[(k + 1) * Rs + i] = gp;                // <g_i, p>
  }
  __syncthreads();
  __shared__ double s_red[128];
  __shared__ double s_thr, s_gamma, s_step, s_succ, s_last;
  __shared__ int s_probes, s_go;
  if (threadIdx.x == 0) {
    double mm = 0.0, mp = 0.0;
    for (int i = 0; i < k; ++i) { mm += (double)gs[i * Rs + k]; mp += (double)gs[i * Rs + k + 1]; }
    gs[k * Rs + k] = (float)(mm * inv);                                // <mu, mu> = mean_i <g_i, mu>
    gs[k * Rs + k + 1] = gs[(k + 1) * Rs + k] = (float)(mp * inv);     // <mu, p>
    gs[(k + 1) * Rs + k + 1] = G[(int64_t)pa * R + pa];
  }
  __syncthreads();
  // The reference's bisection (afl_gamma_search, src/Utils.py:115-130) with the O(k^2) threshold and the O(k) acceptance
  // test of every probe spread over the block; same arithmetic (double), same order of the per-row terms.
  const int rl = rule[a];
  {
    double best = 0.0;
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
      double rowsum = 0.0, rowmax = 0.0;
      for (int j = 0; j < k; ++j) {
        if (i == j) continue;
        double d2 = (double)gs[i * Rs + i] + (double)gs[j * Rs + j] - 2.0 * (double)gs[i * Rs + j];
        d2 = d2 > 0.0 ? d2 : 0.0;
        if (rl == 0) { const double d = sqrt(d2); rowmax = d > rowmax ? d : rowmax; }
        else rowsum += d2;
      }
      const double v = rl == 0 ? rowmax : rowsum;
      best = v > best ? v : best;
    }
    s_red[threadIdx.x] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
      double thr = 0.0;
      for (int t = 0; t < blockDim.x; ++t) thr = s_red[t] > thr ? s_red[t] : thr;
      s_thr = thr; s_gamma = 50.0; s_step = 50.0; s_succ = 0.0; s_last = 50.0; s_probes = 0;
      s_go = fabs(s_succ - s_gamma) > 1.0;
    }
    __syncthreads();
  }
  const double gmm = gs[k * Rs + k], gpp = gs[(k + 1) * Rs + (k + 1)], gmp = gs[k * Rs + (k + 1)];
  while (s_go) {
    const double gamma = s_gamma;
    double worst = 0.0, total = 0.0;
    for (int j = 1 + threadIdx.x; j < k; j += blockDim.x) {           // j = 0 is excluded (aliasing in the reference)
      double d2 = gmm + gamma * gamma * gpp + (double)gs[j * Rs + j] - 2.0 * gamma * gmp
                  - 2.0 * (double)gs[k * Rs + j] + 2.0 * gamma * (double)gs[(k + 1) * Rs + j];
      d2 = d2 > 0.0 ? d2 : 0.0;
      total += d2;
      const double d = sqrt(d2);
      worst = d > worst ? d : worst;
    }
    __syncthreads();
    s_red[threadIdx.x] = rl == 0 ? worst : total;
    __syncthreads();
    if (threadIdx.x == 0) {
      double acc = 0.0;
      for (int t = 0; t < blockDim.x; ++t) acc = rl == 0 ? (s_red[t] > acc ? s_red[t] : acc) : acc + s_red[t];
      s_last = gamma;
      if (acc < s_thr) { s_succ = gamma; s_gamma = gamma + s_step / 2; } else { s_gamma = gamma - s_step / 2; }
      s_step /= 2; ++s_probes;
      s_go = fabs(s_succ - s_gamma) > 1.0 && s_probes < 64;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out[4 * a] = (float)(paper ? s_succ : s_last); out[4 * a + 1] = (float)s_succ; out[4 * a + 2] = (float)s_thr;
    out[4 * a + 3] = (float)s_probes;
  }
}

__global__ void __launch_bounds__(kThreads) multi_craft_kernel(float* const* dst, int world, const float* mu, const float* p,
                                                               int64_t ld, const float* gammas, int64_t lo, int64_t hi) {
  const int a = blockIdx.y;
  const float gamma = gammas[4 * a];
  const int64_t n4 = (hi - lo) / 4;
  for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < n4; q += (int64_t)gridDim.x * kThreads) {
    const float4 m = afl_ld4(mu + (int64_t)a * ld + 4 * q), d = afl_ld4(p + (int64_t)a * ld + 4 * q);
    float4 r; AFL_F4_OP(r, m.x - gamma * d.x, m.y - gamma * d.y, m.z - gamma * d.z, m.w - gamma * d.w);
    for (int w = 0; w < world; ++w) {
      float* t = dst[a * world + w];
      if (t) afl_st4(t + lo + 4 * q, r);
    }
  }
}

}  // namespace

extern "C" {

int afl_multi_moments(const float* const* row_ptrs_dev, int n, const float* center, int64_t lo, int64_t hi, int64_t ld,
                      const float* member_dev, const float* kcount_dev, const int* sign_mode_dev, int A, float* X,
                      float* mu_out, float* p_out, void* stream) {
  AFL_CHECK_ARG((hi - lo) % 4 == 0 && lo % 4 == 0 && ld % 4 == 0, "afl_multi_moments: slice bounds must be multiples of 4");
  AFL_CHECK_ARG(n >= 1 && A >= 1 && (size_t)A * n * 4 <= 160 * 1024, "afl_multi_moments: membership table too large");
  const size_t smem = sizeof(float) * (size_t)A * n;
  AFL_CUDA(cudaFuncSetAttribute(multi_moments_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t work = ((hi - lo) / 4 + kThreads - 1) / kThreads;
  int grid = (int)(work < 4 * afl_rt::num_sms() ? work : 4 * afl_rt::num_sms());
  if (grid < 1) grid = 1;
  multi_moments_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(row_ptrs_dev, n, center, lo, hi, ld, member_dev,
                                                                       kcount_dev, sign_mode_dev, A, X, mu_out, p_out);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_multi_moments");
}

int afl_multi_gamma(const float* G, int R, int n, const int* subsets_dev, const int* k_dev, int kmax, const int* rule_dev,
                    int A, int paper, float* out, void* stream) {
  AFL_CHECK_ARG(A >= 1 && kmax >= 1 && kmax + 2 <= 200, "afl_multi_gamma: subset size must be <= 198");
  const size_t smem = sizeof(float) * (size_t)(kmax + 2) * (kmax + 2);
  AFL_CUDA(cudaFuncSetAttribute(multi_gamma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  multi_gamma_kernel<<<A, 128, smem, (cudaStream_t)stream>>>(G, R, n, subsets_dev, k_dev, kmax, rule_dev, paper, out);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_multi_gamma");
}

int afl_multi_craft(float* const* dst_dev, int world, const float* mu, const float* p, int64_t ld, const float* gammas,
                    int A, int64_t lo, int64_t hi, void* stream) {
  AFL_CHECK_ARG(world >= 1 && A >= 1, "afl_multi_craft: world, A >= 1");
  int64_t work = ((hi - lo) / 4 + kThreads - 1) / kThreads;
  int gx = (int)(work < 2 * afl_rt::num_sms() ? work : 2 * afl_rt::num_sms());
  if (gx < 1) gx = 1;
  multi_craft_kernel<<<dim3(gx, A), kThreads, 0, (cudaStream_t)stream>>>(dst_dev, world, mu, p, ld, gammas, lo, hi);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_multi_craft");
}

}  // extern "C"
