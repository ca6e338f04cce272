// HAR TransformerClassifier training kernels around the tcgen05 linears (gemm_tcgen05.cu): Conv1d(1->64,k3)+positional
// encoding, multi-head attention over T = 561 tokens (flash-style: scores never leave the SM), residual + dropout +
// LayerNorm, mean pooling and the classifier head with cross-entropy — forward and backward.
//
// Reference: TransformerClassifier / PositionalEncoding (/root/reference/src/Model.py:420-458) trained by train_HAR
// (/root/reference/client.py:114-131): Conv1d(1,64,3,p=1) -> +pe -> 2 x nn.TransformerEncoderLayer(d=64, heads=4,
// ff=256, dropout 0.1, post-norm, ReLU) -> mean over T -> Linear(64,64)-ReLU-Dropout(0.3)-Linear(64,6) -> CrossEntropy.
// Activations are [B*T, 64] row-major with row = b*T + t.  Every dropout mask is a pure function of
// (seed, client, step, row, site, column) (Philox4x32-10, common.h) and is recomputed in the backward pass.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "common.h"
#include "har_rng.cuh"
#include "runtime.h"

namespace {

constexpr int D = 64;          // d_model
constexpr int NH = 4;          // heads
constexpr int DH = 16;         // head dim
constexpr float kLnEps = 1e-5f;

using namespace afl_har;

// ------------------------------------------------------------------------------------------ conv embedding + PE
__global__ void __launch_bounds__(256) embed_fwd_kernel(const float* __restrict__ x, const float* __restrict__ cw,
                                                        const float* __restrict__ cb, const float* __restrict__ pe,
                                                        int B, int T, float* __restrict__ out) {
  const int64_t n4 = (int64_t)B * T * (D / 4);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const int64_t row = i / (D / 4);
    const int c0 = (int)(i - row * (D / 4)) * 4;
    const int b = (int)(row / T), t = (int)(row - (int64_t)b * T);
    const float* xr = x + (int64_t)b * T;
    const float xm = t > 0 ? xr[t - 1] : 0.f, xc = xr[t], xp = t + 1 < T ? xr[t + 1] : 0.f;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + j;
      o[j] = cw[c * 3] * xm + cw[c * 3 + 1] * xc + cw[c * 3 + 2] * xp + cb[c] + pe[(int64_t)t * D + c];
    }
    *reinterpret_cast<float4*>(out + row * D + c0) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// dconv_w[c][k] += sum_rows dh[row][c] * x[b][t+k-1];  dconv_b[c] += sum_rows dh[row][c]
__global__ void __launch_bounds__(256) embed_bwd_kernel(const float* __restrict__ dh, const float* __restrict__ x, int B, int T,
                                                        float* __restrict__ dcw, float* __restrict__ dcb) {
  const int c = threadIdx.x & 63, sub = threadIdx.x >> 6;
  const int64_t M = (int64_t)B * T;
  const int64_t per = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(M, r0 + per);
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, ab = 0.f;
  for (int64_t row = r0 + sub; row < r1; row += 4) {
    const int b = (int)(row / T), t = (int)(row - (int64_t)b * T);
    const float* xr = x + (int64_t)b * T;
    const float g = dh[row * D + c];
    a0 += g * (t > 0 ? xr[t - 1] : 0.f);
    a1 += g * xr[t];
    a2 += g * (t + 1 < T ? xr[t + 1] : 0.f);
    ab += g;
  }
  __shared__ float red[4][4][64];
  red[0][sub][c] = a0; red[1][sub][c] = a1; red[2][sub][c] = a2; red[3][sub][c] = ab;
  __syncthreads();
  if (sub == 0) {
    float s[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k] = red[k][0][c] + red[k][1][c] + red[k][2][c] + red[k][3][c];
    atomicAdd(dcw + c * 3, s[0]); atomicAdd(dcw + c * 3 + 1, s[1]); atomicAdd(dcw + c * 3 + 2, s[2]);
    atomicAdd(dcb + c, s[3]);
  }
}

// ------------------------------------------------------------------------------------------ residual + dropout + LN
// y = LN(x + dropout(r)) * gamma + beta;  one warp per row, two columns per lane.
__global__ void __launch_bounds__(256) add_ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ r, Rng rng,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         int64_t M, float* __restrict__ y, float* __restrict__ xhat,
                                                         float* __restrict__ rstd) {
  const int lane = threadIdx.x & 31;
  const float2 ga = *reinterpret_cast<const float2*>(gamma + 2 * lane), be = *reinterpret_cast<const float2*>(beta + 2 * lane);
  const float inv = rng.p > 0.f ? 1.0f / (1.0f - rng.p) : 1.f;
  for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < M; row += (int64_t)gridDim.x * 8) {
    float2 v = *reinterpret_cast<const float2*>(x + row * D + 2 * lane);
    if (r) {
      float2 rv = *reinterpret_cast<const float2*>(r + row * D + 2 * lane);
      if (rng.p > 0.f) {
        const AflU4 q = rng4(rng, (uint32_t)row, (uint32_t)(2 * lane));
        const uint32_t w0 = (lane & 1) ? q.z : q.x, w1 = (lane & 1) ? q.w : q.y;
        rv.x = u01(w0) >= rng.p ? rv.x * inv : 0.f;
        rv.y = u01(w1) >= rng.p ? rv.y * inv : 0.f;
      }
      v.x += rv.x; v.y += rv.y;
    }
    float s = v.x + v.y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.0f / D);
    const float dx = v.x - mean, dy = v.y - mean;
    float q2 = dx * dx + dy * dy;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q2 += __shfl_xor_sync(0xffffffffu, q2, o);
    const float rs = rsqrtf(q2 * (1.0f / D) + kLnEps);
    const float2 xh = make_float2(dx * rs, dy * rs);
    *reinterpret_cast<float2*>(xhat + row * D + 2 * lane) = xh;
    *reinterpret_cast<float2*>(y + row * D + 2 * lane) = make_float2(xh.x * ga.x + be.x, xh.y * ga.y + be.y);
    if (lane == 0) rstd[row] = rs;
  }
}

// dres = LN backward of dy (dy read from memory, or = pooled_grad[row / T] * bscale when dy_bcast is given);
// dr = dres * dropout mask (the gradient that flows into the residual branch r); dgamma / dbeta accumulated atomically.
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ dy_bcast, int T,
                                                     float bscale, const float* __restrict__ xhat,
                                                     const float* __restrict__ rstd, const float* __restrict__ gamma, Rng rng,
                                                     int64_t M, float* __restrict__ dres, float* __restrict__ dr,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float2 ga = *reinterpret_cast<const float2*>(gamma + 2 * lane);
  const float inv = rng.p > 0.f ? 1.0f / (1.0f - rng.p) : 1.f;
  float2 ag = make_float2(0.f, 0.f), ab = make_float2(0.f, 0.f);
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < M; row += (int64_t)gridDim.x * 8) {
    float2 g;
    if (dy_bcast) {
      g = *reinterpret_cast<const float2*>(dy_bcast + (row / T) * D + 2 * lane);
      g.x *= bscale; g.y *= bscale;
    } else {
      g = *reinterpret_cast<const float2*>(dy + row * D + 2 * lane);
    }
    const float2 xh = *reinterpret_cast<const float2*>(xhat + row * D + 2 * lane);
    ag.x += g.x * xh.x; ag.y += g.y * xh.y; ab.x += g.x; ab.y += g.y;
    const float2 dxh = make_float2(g.x * ga.x, g.y * ga.y);
    float a = dxh.x + dxh.y, b = dxh.x * xh.x + dxh.y * xh.y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    const float rs = rstd[row], m1 = a * (1.0f / D), m2 = b * (1.0f / D);
    const float2 dx = make_float2(rs * (dxh.x - m1 - xh.x * m2), rs * (dxh.y - m1 - xh.y * m2));
    *reinterpret_cast<float2*>(dres + row * D + 2 * lane) = dx;
    if (dr) {
      float2 d2 = dx;
      if (rng.p > 0.f) {
        const AflU4 q = rng4(rng, (uint32_t)row, (uint32_t)(2 * lane));
        const uint32_t w0 = (lane & 1) ? q.z : q.x, w1 = (lane & 1) ? q.w : q.y;
        d2.x = u01(w0) >= rng.p ? d2.x * inv : 0.f;
        d2.y = u01(w1) >= rng.p ? d2.y * inv : 0.f;
      }
      *reinterpret_cast<float2*>(dr + row * D + 2 * lane) = d2;
    }
  }
  __shared__ float red[2][8][64];
  red[0][warp][2 * lane] = ag.x; red[0][warp][2 * lane + 1] = ag.y;
  red[1][warp][2 * lane] = ab.x; red[1][warp][2 * lane + 1] = ab.y;
  __syncthreads();
  if (threadIdx.x < 128) {
    const int k = threadIdx.x >> 6, c = threadIdx.x & 63;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[k][w][c];
    atomicAdd((k == 0 ? dgamma : dbeta) + c, s);
  }
}

// ------------------------------------------------------------------------------------------ attention
// One block = 128 queries of one (sample, head); keys / values stream through shared memory in chunks of 64.
// Scores are kept in base-2 units: s' = (q . k) * scale * log2(e); L = m + log2(l) is stored for the backward pass.
constexpr int kQB = 128;       // queries per block
constexpr int kKC = 64;        // keys per shared-memory chunk

__device__ __forceinline__ void load_kv_chunk(const float* __restrict__ qkv, int64_t row0, int n_valid, int h,
                                              float (*ks)[DH], float (*vs)[DH]) {
  // 64 keys x (4 float4 of K + 4 float4 of V) = 512 float4, 128 threads
  for (int i = threadIdx.x; i < kKC * 8; i += kQB) {
    const int j = i >> 3, part = i & 7;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < n_valid) {
      const float* src = qkv + (row0 + j) * (3 * D) + (part < 4 ? D : 2 * D) + h * DH + (part & 3) * 4;
      v = *reinterpret_cast<const float4*>(src);
    }
    float* dst = (part < 4 ? &ks[j][0] : &vs[j][0]) + (part & 3) * 4;
    *reinterpret_cast<float4*>(dst) = v;
  }
}

__global__ void __launch_bounds__(kQB) attn_fwd_kernel(const float* __restrict__ qkv, int T, Rng rng, float* __restrict__ out,
                                                       float* __restrict__ lse) {
  __shared__ __align__(16) float ks[kKC][DH];
  __shared__ __align__(16) float vs[kKC][DH];
  const int b = blockIdx.z, h = blockIdx.y, qi = blockIdx.x * kQB + threadIdx.x;
  const bool live = qi < T;
  const int64_t base = (int64_t)b * T;
  const float qs = 0.25f * 1.4426950408889634f;            // 1/sqrt(16) * log2(e)
  float q[DH], o[DH];
#pragma unroll
  for (int d = 0; d < DH; ++d) { q[d] = 0.f; o[d] = 0.f; }
  if (live) {
    const float4* src = reinterpret_cast<const float4*>(qkv + (base + qi) * (3 * D) + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4) {
      const float4 v = src[d4];
      q[4 * d4] = v.x * qs; q[4 * d4 + 1] = v.y * qs; q[4 * d4 + 2] = v.z * qs; q[4 * d4 + 3] = v.w * qs;
    }
  }
  float m = -INFINITY, l = 0.f;
  const float inv_keep = rng.p > 0.f ? 1.0f / (1.0f - rng.p) : 1.f;
  Rng hr = rng; hr.site = rng.site + (uint32_t)h;
  const uint32_t thr = attn_thr(rng.p);
  for (int k0 = 0; k0 < T; k0 += kKC) {
    const int nv = min(kKC, T - k0);
    __syncthreads();
    load_kv_chunk(qkv, base + k0, nv, h, ks, vs);
    __syncthreads();
    if (!live) continue;
    for (int j0 = 0; j0 < nv; j0 += 8) {
      float s[8];
      float mx = m;
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        float acc = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) acc = fmaf(q[d], ks[j0 + jj][d], acc);
        s[jj] = (j0 + jj < nv) ? acc : -INFINITY;
        mx = fmaxf(mx, s[jj]);
      }
      if (mx > m) {
        const float c = exp2f(m - mx);
        l *= c;
#pragma unroll
        for (int d = 0; d < DH; ++d) o[d] *= c;
        m = mx;
      }
      AflU4 rw; rw.x = rw.y = rw.z = rw.w = 0u;
      if (rng.p > 0.f) rw = attn_call(hr, (uint32_t)(base + qi), (uint32_t)(k0 + j0));
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        float p = exp2f(s[jj] - m);
        l += p;
        if (rng.p > 0.f) p = attn_bits(rw.x, rw.y, rw.z, rw.w, (uint32_t)jj) >= thr ? p * inv_keep : 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) o[d] = fmaf(p, vs[j0 + jj][d], o[d]);
      }
    }
  }
  if (live) {
    const float il = 1.0f / l;
    float4* dst = reinterpret_cast<float4*>(out + (base + qi) * D + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4)
      dst[d4] = make_float4(o[4 * d4] * il, o[4 * d4 + 1] * il, o[4 * d4 + 2] * il, o[4 * d4 + 3] * il);
    lse[((int64_t)b * NH + h) * lse_stride(T) + qi] = m + log2f(l);
  }
}

// dQ: thread per query, keys / values stream through shared memory (same structure as the forward pass).
__global__ void __launch_bounds__(kQB) attn_bwd_dq_kernel(const float* __restrict__ qkv, const float* __restrict__ out,
                                                          const float* __restrict__ dout, const float* __restrict__ lse, int T,
                                                          Rng rng, float* __restrict__ dqkv) {
  __shared__ __align__(16) float ks[kKC][DH];
  __shared__ __align__(16) float vs[kKC][DH];
  const int b = blockIdx.z, h = blockIdx.y, qi = blockIdx.x * kQB + threadIdx.x;
  const bool live = qi < T;
  const int64_t base = (int64_t)b * T;
  const float qs = 0.25f * 1.4426950408889634f;
  float q[DH], go[DH], dq[DH];
  float delta = 0.f, L = 0.f;
#pragma unroll
  for (int d = 0; d < DH; ++d) { q[d] = 0.f; go[d] = 0.f; dq[d] = 0.f; }
  if (live) {
    const float4* qsrc = reinterpret_cast<const float4*>(qkv + (base + qi) * (3 * D) + h * DH);
    const float4* osrc = reinterpret_cast<const float4*>(out + (base + qi) * D + h * DH);
    const float4* gsrc = reinterpret_cast<const float4*>(dout + (base + qi) * D + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4) {
      const float4 v = qsrc[d4], ov = osrc[d4], gv = gsrc[d4];
      q[4 * d4] = v.x * qs; q[4 * d4 + 1] = v.y * qs; q[4 * d4 + 2] = v.z * qs; q[4 * d4 + 3] = v.w * qs;
      go[4 * d4] = gv.x; go[4 * d4 + 1] = gv.y; go[4 * d4 + 2] = gv.z; go[4 * d4 + 3] = gv.w;
      delta += ov.x * gv.x + ov.y * gv.y + ov.z * gv.z + ov.w * gv.w;
    }
    L = lse[((int64_t)b * NH + h) * lse_stride(T) + qi];
  }
  const float inv_keep = rng.p > 0.f ? 1.0f / (1.0f - rng.p) : 1.f;
  Rng hr = rng; hr.site = rng.site + (uint32_t)h;
  const uint32_t thr = attn_thr(rng.p);
  for (int k0 = 0; k0 < T; k0 += kKC) {
    const int nv = min(kKC, T - k0);
    __syncthreads();
    load_kv_chunk(qkv, base + k0, nv, h, ks, vs);
    __syncthreads();
    if (!live) continue;
    for (int j0 = 0; j0 < nv; j0 += 8) {
      AflU4 rw; rw.x = rw.y = rw.z = rw.w = 0u;
      if (rng.p > 0.f) rw = attn_call(hr, (uint32_t)(base + qi), (uint32_t)(k0 + j0));
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        if (j0 + jj >= nv) break;
        float s = 0.f, dpd = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) { s = fmaf(q[d], ks[j0 + jj][d], s); dpd = fmaf(go[d], vs[j0 + jj][d], dpd); }
        const float p = exp2f(s - L);
        const float keep = (rng.p > 0.f) ? (attn_bits(rw.x, rw.y, rw.z, rw.w, (uint32_t)jj) >= thr ? inv_keep : 0.f) : 1.f;
        const float ds = p * (keep * dpd - delta);
#pragma unroll
        for (int d = 0; d < DH; ++d) dq[d] = fmaf(ds, ks[j0 + jj][d], dq[d]);
      }
    }
  }
  if (live) {
    float4* dst = reinterpret_cast<float4*>(dqkv + (base + qi) * (3 * D) + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4)
      dst[d4] = make_float4(dq[4 * d4] * 0.25f, dq[4 * d4 + 1] * 0.25f, dq[4 * d4 + 2] * 0.25f, dq[4 * d4 + 3] * 0.25f);
  }
}

// dK, dV: thread per key; queries (q, dO, L, delta) stream through shared memory in chunks of 64.
__global__ void __launch_bounds__(kQB) attn_bwd_dkv_kernel(const float* __restrict__ qkv, const float* __restrict__ out,
                                                           const float* __restrict__ dout, const float* __restrict__ lse, int T,
                                                           Rng rng, float* __restrict__ dqkv) {
  __shared__ __align__(16) float qsm[kKC][DH];
  __shared__ __align__(16) float gsm[kKC][DH];
  __shared__ float lsm[kKC], dsm[kKC];
  const int b = blockIdx.z, h = blockIdx.y, kj = blockIdx.x * kQB + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const bool live = kj < T;
  const int64_t base = (int64_t)b * T;
  const float qs = 0.25f * 1.4426950408889634f;
  float k[DH], v[DH], dk[DH], dv[DH];
#pragma unroll
  for (int d = 0; d < DH; ++d) { k[d] = 0.f; v[d] = 0.f; dk[d] = 0.f; dv[d] = 0.f; }
  if (live) {
    const float4* ksrc = reinterpret_cast<const float4*>(qkv + (base + kj) * (3 * D) + D + h * DH);
    const float4* vsrc = reinterpret_cast<const float4*>(qkv + (base + kj) * (3 * D) + 2 * D + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4) {
      const float4 a = ksrc[d4], c = vsrc[d4];
      k[4 * d4] = a.x * qs; k[4 * d4 + 1] = a.y * qs; k[4 * d4 + 2] = a.z * qs; k[4 * d4 + 3] = a.w * qs;
      v[4 * d4] = c.x; v[4 * d4 + 1] = c.y; v[4 * d4 + 2] = c.z; v[4 * d4 + 3] = c.w;
    }
  }
  const float inv_keep = rng.p > 0.f ? 1.0f / (1.0f - rng.p) : 1.f;
  Rng hr = rng; hr.site = rng.site + (uint32_t)h;
  const uint32_t thr = attn_thr(rng.p);
  for (int q0 = 0; q0 < T; q0 += kKC) {
    const int nv = min(kKC, T - q0);
    __syncthreads();
    for (int i = threadIdx.x; i < kKC * 8; i += kQB) {
      const int j = i >> 3, part = i & 7;
      float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j < nv) {
        const float* src = part < 4 ? qkv + (base + q0 + j) * (3 * D) + h * DH + part * 4
                                    : dout + (base + q0 + j) * D + h * DH + (part - 4) * 4;
        val = *reinterpret_cast<const float4*>(src);
      }
      *reinterpret_cast<float4*>((part < 4 ? &qsm[j][0] : &gsm[j][0]) + (part & 3) * 4) = val;
    }
    if (threadIdx.x < kKC) {
      const int j = threadIdx.x;
      float de = 0.f, Lv = 0.f;
      if (j < nv) {
        const float4* osrc = reinterpret_cast<const float4*>(out + (base + q0 + j) * D + h * DH);
        const float4* gsrc = reinterpret_cast<const float4*>(dout + (base + q0 + j) * D + h * DH);
#pragma unroll
        for (int d4 = 0; d4 < 4; ++d4) {
          const float4 ov = osrc[d4], gv = gsrc[d4];
          de += ov.x * gv.x + ov.y * gv.y + ov.z * gv.z + ov.w * gv.w;
        }
        Lv = lse[((int64_t)b * NH + h) * lse_stride(T) + q0 + j];
      }
      dsm[j] = de; lsm[j] = Lv;
    }
    __syncthreads();
    // The dropout bits of (query i, key j) come from the Philox call (row i, j>>3): the eight lanes that share j>>3 each
    // evaluate the call for one of eight consecutive queries and exchange the words with shuffles.
    for (int i0 = 0; i0 < nv; i0 += 8) {
      AflU4 mine; mine.x = mine.y = mine.z = mine.w = 0xffffffffu;
      if (rng.p > 0.f) mine = attn_call(hr, (uint32_t)(base + q0 + i0 + (lane & 7)), (uint32_t)kj);
#pragma unroll
      for (int ii = 0; ii < 8; ++ii) {
        const int src = (lane & ~7) | ii;
        const uint32_t wx = __shfl_sync(0xffffffffu, mine.x, src), wy = __shfl_sync(0xffffffffu, mine.y, src);
        const uint32_t wz = __shfl_sync(0xffffffffu, mine.z, src), ww = __shfl_sync(0xffffffffu, mine.w, src);
        const int i = i0 + ii;
        if (i >= nv || !live) continue;
        float s = 0.f, dpd = 0.f;
#pragma unroll
        for (int d = 0; d < DH; ++d) { s = fmaf(qsm[i][d], k[d], s); dpd = fmaf(gsm[i][d], v[d], dpd); }
        const float p = exp2f(s - lsm[i]);
        const float keep = (rng.p > 0.f) ? (attn_bits(wx, wy, wz, ww, (uint32_t)kj) >= thr ? inv_keep : 0.f) : 1.f;
        const float pd = p * keep;
        const float ds = p * (keep * dpd - dsm[i]);
#pragma unroll
        for (int d = 0; d < DH; ++d) { dv[d] = fmaf(pd, gsm[i][d], dv[d]); dk[d] = fmaf(ds, qsm[i][d], dk[d]); }
      }
    }
  }
  if (live) {
    float4* kd = reinterpret_cast<float4*>(dqkv + (base + kj) * (3 * D) + D + h * DH);
    float4* vd = reinterpret_cast<float4*>(dqkv + (base + kj) * (3 * D) + 2 * D + h * DH);
#pragma unroll
    for (int d4 = 0; d4 < 4; ++d4) {
      kd[d4] = make_float4(dk[4 * d4] * 0.25f, dk[4 * d4 + 1] * 0.25f, dk[4 * d4 + 2] * 0.25f, dk[4 * d4 + 3] * 0.25f);
      vd[d4] = make_float4(dv[4 * d4], dv[4 * d4 + 1], dv[4 * d4 + 2], dv[4 * d4 + 3]);
    }
  }
}

// ------------------------------------------------------------------------------------------ pooling + head
__global__ void __launch_bounds__(256) pool_fwd_kernel(const float* __restrict__ h, int T, float* __restrict__ pooled) {
  const int b = blockIdx.x, c = threadIdx.x & 63, sub = threadIdx.x >> 6;
  float a = 0.f;
  for (int t = sub; t < T; t += 4) a += h[((int64_t)b * T + t) * D + c];
  __shared__ float red[4][64];
  red[sub][c] = a;
  __syncthreads();
  if (sub == 0) pooled[b * D + c] = (red[0][c] + red[1][c] + red[2][c] + red[3][c]) / (float)T;
}

// One block per sample: c1 = relu(W1 p + b1), dropout, logits = W2 c1 + b2, softmax cross-entropy (mean over the batch),
// and the whole backward of the head: dW2, db2, dW1, db1 (atomics) and dpooled.  out[0] += loss_b, out[1] += NaN flag,
// out[2] += correct prediction.
__global__ void __launch_bounds__(64) head_kernel(const float* __restrict__ pooled, const float* __restrict__ w1,
                                                  const float* __restrict__ b1, const float* __restrict__ w2,
                                                  const float* __restrict__ b2, const int64_t* __restrict__ labels, int B,
                                                  int n_cls, Rng rng, int train, float* __restrict__ logits_out,
                                                  float* __restrict__ out, float* __restrict__ dpooled, float* __restrict__ dw1,
                                                  float* __restrict__ db1, float* __restrict__ dw2, float* __restrict__ db2) {
  __shared__ float p[64], c1[64], dc1[64], lg[16], dl[16];
  const int b = blockIdx.x, j = threadIdx.x;
  p[j] = pooled[b * D + j];
  __syncthreads();
  float a = b1[j];
  for (int i = 0; i < D; ++i) a = fmaf(w1[j * D + i], p[i], a);
  a = fmaxf(a, 0.f);
  float keep = 1.f;
  if (train && rng.p > 0.f) {
    const AflU4 q = rng4(rng, (uint32_t)b, (uint32_t)j);
    const uint32_t w = (j & 3) == 0 ? q.x : (j & 3) == 1 ? q.y : (j & 3) == 2 ? q.z : q.w;
    keep = u01(w) >= rng.p ? 1.0f / (1.0f - rng.p) : 0.f;
  }
  c1[j] = a * keep;
  __syncthreads();
  if (j < n_cls) {
    float z = b2[j];
    for (int i = 0; i < D; ++i) z = fmaf(w2[j * D + i], c1[i], z);
    lg[j] = z;
    if (logits_out) logits_out[b * n_cls + j] = z;
  }
  __syncthreads();
  if (j == 0) {
    float mx = -INFINITY;
    for (int c = 0; c < n_cls; ++c) mx = fmaxf(mx, lg[c]);
    float se = 0.f;
    for (int c = 0; c < n_cls; ++c) se += expf(lg[c] - mx);
    const int y = (int)labels[b];
    const float loss = -(lg[y] - mx - logf(se));
    int arg = 0;
    for (int c = 1; c < n_cls; ++c) if (lg[c] > lg[arg]) arg = c;
    for (int c = 0; c < n_cls; ++c) dl[c] = (expf(lg[c] - mx) / se - (c == y ? 1.f : 0.f)) / (float)B;
    atomicAdd(out, loss);
    if (!(loss == loss) || isinf(loss)) atomicAdd(out + 1, 1.f);
    if (arg == y) atomicAdd(out + 2, 1.f);
  }
  if (!train) return;
  __syncthreads();
  // dW2[c][j] += dl[c] * c1[j]; db2[c] += dl[c]; dc1[j] = sum_c dl[c] W2[c][j], through dropout and ReLU
  float g = 0.f;
  for (int c = 0; c < n_cls; ++c) {
    atomicAdd(dw2 + c * D + j, dl[c] * c1[j]);
    g = fmaf(dl[c], w2[c * D + j], g);
  }
  if (j < n_cls) atomicAdd(db2 + j, dl[j]);
  g = (a > 0.f) ? g * keep : 0.f;
  dc1[j] = g;
  atomicAdd(db1 + j, g);
  __syncthreads();
  // dW1[r][j] += dc1[r] * p[j];  dpooled[j] = sum_r dc1[r] W1[r][j]
  float dp = 0.f;
  for (int r = 0; r < D; ++r) {
    const float gr = dc1[r];
    if (gr != 0.f) atomicAdd(dw1 + r * D + j, gr * p[j]);
    dp = fmaf(gr, w1[r * D + j], dp);
  }
  dpooled[b * D + j] = dp;
}

inline int grid_rows(int64_t rows, int rows_per_block, int per_sm) {
  int64_t g = (rows + rows_per_block - 1) / rows_per_block;
  const int64_t cap = (int64_t)afl_rt::num_sms() * per_sm;
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

inline Rng make_rng(uint64_t seed, uint32_t client, uint32_t step, uint32_t site, float p) {
  Rng r; r.s0 = (uint32_t)seed; r.s1 = (uint32_t)(seed >> 32); r.client = client; r.step = step; r.site = site; r.p = p;
  return r;
}

}  // namespace

extern "C" {

int afl_har_embed_fwd(const float* x, const float* conv_w, const float* conv_b, const float* pe, int B, int T, float* out,
                      void* stream) {
  AFL_CHECK_ARG(B >= 1 && T >= 1 && T <= 600, "afl_har_embed_fwd: B >= 1, 1 <= T <= 600");
  embed_fwd_kernel<<<grid_rows((int64_t)B * T * 16, 256, 8), 256, 0, (cudaStream_t)stream>>>(x, conv_w, conv_b, pe, B, T, out);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_embed_fwd");
}

int afl_har_embed_bwd(const float* dh, const float* x, int B, int T, float* dconv_w, float* dconv_b, void* stream) {
  embed_bwd_kernel<<<grid_rows((int64_t)B * T, 64, 8), 256, 0, (cudaStream_t)stream>>>(dh, x, B, T, dconv_w, dconv_b);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_embed_bwd");
}

int afl_har_add_ln_fwd(const float* x, const float* r, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                       uint32_t site, const float* gamma, const float* beta, int64_t M, float* y, float* xhat, float* rstd,
                       void* stream) {
  add_ln_fwd_kernel<<<grid_rows(M, 8, 8), 256, 0, (cudaStream_t)stream>>>(x, r, make_rng(seed, client, step, site, drop_p),
                                                                         gamma, beta, M, y, xhat, rstd);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_add_ln_fwd");
}

int afl_har_ln_bwd(const float* dy, const float* dy_bcast, int T, float bscale, const float* xhat, const float* rstd,
                   const float* gamma, float drop_p, uint64_t seed, uint32_t client, uint32_t step, uint32_t site, int64_t M,
                   float* dres, float* dr, float* dgamma, float* dbeta, void* stream) {
  AFL_CHECK_ARG((dy != nullptr) != (dy_bcast != nullptr), "afl_har_ln_bwd: exactly one of dy / dy_bcast");
  ln_bwd_kernel<<<grid_rows(M, 64, 4), 256, 0, (cudaStream_t)stream>>>(dy, dy_bcast, T, bscale, xhat, rstd, gamma,
                                                                      make_rng(seed, client, step, site, drop_p), M, dres, dr,
                                                                      dgamma, dbeta);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_ln_bwd");
}

int afl_har_attn_fwd(const float* qkv, int B, int T, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                     uint32_t site, float* out, float* lse, void* stream) {
  AFL_CHECK_ARG(B >= 1 && T >= 1, "afl_har_attn_fwd: B, T >= 1");
  dim3 grid((T + kQB - 1) / kQB, NH, B);
  attn_fwd_kernel<<<grid, kQB, 0, (cudaStream_t)stream>>>(qkv, T, make_rng(seed, client, step, site, drop_p), out, lse);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_attn_fwd");
}

int afl_har_attn_bwd(const float* qkv, const float* out, const float* dout, const float* lse, int B, int T, float drop_p,
                     uint64_t seed, uint32_t client, uint32_t step, uint32_t site, float* dqkv, void* stream) {
  dim3 grid((T + kQB - 1) / kQB, NH, B);
  const Rng r = make_rng(seed, client, step, site, drop_p);
  attn_bwd_dq_kernel<<<grid, kQB, 0, (cudaStream_t)stream>>>(qkv, out, dout, lse, T, r, dqkv);
  attn_bwd_dkv_kernel<<<grid, kQB, 0, (cudaStream_t)stream>>>(qkv, out, dout, lse, T, r, dqkv);
  afl_rt::count_launch(2);
  return afl_rt::check_last("afl_har_attn_bwd");
}

int afl_har_lse_stride(int T) { return lse_stride(T); }

int afl_har_pool_fwd(const float* h, int B, int T, float* pooled, void* stream) {
  pool_fwd_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(h, T, pooled);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_pool_fwd");
}

int afl_har_head(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                 const int64_t* labels, int B, int n_cls, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                 uint32_t site, int train, float* logits, float* out3, float* dpooled, float* dw1, float* db1, float* dw2,
                 float* db2, void* stream) {
  AFL_CHECK_ARG(n_cls >= 2 && n_cls <= 16, "afl_har_head: 2 <= classes <= 16");
  head_kernel<<<B, 64, 0, (cudaStream_t)stream>>>(pooled, w1, b1, w2, b2, labels, B, n_cls,
                                                  make_rng(seed, client, step, site, drop_p), train, logits, out3, dpooled, dw1,
                                                  db1, dw2, db2);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_har_head");
}

}  // extern "C"
