// Fused compute + collective kernels over NVLink / NVSwitch peer memory (sm_100a).
//
// K1-fused: FedAvg reduce + scale + broadcast + local fan-out in ONE kernel per rank, no NCCL on the data path.
//   phase 0  every rank pre-reduces its own client rows (sample-weighted sum and, in the same pass, the per-attacker
//            shifted moment sums, K2) over all coordinates into its symmetric partial buffer (local HBM only);
//   barrier  flag barrier in a symmetric signal buffer (st.release.sys / ld.acquire.sys);
//   phase 1  rank r owns coordinate slice r: it sums the G partials by loading them where they live (P2P loads over
//            NVLink), finalises mean / std / LIE rows, and stores the result straight into EVERY rank's global-model
//            (and LIE) buffer (P2P stores);
//   barrier  second flag barrier;
//   phase 2  every rank fans the complete new global model out to its own client rows (local HBM writes only).
//   NVLink bytes per rank and direction: (G-1)/G * D * 4 * (1 + 2A) in, (G-1)/G * D * 4 * (1 + A) out — independent
//   of the number of clients per rank (a one-level version that pulled every remote row measured 3x slower than the
//   NCCL path at 8 rows/rank).  The reference moves 2*C*D*4 bytes through a broker per round
//   (/root/reference/server.py:588-614,751-775).
// All blocks are co-resident (cooperative launch); spins are bounded by a clock64 timeout that raises an error flag
// instead of hanging the GPU.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../include/attackfl_b200.h"
#include "flat_core.h"
#include "runtime.h"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxWorld = 16;
constexpr long long kSpinTimeout = 4000000000LL;   // ~2 s at 2 GHz

struct FusedArgs {
  const float* local_rows_in;     // [CL][d_pad] this rank's client rows
  const float* weights;           // [CL] aggregation weights of the local clients (already normalised globally)
  int CL;
  int64_t d_pad, lo, hi;          // owned slice
  float* part[kMaxWorld];         // every rank's partial buffer [1 + 2A][d_pad] (peer-mapped)
  float* glob[kMaxWorld];         // every rank's global-model buffer (peer-mapped)
  float* lie[kMaxWorld];          // every rank's LIE buffer [A][d_pad] (peer-mapped) or null
  uint32_t* sig[kMaxWorld];       // every rank's signal buffer (peer-mapped), [world] words used
  float* part_mc;                 // NVLS multicast address of the partial buffers (null: use P2P loads)
  float* glob_mc;                 // NVLS multicast address of the global buffers (null: use P2P stores)
  float* lie_mc;                  // NVLS multicast address of the LIE buffers
  int world, rank;
  uint32_t epoch;                 // flag value of this call (monotonic)
  int n_chunks;                   // pipeline depth: sub-ranges per owned slice (<= 32, divides the slice into multiples of 4)
  // moments (K2)
  int A;
  const float* ref;               // [d_pad] shift (previous global model, identical on all ranks)
  const float* member;            // [A][CL] membership of the LOCAL clients in each attacker's subset
  const float* kcount;            // [A] global subset sizes
  const float* z;                 // [A]
  int strict_k1;
  // local phase
  float* local_rows_out;          // [CL][d_pad] fan-out target (nullable)
  uint32_t* grid_ctr;             // local: [2][32] blocks-arrived counters per phase and chunk (reset by the last block)
  int* err;                       // local: set to 1 on spin timeout
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

// NVLS (NVSwitch in-network reduction / multicast): one load returns the SUM over all ranks' copies, one store
// lands in all ranks' copies.  Both are weak operations; the flag barriers around phase 1 order them.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---- chunk-pipelined schedule ------------------------------------------------------------------------------------------
// Every rank's slice [lo, hi) is cut into NC chunks; "chunk j" is the union over the G owners of sub-range j of their
// slice, so the three phases of a chunk touch the same coordinates on every rank:
//   P0(j)  local pre-reduce of this rank's rows over chunk j -> own partial buffer          (local HBM)
//   P1(j)  owner work on sub-range j of the OWN slice: sum of the G partials (NVLS multimem.ld_reduce or P2P loads),
//          mean / std / LIE rows, result stored into every rank's global (and LIE) buffer  (NVLink)
//   P2(j)  fan-out of chunk j of the complete global model to the local client rows         (local HBM)
// A block runs P0(t), P1(t-1), P2(t-2) in iteration t, so the NVLink phase of one chunk overlaps the HBM phases of its
// neighbours.  There is no grid-wide or cross-GPU barrier: the last block of a rank to finish P0(j) / P1(j) publishes one
// flag per peer (st.release.sys, value = epoch) and consumers wait for exactly the G flags they need (ld.acquire.sys,
// bounded spins).  sig layout per rank: [2][kMaxChunks][kMaxWorld] words; ctrl: [2][kMaxChunks] local counters, then err.
constexpr int kMaxChunks = 32;

__device__ __forceinline__ bool wait_flags(const uint32_t* flags, int world, uint32_t ep, int* err) {
  // one thread per block polls; flags[p] >= ep once rank p has published
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    int ok = 1;
    const long long t0 = clock64();
    for (int p = 0; p < world && ok; ++p) {
      while (ld_acquire_sys(flags + p) < ep) {
        if (clock64() - t0 > kSpinTimeout || *reinterpret_cast<volatile int*>(err) != 0) { *err = 1; ok = 0; break; }
      }
    }
    s_ok = ok;
  }
  __syncthreads();
  const bool ok = s_ok != 0;
  __syncthreads();
  return ok;
}

// A block finished its share of a phase of chunk j: the last one on this rank tells every rank (itself included).
__device__ __forceinline__ void publish(const FusedArgs& a, uint32_t* counter, int which, int j) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (atomicAdd(counter, 1u) + 1u == gridDim.x) {
      *counter = 0u;
      __threadfence_system();
      for (int p = 0; p < a.world; ++p)
        st_release_sys(a.sig[p] + ((which * kMaxChunks + j) * kMaxWorld + a.rank), a.epoch);
    }
  }
}

template <int A>
__global__ void __launch_bounds__(kThreads) fedavg_fused_kernel(FusedArgs a) {
  extern __shared__ float sm[];
  float* wsm = sm;                                   // [CL]
  float* msm = sm + a.CL;                            // [A][CL]
  for (int i = threadIdx.x; i < a.CL; i += kThreads) wsm[i] = a.weights[i];
  for (int i = threadIdx.x; i < A * a.CL; i += kThreads) msm[i] = a.member[i];
  __syncthreads();
  const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x, nt = (int64_t)gridDim.x * kThreads;
  constexpr int U = A == 0 ? 4 : (A == 1 ? 2 : 1);   // independent float4 columns per thread (bytes in flight)
  const int64_t slice = a.hi - a.lo;
  const int NC = a.n_chunks;
  const int64_t csz = slice / NC;                    // floats per sub-range (multiple of 4)
  float* mypart = a.part[a.rank];
  uint32_t* cnt0 = a.grid_ctr;
  uint32_t* cnt1 = a.grid_ctr + kMaxChunks;
  const uint32_t* myflags = a.sig[a.rank];

  for (int t = 0; t < NC + 2; ++t) {
    // ---------------------------------------------------------------- P0(t): local pre-reduce over chunk t
    if (t < NC) {
      // element e of the chunk: owner o = e / csz, offset e % csz inside the sub-range.  One division when the loop starts,
      // then (owner, offset) advance incrementally: a 64-bit division per 16 bytes would make this streaming loop ALU-bound.
      const int64_t n = csz * a.world, step = 4 * nt * U;
      int64_t own[U], off[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { const int64_t e = 4 * tid + (int64_t)u * 4 * nt; own[u] = e / csz; off[u] = e - own[u] * csz; }
      for (int64_t e0 = 4 * tid; e0 < n; e0 += step) {
        float4 acc[U], s1[A > 0 ? A : 1][U], s2[A > 0 ? A : 1][U], rf[U];
        int64_t idx[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          idx[u] = own[u] < a.world ? own[u] * slice + (int64_t)t * csz + off[u] : -1;
          off[u] += step;
          while (off[u] >= csz) { off[u] -= csz; ++own[u]; }
          acc[u] = afl_f4(0.f);
          if (A > 0) {
            rf[u] = idx[u] >= 0 ? afl_ld4(a.ref + idx[u]) : afl_f4(0.f);
#pragma unroll
            for (int q = 0; q < A; ++q) { s1[q][u] = afl_f4(0.f); s2[q][u] = afl_f4(0.f); }
          }
        }
        for (int c = 0; c < a.CL; ++c) {
          const float* row = a.local_rows_in + (int64_t)c * a.d_pad;
          const float wc = wsm[c];
          float4 x[U];
#pragma unroll
          for (int u = 0; u < U; ++u) x[u] = idx[u] >= 0 ? afl_ld4(row + idx[u]) : afl_f4(0.f);
#pragma unroll
          for (int u = 0; u < U; ++u) {
            acc[u].x += wc * x[u].x; acc[u].y += wc * x[u].y; acc[u].z += wc * x[u].z; acc[u].w += wc * x[u].w;
            if (A > 0) {
              const float dx = x[u].x - rf[u].x, dy = x[u].y - rf[u].y, dz = x[u].z - rf[u].z, dw = x[u].w - rf[u].w;
#pragma unroll
              for (int q = 0; q < A; ++q) {
                const float mq = msm[q * a.CL + c];
                s1[q][u].x += mq * dx; s1[q][u].y += mq * dy; s1[q][u].z += mq * dz; s1[q][u].w += mq * dw;
                s2[q][u].x += mq * dx * dx; s2[q][u].y += mq * dy * dy; s2[q][u].z += mq * dz * dz; s2[q][u].w += mq * dw * dw;
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (idx[u] < 0) continue;
          afl_st4(mypart + idx[u], acc[u]);
          if (A > 0) {
#pragma unroll
            for (int q = 0; q < A; ++q) {
              afl_st4(mypart + (int64_t)(1 + 2 * q) * a.d_pad + idx[u], s1[q][u]);
              afl_st4(mypart + (int64_t)(2 + 2 * q) * a.d_pad + idx[u], s2[q][u]);
            }
          }
        }
      }
      publish(a, cnt0 + t, 0, t);
    }
    // ---------------------------------------------------------------- P1(t-1): owner reduce + broadcast of sub-range t-1
    if (t >= 1 && t <= NC) {
      const int j = t - 1;
      if (!wait_flags(myflags + (0 * kMaxChunks + j) * kMaxWorld, a.world, a.epoch, a.err)) return;
      const int64_t b0 = a.lo + (int64_t)j * csz, b1 = b0 + csz;
      for (int64_t i0 = b0 + 4 * tid; i0 < b1; i0 += 4 * nt * U) {
        float4 acc[U], s1[A > 0 ? A : 1][U], s2[A > 0 ? A : 1][U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u] = afl_f4(0.f);
#pragma unroll
          for (int q = 0; q < (A > 0 ? A : 1); ++q) { s1[q][u] = afl_f4(0.f); s2[q][u] = afl_f4(0.f); }
        }
        if (a.part_mc) {
          // in-switch reduction: one multimem.ld_reduce per column returns the sum of the G partials
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + (int64_t)u * 4 * nt;
            if (i < b1) {
              acc[u] = multimem_ld_reduce_add(a.part_mc + i);
#pragma unroll
              for (int q = 0; q < A; ++q) {
                s1[q][u] = multimem_ld_reduce_add(a.part_mc + (int64_t)(1 + 2 * q) * a.d_pad + i);
                s2[q][u] = multimem_ld_reduce_add(a.part_mc + (int64_t)(2 + 2 * q) * a.d_pad + i);
              }
            }
          }
        } else {
          for (int r = 0; r < a.world; ++r) {
            const float* pr = a.part[r];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int64_t i = i0 + (int64_t)u * 4 * nt;
              if (i < b1) {
                const float4 x = afl_ld4(pr + i);
                acc[u].x += x.x; acc[u].y += x.y; acc[u].z += x.z; acc[u].w += x.w;
#pragma unroll
                for (int q = 0; q < A; ++q) {
                  const float4 p1 = afl_ld4(pr + (int64_t)(1 + 2 * q) * a.d_pad + i);
                  const float4 p2 = afl_ld4(pr + (int64_t)(2 + 2 * q) * a.d_pad + i);
                  s1[q][u].x += p1.x; s1[q][u].y += p1.y; s1[q][u].z += p1.z; s1[q][u].w += p1.w;
                  s2[q][u].x += p2.x; s2[q][u].y += p2.y; s2[q][u].z += p2.z; s2[q][u].w += p2.w;
                }
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = i0 + (int64_t)u * 4 * nt;
          if (i >= b1) continue;
          if (a.glob_mc) multimem_st(a.glob_mc + i, acc[u]);
          else
            for (int r = 0; r < a.world; ++r) afl_st4(a.glob[r] + i, acc[u]);
          if (A > 0) {
            const float4 rf = afl_ld4(a.ref + i);
#pragma unroll
            for (int q = 0; q < A; ++q) {
              const float k = a.kcount[q], zz = a.z[q];
              float4 l;
#define AFL_LIE1(c) { const float mu = rf.c + s1[q][u].c / k; float sd; \
          if (k < 2.f) sd = a.strict_k1 ? NAN : 0.f; \
          else { const float var = (s2[q][u].c - s1[q][u].c * s1[q][u].c / k) / (k - 1.f); sd = sqrtf(var > 0.f ? var : 0.f); } \
          l.c = mu + zz * sd; }
              AFL_LIE1(x) AFL_LIE1(y) AFL_LIE1(z) AFL_LIE1(w)
#undef AFL_LIE1
              if (a.lie_mc) multimem_st(a.lie_mc + (int64_t)q * a.d_pad + i, l);
              else
                for (int r = 0; r < a.world; ++r) afl_st4(a.lie[r] + (int64_t)q * a.d_pad + i, l);
            }
          }
        }
      }
      publish(a, cnt1 + j, 1, j);
    }
    // ---------------------------------------------------------------- P2(t-2): fan chunk t-2 out to the local client rows
    if (t >= 2) {
      const int j = t - 2;
      // the flags are awaited even without a fan-out target: the call must not return before every owner has delivered
      if (!wait_flags(myflags + (1 * kMaxChunks + j) * kMaxWorld, a.world, a.epoch, a.err)) return;
      if (a.local_rows_out != nullptr) {
        const float* g = a.glob[a.rank];
        const int64_t n = csz * a.world, step = 4 * nt;
        int64_t own = (4 * tid) / csz, off = 4 * tid - own * csz;
        for (int64_t e = 4 * tid; e < n; e += step) {
          const int64_t i = own * slice + (int64_t)j * csz + off;
          off += step;
          while (off >= csz) { off -= csz; ++own; }
          const float4 v = afl_ld4(g + i);
          for (int c = 0; c < a.CL; ++c) afl_st4(a.local_rows_out + (int64_t)c * a.d_pad + i, v);
        }
      }
    }
  }
}

// K3 multi-GPU, step 1: for the owned K-slice [lo, hi) pull the k genuine rows from wherever they live (peer-mapped
// pointers => P2P loads over NVLink), and in the same pass form mu, sigma (unbiased, shifted sums) and the centred
// operand rows of the Gram GEMM:  out[j] = x_j - c (j < k), out[k] = mu - c, out[k+1] = p (sigma, or sign(mu)).
// out: [k+2][ld] fp32 (ld >= hi-lo, multiple of 4); mu_out / p_out: [ld] (kept for the craft step).
__global__ void __launch_bounds__(kThreads) slice_moments_kernel(const float* const* row_ptrs, int k, const float* center,
                                                                 int64_t lo, int64_t hi, int64_t ld, int sign_mode,
                                                                 float* out, float* mu_out, float* p_out) {
  const int64_t n4 = (hi - lo) / 4;
  for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < n4; q += (int64_t)gridDim.x * kThreads) {
    const int64_t i = lo + 4 * q, o = 4 * q;
    const float4 c = afl_ld4(center + i);
    float4 s1 = afl_f4(0.f), s2 = afl_f4(0.f);
    for (int j = 0; j < k; ++j) {
      const float4 x = afl_ld4(row_ptrs[j] + i);
      float4 d; AFL_F4_OP(d, x.x - c.x, x.y - c.y, x.z - c.z, x.w - c.w);
      afl_st4(out + (int64_t)j * ld + o, d);
      s1.x += d.x; s1.y += d.y; s1.z += d.z; s1.w += d.w;
      s2.x += d.x * d.x; s2.y += d.y * d.y; s2.z += d.z * d.z; s2.w += d.w * d.w;
    }
    const float kf = (float)k;
    float4 mc, sd, mu, p;
#define AFL_M1(e) { mc.e = s1.e / kf; const float var = k > 1 ? (s2.e - s1.e * s1.e / kf) / (kf - 1.f) : 0.f; \
      sd.e = sqrtf(var > 0.f ? var : 0.f); mu.e = c.e + mc.e; \
      p.e = sign_mode ? (float)((mu.e > 0.f) - (mu.e < 0.f)) : sd.e; }
    AFL_M1(x) AFL_M1(y) AFL_M1(z) AFL_M1(w)
#undef AFL_M1
    afl_st4(out + (int64_t)k * ld + o, mc);
    afl_st4(out + (int64_t)(k + 1) * ld + o, p);
    afl_st4(mu_out + o, mu);
    afl_st4(p_out + o, p);
  }
}

// step 3: crafted[lo:hi] = mu - gamma * p, stored into EVERY rank's crafted row (P2P stores); gamma from device memory
struct ScatterArgs { float* dst[kMaxWorld]; int world; };
__global__ void __launch_bounds__(kThreads) craft_scatter_kernel(ScatterArgs a, const float* mu, const float* p,
                                                                 const float* gamma_ptr, int64_t lo, int64_t hi) {
  const float gamma = gamma_ptr[0];
  const int64_t n4 = (hi - lo) / 4;
  for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < n4; q += (int64_t)gridDim.x * kThreads) {
    const float4 m = afl_ld4(mu + 4 * q), d = afl_ld4(p + 4 * q);
    float4 r; AFL_F4_OP(r, m.x - gamma * d.x, m.y - gamma * d.y, m.z - gamma * d.z, m.w - gamma * d.w);
    for (int w = 0; w < a.world; ++w) afl_st4(a.dst[w] + lo + 4 * q, r);
  }
}

}  // namespace

extern "C" {

int afl_slice_moments(const float* const* row_ptrs_dev, int k, const float* center, int64_t lo, int64_t hi, int64_t ld,
                      int sign_mode, float* out, float* mu_out, float* p_out, void* stream) {
  AFL_CHECK_ARG((hi - lo) % 4 == 0 && lo % 4 == 0 && ld % 4 == 0, "afl_slice_moments: slice bounds must be multiples of 4");
  int64_t work = ((hi - lo) / 4 + kThreads - 1) / kThreads;
  int grid = (int)(work < 4 * afl_rt::num_sms() ? work : 4 * afl_rt::num_sms());
  if (grid < 1) grid = 1;
  slice_moments_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(row_ptrs_dev, k, center, lo, hi, ld, sign_mode, out,
                                                                    mu_out, p_out);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_slice_moments");
}

int afl_craft_scatter(float* const* dst, int world, const float* mu, const float* p, const float* gamma_dev, int64_t lo,
                      int64_t hi, void* stream) {
  AFL_CHECK_ARG(world >= 1 && world <= kMaxWorld, "afl_craft_scatter: world size must be 1..16");
  ScatterArgs a;
  for (int r = 0; r < kMaxWorld; ++r) a.dst[r] = r < world ? dst[r] : nullptr;
  a.world = world;
  int64_t work = ((hi - lo) / 4 + kThreads - 1) / kThreads;
  int grid = (int)(work < 2 * afl_rt::num_sms() ? work : 2 * afl_rt::num_sms());
  if (grid < 1) grid = 1;
  craft_scatter_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(a, mu, p, gamma_dev, lo, hi);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_craft_scatter");
}


// Fused multi-GPU FedAvg (+ LIE moments), two levels in one kernel.  local_rows [CL][d_pad] and weights/member cover
// the LOCAL clients; part/glob/lie/sig are host arrays of `world` peer-mapped device pointers (part: [1+2A][d_pad]
// per rank).  ctrl: local device memory, 65 x uint32 {[2][32] phase counters, err}, zero-initialised once; sig: symmetric
// uint32 buffers of at least 2 * 32 * 16 words (zero-initialised once).  epoch must increase by one per call on every rank.
// A <= 4 per call.  part_mc / glob_mc / lie_mc: NVLS multicast addresses of the
// same symmetric buffers (null => P2P loads / stores); with them phase 1 uses multimem.ld_reduce / multimem.st.
int afl_fedavg_fused(const float* local_rows, const float* weights, int CL, int64_t d_pad, int world, int rank,
                     float* const* part, float* const* glob, float* const* lie, uint32_t* const* sig, uint32_t epoch,
                     int A, const float* ref, const float* member, const float* kcount, const float* z, int strict_k1,
                     float* local_rows_out, uint32_t* ctrl, float* part_mc, float* glob_mc, float* lie_mc, void* stream) {
  AFL_CHECK_ARG(world >= 1 && world <= kMaxWorld, "afl_fedavg_fused: world size must be 1..16");
  AFL_CHECK_ARG(d_pad % (4 * (int64_t)world) == 0, "afl_fedavg_fused: d_pad must be divisible by 4*world");
  AFL_CHECK_ARG(A >= 0 && A <= AFL_MAX_ATT, "afl_fedavg_fused: at most 4 moment sets per call");
  FusedArgs a;
  a.local_rows_in = local_rows; a.weights = weights; a.CL = CL; a.d_pad = d_pad;
  const int64_t slice = d_pad / world;
  a.lo = rank * slice; a.hi = a.lo + slice;
  for (int r = 0; r < kMaxWorld; ++r) {
    a.part[r] = r < world ? part[r] : nullptr;
    a.glob[r] = r < world ? glob[r] : nullptr;
    a.lie[r] = (r < world && lie) ? lie[r] : nullptr;
    a.sig[r] = r < world ? sig[r] : nullptr;
  }
  a.part_mc = part_mc; a.glob_mc = glob_mc; a.lie_mc = (lie && A > 0) ? lie_mc : nullptr;
  a.world = world; a.rank = rank; a.epoch = epoch; a.A = A; a.ref = ref; a.member = member; a.kcount = kcount;
  a.z = z; a.strict_k1 = strict_k1; a.local_rows_out = local_rows_out;
  a.grid_ctr = ctrl; a.err = reinterpret_cast<int*>(ctrl + 2 * kMaxChunks);
  // pipeline depth: a chunk costs three flag hand-offs (~10 us), so sub-ranges are kept >= 8 MiB; at most 8 chunks
  // (ATTACKFL_FUSED_CHUNKS overrides, 1..32); sub-ranges stay multiples of 4 floats
  int nc = (int)(slice * 4 / (8 << 20));
  if (nc > 8) nc = 8;
  if (const char* ev = getenv("ATTACKFL_FUSED_CHUNKS")) nc = atoi(ev);
  if (nc < 1) nc = 1;
  if (nc > kMaxChunks) nc = kMaxChunks;
  while (nc > 1 && (slice % (4 * (int64_t)nc)) != 0) --nc;
  a.n_chunks = nc;
  const size_t smem = sizeof(float) * (size_t)(CL + A * CL + 2);
  void* args[] = {&a};
  cudaStream_t st = (cudaStream_t)stream;
  const void* fn = nullptr;
  switch (A) {
    case 0: fn = (const void*)fedavg_fused_kernel<0>; break;
    case 1: fn = (const void*)fedavg_fused_kernel<1>; break;
    case 2: fn = (const void*)fedavg_fused_kernel<2>; break;
    case 3: fn = (const void*)fedavg_fused_kernel<3>; break;
    default: fn = (const void*)fedavg_fused_kernel<4>; break;
  }
  // co-resident grid (the in-kernel barriers need it): up to 4 blocks per SM for bytes in flight
  int per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kThreads, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  const int64_t items = (d_pad / 4 + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)per_sm * afl_rt::num_sms();
  if (const char* ev = getenv("ATTACKFL_FUSED_MAX_BLOCKS")) { const int64_t m = atoi(ev); if (m >= 1 && m < cap) cap = m; }   // tests: several emulated ranks on one GPU
  int grid = (int)(items < cap ? items : cap);
  if (grid < 1) grid = 1;
  AFL_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), args, smem, st));
  afl_rt::count_launch();
  return afl_rt::check_last("afl_fedavg_fused");
}

}  // extern "C"
