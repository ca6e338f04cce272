// Fused compute + collective kernels over NVLink / NVSwitch peer memory (sm_100a).
//
// K1-fused: FedAvg reduce + scale + all-gather + local fan-out in ONE kernel per rank, no NCCL on the data path.
//   rank r owns coordinates [lo, hi) = slice r of D_pad.  For every float4 of its slice it loads the value from
//   EVERY client's row where that row lives (its own HBM or a peer's, through peer-mapped pointers: P2P loads over
//   NVLink), forms the sample-weighted sum (and, in the same pass, the per-attacker shifted moments -> LIE rows,
//   K2), and stores the result straight into EVERY rank's global-model buffer (P2P stores).  The ranks then meet at
//   a flag barrier in a symmetric signal buffer (st.release.sys / ld.acquire.sys) and each one fans the complete new
//   global model out to its own client rows (local HBM writes only).
//   NVLink bytes per rank: reads (C - C_local) * D/G * 4, writes (G-1) * D/G * 4 (+ A LIE rows) — 1/G of an
//   all-gather of the updates; the reference moves 2*C*D*4 bytes through a broker per round
//   (/root/reference/server.py:588-614,751-775).
// All blocks are co-resident (cooperative launch); spins are bounded by a clock64 timeout that raises an error flag
// instead of hanging the GPU.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "flat_core.h"
#include "runtime.h"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxWorld = 16;
constexpr long long kSpinTimeout = 4000000000LL;   // ~2 s at 2 GHz

struct FusedArgs {
  const float* const* row_ptrs;   // [C] device array: pointer to every client's row (local or peer-mapped)
  const float* weights;           // [C]
  int C;
  int64_t d_pad, lo, hi;          // owned slice
  float* glob[kMaxWorld];         // every rank's global-model buffer (peer-mapped)
  float* lie[kMaxWorld];          // every rank's LIE buffer [A][d_pad] (peer-mapped) or null
  uint32_t* sig[kMaxWorld];       // every rank's signal buffer (peer-mapped), [world] words used
  int world, rank;
  uint32_t epoch;
  // moments (K2)
  int A;
  const float* ref;               // [d_pad] shift (previous global model, identical on all ranks)
  const float* member;            // [A][C]
  const float* kcount;            // [A]
  const float* z;                 // [A]
  int strict_k1;
  // local phase
  float* local_rows;              // [CL][d_pad] fan-out target (nullable)
  int CL;
  uint32_t* grid_ctr;             // local: blocks arrived (reset by the last block)
  uint32_t* go_flag;              // local: epoch published by the last block once all ranks have signalled
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

template <int A>
__global__ void __launch_bounds__(kThreads) fedavg_fused_kernel(FusedArgs a) {
  extern __shared__ float sm[];
  float* wsm = sm;                                   // [C]
  float* msm = sm + a.C;                             // [A][C]
  const float** rp = reinterpret_cast<const float**>(msm + A * a.C + ((a.C + A * a.C) & 1));   // 8-byte aligned
  for (int i = threadIdx.x; i < a.C; i += kThreads) { wsm[i] = a.weights[i]; rp[i] = a.row_ptrs[i]; }
  for (int i = threadIdx.x; i < A * a.C; i += kThreads) msm[i] = a.member[i];
  __syncthreads();

  // ---- phase 1: owned slice: gather (P2P loads) -> reduce -> scatter to every rank (P2P stores)
  const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x, nt = (int64_t)gridDim.x * kThreads;
  if (A == 0) {
    // bandwidth path: U independent float4 columns per thread so that U remote loads are in flight per row
    // (one 16-byte peer load per thread caps at ~300 GB/s: 148*256*16 B in flight / ~2 us NVLink latency)
    constexpr int U = 4;
    for (int64_t i0 = a.lo + 4 * tid; i0 < a.hi; i0 += 4 * nt * U) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) acc[u] = afl_f4(0.f);
      for (int c = 0; c < a.C; ++c) {
        const float* row = rp[c];
        const float wc = wsm[c];
        float4 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t i = i0 + (int64_t)u * 4 * nt;
          x[u] = i < a.hi ? afl_ld4(row + i) : afl_f4(0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u].x += wc * x[u].x; acc[u].y += wc * x[u].y; acc[u].z += wc * x[u].z; acc[u].w += wc * x[u].w;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t i = i0 + (int64_t)u * 4 * nt;
        if (i < a.hi)
          for (int r = 0; r < a.world; ++r) afl_st4(a.glob[r] + i, acc[u]);
      }
    }
  }
  for (int64_t i = a.lo + 4 * tid; A > 0 && i < a.hi; i += 4 * nt) {
    float4 acc = afl_f4(0.f), rf = afl_f4(0.f);
    float4 s1[A > 0 ? A : 1], s2[A > 0 ? A : 1];
    if (A > 0) {
      rf = afl_ld4(a.ref + i);
#pragma unroll
      for (int q = 0; q < A; ++q) { s1[q] = afl_f4(0.f); s2[q] = afl_f4(0.f); }
    }
    for (int c = 0; c < a.C; ++c) {
      const float4 x = afl_ld4(rp[c] + i);
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
    for (int r = 0; r < a.world; ++r) afl_st4(a.glob[r] + i, acc);
    if (A > 0) {
#pragma unroll
      for (int q = 0; q < A; ++q) {
        const float k = a.kcount[q], zz = a.z[q];
        float4 l;
#define AFL_LIE1(c) { const float mu = rf.c + s1[q].c / k; float sd; \
          if (k < 2.f) sd = a.strict_k1 ? NAN : 0.f; \
          else { const float var = (s2[q].c - s1[q].c * s1[q].c / k) / (k - 1.f); sd = sqrtf(var > 0.f ? var : 0.f); } \
          l.c = mu + zz * sd; }
        AFL_LIE1(x) AFL_LIE1(y) AFL_LIE1(z) AFL_LIE1(w)
#undef AFL_LIE1
        for (int r = 0; r < a.world; ++r) afl_st4(a.lie[r] + (int64_t)q * a.d_pad + i, l);
      }
    }
  }
  __threadfence_system();
  __syncthreads();

  // ---- phase 2: local grid barrier; the last block signals every peer and waits for every peer
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    int ok = 1;
    const uint32_t arrived = atomicAdd(a.grid_ctr, 1u) + 1u;
    if (arrived == gridDim.x) {
      *a.grid_ctr = 0u;                                            // ready for the next launch
      __threadfence_system();
      for (int p = 0; p < a.world; ++p)
        if (p != a.rank) st_release_sys(a.sig[p] + a.rank, a.epoch);
      const long long t0 = clock64();
      for (int p = 0; p < a.world; ++p) {
        if (p == a.rank) continue;
        while (ld_acquire_sys(a.sig[a.rank] + p) < a.epoch) {
          if (clock64() - t0 > kSpinTimeout) { *a.err = 1; ok = 0; break; }
        }
      }
      __threadfence_system();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a.go_flag), "r"(ok ? a.epoch : 0xFFFFFFFFu) : "memory");
    } else {
      const long long t0 = clock64();
      uint32_t g;
      while ((g = ld_acquire_gpu(a.go_flag)) < a.epoch) {
        if (clock64() - t0 > 2 * kSpinTimeout) { *a.err = 1; ok = 0; break; }
      }
      if (g == 0xFFFFFFFFu) ok = 0;
    }
    s_ok = ok;
  }
  __syncthreads();
  if (!s_ok || a.local_rows == nullptr) return;

  // ---- phase 3: fan the complete global model out to this rank's client rows (local HBM only)
  const float* g = a.glob[a.rank];
  for (int64_t i = 4 * tid; i < a.d_pad; i += 4 * nt) {
    const float4 v = afl_ld4(g + i);
    for (int c = 0; c < a.CL; ++c) afl_st4(a.local_rows + (int64_t)c * a.d_pad + i, v);
  }
}

}  // namespace

extern "C" {

// Fused multi-GPU FedAvg (+ LIE moments).  Pointer tables: row_ptrs (device) [C]; glob/lie/sig (host arrays of
// `world` peer-mapped device pointers).  ctrl: local device memory, 4 x uint32 {grid_ctr, go_flag, err, pad},
// zero-initialised once.  epoch must increase by one per call on every rank.  A <= 4 per call.
int afl_fedavg_fused(const float* const* row_ptrs, const float* weights, int C, int64_t d_pad, int world, int rank,
                     float* const* glob, float* const* lie, uint32_t* const* sig, uint32_t epoch, int A,
                     const float* ref, const float* member, const float* kcount, const float* z, int strict_k1,
                     float* local_rows, int CL, uint32_t* ctrl, void* stream) {
  AFL_CHECK_ARG(world >= 1 && world <= kMaxWorld, "afl_fedavg_fused: world size must be 1..16");
  AFL_CHECK_ARG(d_pad % (4 * (int64_t)world) == 0, "afl_fedavg_fused: d_pad must be divisible by 4*world");
  AFL_CHECK_ARG(A >= 0 && A <= AFL_MAX_ATT, "afl_fedavg_fused: at most 4 moment sets per call");
  FusedArgs a;
  a.row_ptrs = row_ptrs; a.weights = weights; a.C = C; a.d_pad = d_pad;
  const int64_t slice = d_pad / world;
  a.lo = rank * slice; a.hi = a.lo + slice;
  for (int r = 0; r < kMaxWorld; ++r) {
    a.glob[r] = r < world ? glob[r] : nullptr;
    a.lie[r] = (r < world && lie) ? lie[r] : nullptr;
    a.sig[r] = r < world ? sig[r] : nullptr;
  }
  a.world = world; a.rank = rank; a.epoch = epoch; a.A = A; a.ref = ref; a.member = member; a.kcount = kcount;
  a.z = z; a.strict_k1 = strict_k1; a.local_rows = local_rows; a.CL = CL;
  a.grid_ctr = ctrl; a.go_flag = ctrl + 1; a.err = reinterpret_cast<int*>(ctrl + 2);
  const size_t smem = sizeof(float) * (size_t)(C + A * C + 2) + sizeof(void*) * (size_t)C;
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
  const int64_t slice_items = (slice / 4 + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)per_sm * afl_rt::num_sms();
  int grid = (int)(slice_items < cap ? slice_items : cap);
  if (grid < 1) grid = 1;
  AFL_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kThreads), args, smem, st));
  afl_rt::count_launch();
  return afl_rt::check_last("afl_fedavg_fused");
}

}  // extern "C"
