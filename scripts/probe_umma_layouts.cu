// Probe: which shared-memory operand layouts does tcgen05.mma kind::tf32 accept when the operand tiles are written by
// threads (not TMA)?  One CTA, D[128 x 64] = A[128 x 32] . B[64 x 32]^T with small integers (exact in tf32).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -o build/probe_umma scripts/probe_umma_layouts.cu
// Variants (A placement / B placement):
//   0  A K-major SW128          B K-major SW128                       (control)
//   1  A K-major SW128          B K-major, layout type 1 (SW128_32B)  (is the 32B-atom swizzle usable K-major?)
//   2  A K-major type 1         B K-major type 1
//   3  A K-major SW128          B MN-major SW128_32B                  (control, as the HAR kernels use it)
//   4  A K-major no-swizzle     B K-major no-swizzle
//   5  A K-major no-swizzle     B MN-major no-swizzle, SBO = mn stride, LBO = k stride
//   6  A K-major no-swizzle     B MN-major no-swizzle, fields swapped
//   7  A MN-major no-swizzle (SBO = mn, LBO = k)   B K-major no-swizzle
//   8  A MN-major SW128_32B     B K-major SW128                       (control)
//   9  M = 64 (A K-major SW128, B K-major SW128): reports which TMEM lanes hold the 64 rows
// Also times a dependent chain MMA(K=64) -> commit -> wait -> tcgen05.ld -> st.shared -> fence -> next MMA.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../csrc/tc_common.cuh"

namespace afl_rt {
void set_error(const char*, ...) {}
}

using namespace afl_tc;

constexpr int M = 128, N = 64, K = 32;

__device__ __forceinline__ uint64_t desc_generic(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// byte offset of logical element (r = M/N index, k) for each placement kind
__device__ __forceinline__ uint32_t place(int kind, int r, int k, int rows) {
  switch (kind) {
    case 0: return (r / 8) * 1024 + (r % 8) * 128 + (((k / 4) ^ (r % 8)) * 16) + (k % 4) * 4;         // K-major SW128
    case 1: return r * 128 + (((k / 8) ^ (r & 3)) * 32) + (k % 8) * 4;                                  // K-major, 32B-atom swizzle
    case 2: return (r / 8) * 1024 + (k / 4) * 128 + (r % 8) * 16 + (k % 4) * 4;                         // K-major no swizzle (K = 32)
    case 3: return (r / 32) * (K * 128) + k * 128 + ((((r % 32) / 8) ^ (k & 3)) * 32) + (r % 8) * 4;    // MN-major SW128_32B
    case 4: return (k / 8) * ((rows / 4) * 128) + (r / 4) * 128 + (k % 8) * 16 + (r % 4) * 4;           // MN-major no swizzle
  }
  return 0;
}

struct Variant { int a_kind, b_kind, mm; };

__global__ void probe_kernel(float* out, long long* cycles, int variant, int a_kind, int b_kind, int mm, int swap_fields) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;                 // 16 KB
  uint8_t* sb = smem + 16384;         // 8 KB
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int tid = threadIdx.x;
  for (int i = tid; i < (16384 + 8192) / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < M * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    float v = (float)(((r * 31 + k * 7) % 61) - 30);
    if (r < mm) *reinterpret_cast<float*>(sa + place(a_kind, r, k, M)) = v;
  }
  for (int i = tid; i < N * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    float v = (float)(((r * 5 + k * 2) % 13) - 6);
    *reinterpret_cast<float*>(sb + place(b_kind, r, k, N)) = v;
  }
  if (tid == 0) mbar_init(&bar, 1);
  if (tid < 32) tmem_alloc(&tslot, 128);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  if (a_kind == 5) {   // A operand in TMEM columns [64, 96): lane = row, column = k
    uint32_t r[32];
    for (int k = 0; k < 32; ++k) r[k] = __float_as_uint((float)(((tid * 31 + k * 7) % 61) - 30));
    const uint32_t taddr = tmem + 64 + ((uint32_t)(tid & ~31) << 16);
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
           "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
           "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
           "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (tid == 0) {
    const uint32_t a_addr = smem_u32(sa), b_addr = smem_u32(sb);
    const int a_mn = (a_kind == 3 || a_kind == 4), b_mn = (b_kind >= 3);
    const uint32_t idesc = idesc_tf32(mm, N, a_mn, b_mn);
    for (int ks = 0; ks < K / 8; ++ks) {
      uint64_t ad, bd;
      auto mk = [&](int kind, uint32_t addr, int rows) -> uint64_t {
        switch (kind) {
          case 0: return desc_add(desc_k_sw128(addr), ks * 32);
          case 1: return desc_add(desc_generic(addr, 16, 1024, 1), ks * 32);
          case 2: return desc_add(desc_generic(addr, 128, 1024, 0), ks * 256);
          case 3: return desc_add(desc_mn_sw128(addr, K * 128), ks * 1024);
          case 4: {
            uint32_t kstride = (rows / 4) * 128, mnstride = 128;
            return swap_fields ? desc_add(desc_generic(addr, mnstride, kstride, 0), ks * kstride)
                               : desc_add(desc_generic(addr, kstride, mnstride, 0), ks * kstride);
          }
        }
        return 0;
      };
      bd = mk(b_kind, b_addr, N);
      if (a_kind == 5) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
            ::"r"(tmem), "r"(tmem + 64 + ks * 8), "l"(bd), "r"(idesc), "r"((uint32_t)(ks > 0)) : "memory");
      } else {
        ad = mk(a_kind, a_addr, M);
        umma_tf32(tmem, ad, bd, idesc, ks > 0);
      }
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  if (tid < 128) {
    float v[64];
    const uint32_t taddr = tmem + ((uint32_t)(tid & ~31) << 16);
    tmem_ld32_async(taddr, v);
    tmem_ld32_async(taddr + 32, v + 32);
    tmem_ld_wait();
    for (int j = 0; j < 64; ++j) out[tid * 64 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();

  if (variant == 10 || variant == 0) {   // issue time of 8 MMAs (K = 64) + commit by one thread, nothing else running
    tc_fence_after();
    long long t0 = 0, t1 = 0, t2 = 0;
    if (tid == 0) {
      const uint32_t idesc = idesc_tf32(128, 64, 0, 0);
      const uint64_t ad = desc_k_sw128(smem_u32(sa)), bd = desc_k_sw128(smem_u32(sb));
      t0 = clock64();
      for (int ks = 0; ks < 8; ++ks) {
        if (variant == 10)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
                       ::"r"(tmem), "r"(tmem + 64 + (ks & 3) * 8), "l"(desc_add(bd, (ks & 3) * 32)), "r"(idesc), "r"((uint32_t)(ks > 0)) : "memory");
        else umma_tf32(tmem, desc_add(ad, (ks & 3) * 32), desc_add(bd, (ks & 3) * 32), idesc, ks > 0);
      }
      umma_commit(&bar);
      t1 = clock64();
    }
    mbar_wait(&bar, 1);
    if (tid == 0) { t2 = clock64(); cycles[1] = t1 - t0; cycles[2] = t2 - t0; }
    tc_fence_after();
    __syncthreads();
  }
  // ---- latency chain (variant 0 only): 16 dependent phases of [MMA K=64 (8 instr) -> commit -> wait -> ld 64 cols -> st.shared -> fence]
  if (variant == 0) {
    tc_fence_after();
    uint32_t parity = 0;
    long long t0 = clock64();
    for (int it = 0; it < 16; ++it) {
      if (tid == 0) {
        const uint32_t idesc = idesc_tf32(128, 64, 0, 0);
        const uint64_t ad = desc_k_sw128(smem_u32(sa)), bd = desc_k_sw128(smem_u32(sb));
        for (int ks = 0; ks < 8; ++ks) umma_tf32(tmem, desc_add(ad, (ks & 3) * 32), desc_add(bd, (ks & 3) * 32), idesc, ks > 0);
        umma_commit(&bar);
      }
      mbar_wait(&bar, parity);
      parity ^= 1;
      tc_fence_after();
      if (tid < 128) {
        float v[64];
        const uint32_t taddr = tmem + ((uint32_t)(tid & ~31) << 16);
        tmem_ld32_async(taddr, v);
        tmem_ld32_async(taddr + 32, v + 32);
        tmem_ld_wait();
        float4* dst = reinterpret_cast<float4*>(sa + (tid / 8) * 1024 + (tid % 8) * 128);
        for (int j = 0; j < 8; ++j) dst[j ^ (tid % 8)] = make_float4(v[4 * j] * 1e-3f, v[4 * j + 1] * 1e-3f, v[4 * j + 2] * 1e-3f, v[4 * j + 3] * 1e-3f);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
    }
    long long t1 = clock64();
    if (tid == 0) cycles[0] = (t1 - t0) / 16;
  }
  __syncthreads();
  if (tid < 32) tmem_dealloc(tmem, 128);
}

int main(int argc, char** argv) {
  const int only = argc > 1 ? atoi(argv[1]) : -1;
  float* d_out; long long* d_cyc;
  cudaMalloc(&d_out, sizeof(float) * 128 * 64);
  cudaMalloc(&d_cyc, 4 * sizeof(long long));
  std::vector<float> ref(M * N), got(128 * 64);
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      float s = 0;
      for (int k = 0; k < K; ++k) s += (float)(((m * 31 + k * 7) % 61) - 30) * (float)(((n * 5 + k * 2) % 13) - 6);
      ref[m * N + n] = s;
    }
  struct V { int a, b, mm, swap; const char* name; } vs[] = {
      {0, 0, 128, 0, "A K-sw128 / B K-sw128 (control)"},
      {0, 1, 128, 0, "A K-sw128 / B K-major layout type 1 (32B-atom swizzle)"},
      {1, 1, 128, 0, "A K-major type 1 / B K-major type 1"},
      {0, 3, 128, 0, "A K-sw128 / B MN-sw128_32B (control)"},
      {2, 2, 128, 0, "A K-noswz / B K-noswz"},
      {2, 4, 128, 0, "A K-noswz / B MN-noswz (LBO = k stride, SBO = mn stride)"},
      {2, 4, 128, 1, "A K-noswz / B MN-noswz (fields swapped)"},
      {4, 2, 128, 0, "A MN-noswz / B K-noswz"},
      {3, 0, 128, 0, "A MN-sw128_32B / B K-sw128 (control)"},
      {0, 0, 64, 0, "M = 64: lane map"},
      {5, 0, 128, 0, "A in TMEM (tcgen05.st) / B K-sw128"},
      {5, 3, 128, 0, "A in TMEM (tcgen05.st) / B MN-sw128_32B"},
  };
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  int vi = 0;
  for (auto& v : vs) {
    if (only >= 0 && vi != only) { ++vi; continue; }
    cudaMemset(d_out, 0xff, sizeof(float) * 128 * 64);
    probe_kernel<<<1, 128, 16384 + 8192 + 1024>>>(d_out, d_cyc, vi, v.a, v.b, v.mm, v.swap);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("variant %d (%s): CUDA error %s\n", vi, v.name, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(got.data(), d_out, sizeof(float) * 128 * 64, cudaMemcpyDeviceToHost);
    if (v.mm == 128) {
      double maxerr = 0; int bad = 0;
      for (int i = 0; i < M * N; ++i) { double d = fabs((double)got[i] - ref[i]); if (!(d <= 1e-3)) ++bad; if (d > maxerr) maxerr = d; }
      printf("variant %d: %-62s max_err %.3g bad %d/%d  %s\n", vi, v.name, maxerr, bad, M * N, bad == 0 ? "OK" : "MISMATCH");
    } else {
      printf("variant %d: %s\n  lanes holding logical row r (first col match): ", vi, v.name);
      for (int r = 0; r < 64; ++r) {
        int found = -1;
        for (int lane = 0; lane < 128; ++lane) {
          bool ok = true;
          for (int n = 0; n < N; ++n) if (fabs(got[lane * 64 + n] - ref[r * N + n]) > 1e-3) { ok = false; break; }
          if (ok) { found = lane; break; }
        }
        printf("%d ", found);
      }
      printf("\n");
    }
    if (vi == 0 || vi == 10) { long long cc[3]; cudaMemcpy(cc, d_cyc, sizeof(cc), cudaMemcpyDeviceToHost); printf("  8 MMAs (128x64x8 tf32, A %s) + commit: issue %lld cycles, issue->complete %lld cycles\n", vi == 10 ? "in TMEM" : "in smem", cc[1], cc[2]); }
    if (vi == 0) { long long c; cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost); printf("  dependent phase latency: %lld cycles (MMA 128x64x64 tf32 + commit + wait + ld 64 cols + st.shared + fences + bar)\n", c); }
    ++vi;
  }
  return 0;
}
