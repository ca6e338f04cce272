// K3 — split-K Gram matrix G = A A^T on 5th-gen tensor cores (sm_100a): TMA -> smem (128B swizzle) ->
// tcgen05.mma (kind::f16, bf16 operands, fp32 accumulate in TMEM) -> tcgen05.ld epilogue.
//
// What it replaces: the reference's O(k^2) Python loops of per-tensor torch.linalg.norm on the CPU
// (/root/reference/src/Utils.py:30-49 compute_distance; :108-113,142-147,176-185 pairwise loops; :326-342 Krum;
// server.py:310-350 ShieldFL cosines; :703-743 FLTrust) — 9-18 s at k=48.  One Gram of the centred
// rows [g_1-c .. g_k-c, mu-c, p] gives every distance the gamma search needs.
//
// Shape of the problem: R <= 256 rows, D = model size (1e5 .. 1e8) => the output is a single 128/256-row tile
// and ALL parallelism is split-K: CTA b owns a contiguous range of 64-wide K-blocks, accumulates its partial
// Gram in TMEM and writes it to `partials[b]` (scratch); a second kernel reduces the partials in fp32.
// A and B operands are the SAME smem tile (rows x 64 bf16, K-major, SWIZZLE_128B): one TMA load feeds both.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA issuer, warps 2..5 = epilogue
// (each reads its 32-lane quarter of TMEM with tcgen05.ld.32x32b).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/attackfl_b200.h"
#include "runtime.h"

namespace {

constexpr int kBlockK = 64;            // bf16 elements per K-block = 128 bytes = one swizzle row
constexpr int kUmmaK = 16;
constexpr int kThreads = 192;
constexpr int kMaxRows = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, sm100 version bit set):
//   start_address [0,14) = addr >> 4;  LBO [16,30) unused for swizzled K-major (1);  SBO [32,46) = 1024 B >> 4
//   (8 rows x 128 B per swizzle atom);  version [46,48) = 1;  layout_type [61,64) = 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format[4,6)=1 (F32), a_format[7,10)=1 (BF16),
// b_format[10,13)=1 (BF16), a_major[15]=0, b_major[16]=0 (K-major), n_dim[17,23)=N>>3, m_dim[24,29)=M>>4.
__device__ __forceinline__ uint32_t make_idesc(int M, int N, uint32_t fmt /* 1 = BF16, 2 = TF32 */) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

struct GramParams {
  int rows_box;          // 128 or 256: rows of the TMA box / smem tile
  int n_cols;            // UMMA N (multiple of 16, >= 16): padded number of Gram rows
  int m_tiles;           // 1 or 2
  int n_kblocks;         // total 128-byte-wide K blocks (64 bf16 or 32 tf32 elements)
  int kblock_elems;      // 64 (bf16) or 32 (tf32)
  int tf32;              // 1: fp32 rows read as tf32 operands (kind::tf32, K = 8 per MMA)
  int stages;
  float* partials;       // [gridDim.x][m_tiles*128][n_cols]
};

__global__ void __launch_bounds__(kThreads, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmap, GramParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: tiles (1024-aligned), then barriers
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* tiles = smem_raw + (base - smem_u32(smem_raw));
  const int tile_bytes = P.rows_box * 128;
  uint64_t* full = reinterpret_cast<uint64_t*>(tiles + (size_t)P.stages * tile_bytes);
  uint64_t* empty = full + P.stages;
  uint64_t* tmem_full = empty + P.stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // K-block range of this CTA
  const int per = (P.n_kblocks + gridDim.x - 1) / gridDim.x;
  const int kb0 = min(P.n_kblocks, (int)blockIdx.x * per);
  const int kb1 = min(P.n_kblocks, kb0 + per);
  const int nkb = kb1 - kb0;
  const uint32_t tmem_cols = (uint32_t)(P.m_tiles * P.n_cols <= 32 ? 32 : P.m_tiles * P.n_cols <= 64 ? 64
                                        : P.m_tiles * P.n_cols <= 128 ? 128 : P.m_tiles * P.n_cols <= 256 ? 256 : 512);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmap)) : "memory");
    for (int s = 0; s < P.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer =====
      for (int i = 0; i < nkb; ++i) {
        const int st = i % P.stages;
        const uint32_t ph = (uint32_t)(i / P.stages) & 1u;
        mbar_wait(empty + st, ph ^ 1u);
        mbar_expect_tx(full + st, (uint32_t)tile_bytes);
        tma_load_2d(tiles + (size_t)st * tile_bytes, &tmap, (kb0 + i) * P.kblock_elems, 0, full + st);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nkb > 0) {
      // ===== MMA issuer (single thread) =====
      const uint32_t idesc = make_idesc(128, P.n_cols, P.tf32 ? 2u : 1u);
      for (int i = 0; i < nkb; ++i) {
        const int st = i % P.stages;
        const uint32_t ph = (uint32_t)(i / P.stages) & 1u;
        mbar_wait(full + st, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t tile_addr = base + (uint32_t)(st * tile_bytes);
        for (int mt = 0; mt < P.m_tiles; ++mt) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // advance 32 bytes (16 bf16) along K inside the 128B swizzle row: +2 in the (addr>>4) field
            const uint64_t adesc = make_desc(tile_addr + (uint32_t)(mt * 128 * 128) + (uint32_t)(k * 32));
            const uint64_t bdesc = make_desc(tile_addr + (uint32_t)(k * 32));
            // 32 bytes per MMA along K in both modes: 16 bf16 (kind::f16) or 8 tf32 (kind::tf32)
            if (P.tf32) umma_tf32(tmem_base + (uint32_t)(mt * P.n_cols), adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
            else umma_bf16(tmem_base + (uint32_t)(mt * P.n_cols), adesc, bdesc, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(empty + st);            // frees the smem stage when these MMAs retire
      }
      umma_commit(tmem_full);               // accumulator complete
    }
  } else {
    // ===== epilogue warps 2..5: TMEM -> registers -> partial Gram in global memory =====
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    float* out = P.partials + (size_t)blockIdx.x * (size_t)(P.m_tiles * 128) * P.n_cols;
    if (nkb > 0) {
      mbar_wait(tmem_full, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      for (int mt = 0; mt < P.m_tiles; ++mt) {
        const int row = mt * 128 + q * 32 + lane;
        for (int c0 = 0; c0 < P.n_cols; c0 += 16) {
          uint32_t r[16];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * P.n_cols + c0);
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
              : "r"(taddr));
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          float4* dst = reinterpret_cast<float4*>(out + (size_t)row * P.n_cols + c0);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                 __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        }
      }
    } else {
      for (int mt = 0; mt < P.m_tiles; ++mt) {
        const int row = mt * 128 + q * 32 + lane;
        for (int c = 0; c < P.n_cols; ++c) out[(size_t)row * P.n_cols + c] = 0.f;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
  }
}

// G[i][j] = sum_b partials[b][i][j]  (fp32, deterministic order);  writes the top-left R x R block.
__global__ void gram_reduce_kernel(const float* partials, int n_parts, int rows_p, int n_cols, int R, float* G) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= R * R) return;
  const int i = idx / R, j = idx - i * R;
  float acc = 0.f;
  const size_t stride = (size_t)rows_p * n_cols;
  for (int b = 0; b < n_parts; ++b) acc += partials[b * stride + (size_t)i * n_cols + j];
  G[idx] = acc;
}

// fp32 rows -> bf16 (hi) and optionally the bf16 residual (lo) for the 3-MMA exactness mode; zero padding.
__global__ void cast_rows_kernel(const float* a, int R, int64_t D, int64_t lda, __nv_bfloat16* hi, __nv_bfloat16* lo,
                                 int rows_box, int64_t Dp) {
  const int64_t n = (int64_t)rows_box * Dp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / Dp, c = i - r * Dp;
    float v = (r < R && c < D) ? a[r * lda + c] : 0.f;
    const __nv_bfloat16 h = __float2bfloat16(v);
    hi[i] = h;
    if (lo) lo[i] = __float2bfloat16(v - __bfloat162float(h));
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode() {
  static EncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && p) fn = (EncodeFn)p;
  }
  return fn;
}

// bf16: operand copy [rows_box][Dp] (zero padded).  fp32/tf32: the caller's matrix itself, [R][D] with row stride lda;
// rows >= R and columns >= D of a box are zero-filled by TMA (out-of-bounds fill), so no padded copy is needed.
int make_tmap(CUtensorMap* map, const void* base, int rows_box, int64_t cols, int64_t rows, int64_t ld_elems, bool f32) {
  EncodeFn enc = get_encode();
  if (!enc) { afl_rt::set_error("cuTensorMapEncodeTiled not available"); return -5; }
  const int esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld_elems * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)rows_box};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { afl_rt::set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return -6; }
  return 0;
}

struct Geometry { int rows_box, n_cols, m_tiles, stages, grid, kelems; int64_t Dp; size_t smem, partial_floats; };

int geometry(int R, int64_t D, Geometry* g, int kelems = kBlockK) {
  if (R < 1 || R > kMaxRows) { afl_rt::set_error("gram: 1 <= R <= 256 required (got %d)", R); return -1; }
  g->rows_box = R <= 128 ? 128 : 256;
  g->m_tiles = g->rows_box / 128;
  g->n_cols = ((R + 15) / 16) * 16;
  g->kelems = kelems;
  g->Dp = ((D + kelems - 1) / kelems) * kelems;
  g->stages = g->rows_box == 128 ? 8 : 4;
  const int nkb = (int)(g->Dp / kelems);
  const int sms = afl_rt::num_sms();
  g->grid = nkb < sms ? nkb : sms;
  g->smem = 1024 + (size_t)g->stages * g->rows_box * 128 + (2 * g->stages + 1) * 8 + 16;
  g->partial_floats = (size_t)g->grid * g->rows_box * g->n_cols;
  return 0;
}

int launch_gram(const void* base, int64_t cols, int64_t rows, int64_t ld, bool f32, const Geometry& g, float* partials,
                cudaStream_t st) {
  CUtensorMap map;
  int rc = make_tmap(&map, base, g.rows_box, cols, rows, ld, f32);
  if (rc) return rc;
  AFL_CUDA(cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
  GramParams P{g.rows_box, g.n_cols, g.m_tiles, (int)(g.Dp / g.kelems), g.kelems, f32 ? 1 : 0, g.stages, partials};
  gram_kernel<<<g.grid, kThreads, g.smem, st>>>(map, P);
  afl_rt::count_launch();
  return afl_rt::check_last("gram_kernel");
}

}  // namespace

extern "C" {

// Workspace sizes for R rows of length D: bf16 operand elements (per copy) and fp32 partial elements.
int afl_gram_workspace(int R, int64_t D, int64_t* bf16_elems, int64_t* partial_floats) {
  Geometry g;
  int rc = geometry(R, D, &g);
  if (rc) return rc;
  *bf16_elems = (int64_t)g.rows_box * g.Dp;
  *partial_floats = (int64_t)g.partial_floats;
  return 0;
}

// G [R][R] = A A^T with A [R][lda>=D] fp32 (centred rows).  a16: bf16 workspace (>= bf16_elems), partials: fp32
// scratch (>= partial_floats).  Single bf16 pass (full tensor-core rate; SURVEY §2.8: 3e-5 relative error on
// distances of centred rows).
int afl_gram_bf16(const float* A, int R, int64_t D, int64_t lda, void* a16, float* partials, float* G, void* stream) {
  Geometry g;
  int rc = geometry(R, D, &g);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = (int64_t)g.rows_box * g.Dp;
  int cg = (int)((n + 255) / 256);
  if (cg > afl_rt::num_sms() * 16) cg = afl_rt::num_sms() * 16;
  cast_rows_kernel<<<cg, 256, 0, st>>>(A, R, D, lda, (__nv_bfloat16*)a16, nullptr, g.rows_box, g.Dp);
  afl_rt::count_launch();
  rc = launch_gram(a16, g.Dp, g.rows_box, g.Dp, false, g, partials, st);
  if (rc) return rc;
  gram_reduce_kernel<<<(R * R + 255) / 256, 256, 0, st>>>(partials, g.grid, g.rows_box, g.n_cols, R, G);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_gram_bf16");
}

// G = A A^T with the fp32 rows used directly as TF32 tensor-core operands: no cast pass and no operand copy (TMA
// reads the caller's matrix and zero-fills rows >= R / columns >= D), one pass over R*D*4 bytes; SURVEY §2.8 probe:
// 4.6e-6 relative error on distances of centred rows.  A must be 16-byte aligned with lda % 4 == 0.
int afl_gram_tf32(const float* A, int R, int64_t D, int64_t lda, float* partials, float* G, void* stream) {
  Geometry g;
  int rc = geometry(R, D, &g, 32);
  if (rc) return rc;
  AFL_CHECK_ARG((lda % 4) == 0 && (reinterpret_cast<uintptr_t>(A) % 16) == 0, "afl_gram_tf32: A must be 16-byte aligned, lda % 4 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  rc = launch_gram(A, D, R, lda, true, g, partials, st);
  if (rc) return rc;
  gram_reduce_kernel<<<(R * R + 255) / 256, 256, 0, st>>>(partials, g.grid, g.rows_box, g.n_cols, R, G);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_gram_tf32");
}

int afl_gram_tf32_workspace(int R, int64_t D, int64_t* partial_floats) {
  Geometry g;
  int rc = geometry(R, D, &g, 32);
  if (rc) return rc;
  *partial_floats = (int64_t)g.partial_floats;
  return 0;
}

}  // extern "C"
