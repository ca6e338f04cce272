// Persistent multi-client training kernel and batched inference kernel (sm_100a).
//
// One launch = one FL round of local training for a chunk of clients (reference: client.py:75-111
// run once per client process).  Grid = clients x G blocks; the G blocks of a client split every
// batch by samples, keep activations in shared memory, write partial parameter gradients to their
// own row of `gpart` (L2-resident), meet at a per-client barrier, and each block then reduces +
// Adam-updates its 1/G slice of the client's parameter row.  No host involvement inside a round:
// sample indices, step counts, losses and the NaN flag live on the device.
#include <cuda_runtime.h>
#include <stdlib.h>

#include "../include/attackfl_b200.h"
#include "runtime.h"
#include "train_core.h"

namespace {

constexpr int kThreads = 512;

struct TrainParams {
  const AflOp* ops; int n_ops; int pool_floats;
  float* w; float* m; float* v; float* gpart; int64_t d_pad;
  const float* data; int rec;
  const int32_t* idx; int64_t idx_stride_c, idx_stride_e;
  const int32_t* n_data; const int32_t* client_ids;
  int n_clients, G, M, S, batch, epochs;
  int opt_kind; float lr, beta1, beta2, eps, momentum; int clip_proper; float clip;
  uint32_t seed_lo, seed_hi; int dropout;
  float* loss_out;        // [C][2]
  int32_t* steps_out;     // [C]
  uint32_t* bar;          // [C]
  float* gnorm;           // [C][2]
  int use_cluster;        // 1: the G blocks of a client form one thread-block cluster (hardware barrier)
  int chain;              // 1: chain mode (per-sample warp teams + deferred parameter gradients)
};

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Barrier among the G co-resident blocks of one client (monotonic counter, release/acquire at gpu scope).  The spin is
// bounded like the cross-GPU barriers of comm.cu: after ~2 s without progress the block raises the NaN flag of the
// client (loss_out[2c+1], which fails the round on the host) and traps, so a protocol bug cannot hang the GPU.
__device__ __forceinline__ void client_barrier(uint32_t* ctr, uint32_t target, float* nan_flag) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    const long long t0 = clock64();
    uint32_t it = 0;
    while (ld_acquire(ctr) < target) {
      if ((++it & 1023u) == 0u && clock64() - t0 > 4000000000LL) { atomicAdd(nan_flag, 1.0f); __threadfence_system(); __trap(); }
    }
    __threadfence();
  }
  __syncthreads();
}

// Hardware barrier of a thread-block cluster with release/acquire semantics at cluster scope: global-memory writes made
// before the barrier by any block of the cluster are visible to every block of the cluster after it.
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

#define AFL_CLIENT_BARRIER() do { if (P.use_cluster) { __threadfence(); cluster_barrier(); } \
  else { bar_target += P.G; client_barrier(bar, bar_target, P.loss_out + 2 * c + 1); } } while (0)

__global__ void __launch_bounds__(kThreads, 1) train_kernel(TrainParams P) {
  extern __shared__ float4 smem4[];
  float* pool = reinterpret_cast<float*>(smem4);
  AflOp* ops = reinterpret_cast<AflOp*>(pool + P.pool_floats);
  AflOp* ops_s = ops + P.n_ops;                               // chain mode: [S][n_ops] per-sample shifted programs
  float* red = reinterpret_cast<float*>(ops_s + (P.chain ? P.S * P.n_ops : 0));   // [kThreads] scratch, loss_acc[2]
  float* loss_acc = red + kThreads;
  const int tid = threadIdx.x;
  const int c = blockIdx.x / P.G, g = blockIdx.x - c * P.G;

  {  // program -> shared memory; zero this block's partial-gradient row and the pool
    const int32_t* src = reinterpret_cast<const int32_t*>(P.ops);
    int32_t* dst = reinterpret_cast<int32_t*>(ops);
    for (int i = tid; i < P.n_ops * AFL_OP_WORDS; i += kThreads) dst[i] = src[i];
    for (int i = tid; i < P.pool_floats; i += kThreads) pool[i] = 0.f;
    if (tid < 2) loss_acc[tid] = 0.f;
    if (P.chain) {
      __syncthreads();
      for (int i = tid; i < P.S * P.n_ops; i += kThreads) afl_shift_op(ops[i % P.n_ops], i / P.n_ops, ops_s[i]);
    }
  }
  float* gw = P.gpart + ((int64_t)c * P.G + g) * P.d_pad;
  for (int64_t i = tid; i < P.d_pad; i += kThreads) gw[i] = 0.f;
  __syncthreads();

  float* w = P.w + (int64_t)c * P.d_pad;
  float* m = P.m + (int64_t)c * P.d_pad;
  float* v = P.v ? P.v + (int64_t)c * P.d_pad : nullptr;
  const float* gp = P.gpart + (int64_t)c * P.G * P.d_pad;
  const int n = P.n_data[c];
  const int steps_per_epoch = (n + P.batch - 1) / P.batch;
  const int64_t slice = P.d_pad / P.G;
  const int64_t lo = g * slice, hi = lo + slice;
  uint32_t bar_target = 0;
  uint32_t* bar = P.bar + c;
  double b1t = 1.0, b2t = 1.0;
  int opt_steps = 0, gstep = 0;

  AflTile t;
  t.pool = pool; t.w = w; t.gw = gw; t.data = P.data; t.rec = P.rec; t.probs = nullptr;
  t.loss_acc = loss_acc; t.row0 = 0; t.S = P.S; t.train = P.dropout;
  t.seed_lo = P.seed_lo; t.seed_hi = P.seed_hi; t.client = (uint32_t)P.client_ids[c];

  for (int e = 0; e < P.epochs; ++e) {
    const int32_t* idx_e = P.idx + (int64_t)c * P.idx_stride_c + (int64_t)e * P.idx_stride_e;
    for (int s = 0; s < steps_per_epoch; ++s, ++gstep) {
      const int b = min(P.batch, n - s * P.batch);
      if (b == 1) continue;                      // client.py:86-87: batches of size 1 are skipped
      t.step = (uint32_t)gstep;
      t.inv_b = 1.0f / (float)b;
      for (int mt = 0; mt < P.M; ++mt) {
        const int base = (g * P.M + mt) * P.S;
        const int nv = max(0, min(P.S, b - base));
        if (nv == 0 && mt > 0) break;            // tile 0 always runs so stale partials are overwritten
        t.idx = idx_e + (int64_t)s * P.batch + base;
        t.n_valid = nv; t.sample_base = base; t.accumulate = mt > 0;
        if (P.chain) {
          afl_run_tile_chain(ops, ops_s, P.n_ops, t, tid, kThreads);
        } else {
          afl_run_forward(ops, P.n_ops, t, tid, kThreads);
          afl_run_backward(ops, P.n_ops, t, tid, kThreads);
        }
      }
      if (P.G > 1) AFL_CLIENT_BARRIER();

      ++opt_steps;
      b1t *= (double)P.beta1; b2t *= (double)P.beta2;
      AflOpt o;
      o.kind = P.opt_kind; o.lr = P.lr; o.beta1 = P.beta1; o.beta2 = P.beta2; o.eps = P.eps;
      o.momentum = P.momentum; o.bc1 = (float)(1.0 - b1t); o.bc2 = (float)(1.0 - b2t);
      o.first_step = opt_steps == 1; o.clip_coef = 1.f;
      if (P.clip_proper) {
        // torch.nn.utils.clip_grad_norm_ semantics on the *reduced* gradient (opt-in)
        float part = afl_reduce_sqnorm(gp, P.G, P.d_pad, lo, hi, tid, kThreads);
        red[tid] = part;
        __syncthreads();
        float* gn = P.gnorm + 2 * c;
        const int par = opt_steps & 1;
        if (tid == 0) {
          float sum = 0.f;
          for (int i = 0; i < kThreads; ++i) sum += red[i];
          atomicAdd(gn + par, sum);
        }
        AFL_CLIENT_BARRIER();
        const float norm = sqrtf(*reinterpret_cast<volatile float*>(gn + par));
        float coef = P.clip / (norm + 1e-6f);
        o.clip_coef = coef < 1.f ? coef : 1.f;
        if (g == 0 && tid == 0) gn[par ^ 1] = 0.f;   // next step's accumulator
      }
      afl_reduce_update(w, m, v, gp, P.G, P.d_pad, lo, hi, o, tid, kThreads);
      if (P.G > 1) AFL_CLIENT_BARRIER(); else __syncthreads();
    }
  }
  if (tid == 0) {
    atomicAdd(P.loss_out + 2 * c, loss_acc[0]);
    atomicAdd(P.loss_out + 2 * c + 1, loss_acc[1]);
    if (g == 0) P.steps_out[c] = opt_steps;
  }
}

struct InferParams {
  const AflOp* ops; int n_ops; int pool_floats;
  const float* w; int64_t d_pad; int n_models;
  const float* data; int rec; int64_t n_rows;
  float* probs;            // [n_models][n_rows]
  int S; int tiles_per_model;
};

__global__ void __launch_bounds__(kThreads, 1) infer_kernel(InferParams P) {
  extern __shared__ float4 smem4[];
  float* pool = reinterpret_cast<float*>(smem4);
  AflOp* ops = reinterpret_cast<AflOp*>(pool + P.pool_floats);
  const int tid = threadIdx.x;
  {
    const int32_t* src = reinterpret_cast<const int32_t*>(P.ops);
    int32_t* dst = reinterpret_cast<int32_t*>(ops);
    for (int i = tid; i < P.n_ops * AFL_OP_WORDS; i += kThreads) dst[i] = src[i];
    for (int i = tid; i < P.pool_floats; i += kThreads) pool[i] = 0.f;
  }
  __syncthreads();
  const int64_t total = (int64_t)P.n_models * P.tiles_per_model;
  AflTile t;
  t.pool = pool; t.gw = nullptr; t.data = P.data; t.rec = P.rec; t.idx = nullptr; t.loss_acc = nullptr;
  t.S = P.S; t.train = 0; t.accumulate = 0; t.inv_b = 0.f; t.seed_lo = t.seed_hi = t.client = t.step = 0;
  t.sample_base = 0;
  for (int64_t job = blockIdx.x; job < total; job += gridDim.x) {
    const int mdl = (int)(job / P.tiles_per_model);
    const int64_t tile = job - (int64_t)mdl * P.tiles_per_model;
    t.w = P.w + (int64_t)mdl * P.d_pad;
    t.probs = P.probs + (int64_t)mdl * P.n_rows;
    t.row0 = tile * P.S;
    t.n_valid = (int)min((int64_t)P.S, P.n_rows - t.row0);
    afl_run_forward(ops, P.n_ops, t, tid, kThreads);
  }
}

size_t train_smem(int pool_floats, int n_ops, int chain_S = 0) {
  return sizeof(float) * (size_t)pool_floats + sizeof(AflOp) * (size_t)n_ops * (size_t)(1 + chain_S) +
         sizeof(float) * (kThreads + 4);
}

}  // namespace

extern "C" {

// Number of training blocks that can be co-resident on the current device for this program.
int afl_train_capacity(int pool_floats, int n_ops, int chain_S) {
  const size_t smem = train_smem(pool_floats, n_ops, chain_S);
  if ((int)smem > afl_rt::max_smem_optin()) { afl_rt::set_error("training pool needs %zu B shared memory", smem); return 0; }
  if (cudaFuncSetAttribute(train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, train_kernel, kThreads, smem) != cudaSuccess) return 0;
  return per_sm * afl_rt::num_sms();
}

int afl_train_round(const int32_t* ops_dev, int n_ops, int pool_floats, float* w, float* m, float* v, float* gpart,
                    int64_t d_pad, const float* data, int rec, const int32_t* idx, int64_t idx_stride_c,
                    int64_t idx_stride_e, const int32_t* n_data, const int32_t* client_ids, int n_clients, int G,
                    int M, int S, int batch, int epochs, int opt_kind, float lr, float beta1, float beta2, float eps,
                    float momentum, int clip_proper, float clip, uint64_t seed, int dropout, float* loss_out,
                    int32_t* steps_out, uint32_t* bar, float* gnorm, int chain, void* stream) {
  AFL_CHECK_ARG(sizeof(AflOp) == 4 * AFL_OP_WORDS, "AflOp layout mismatch");
  AFL_CHECK_ARG(!chain || ((kThreads / 32) % S == 0 && S <= 15), "afl_train_round: chain mode needs S | 16 warps, S <= 15");
  AFL_CHECK_ARG(G >= 1 && M >= 1 && S >= 1 && G * M * S >= batch, "afl_train_round: G*M*S must cover the batch");
  AFL_CHECK_ARG(d_pad % (4 * (int64_t)G) == 0, "afl_train_round: d_pad must be divisible by 4*G");
  const size_t smem = train_smem(pool_floats, n_ops, chain ? S : 0);
  const int cap = afl_train_capacity(pool_floats, n_ops, chain ? S : 0);
  if (cap <= 0) return -2;
  if (n_clients * G > cap) {
    afl_rt::set_error("afl_train_round: %d blocks requested but only %d can be co-resident (chunk the clients)",
                      n_clients * G, cap);
    return -3;
  }
  TrainParams P;
  P.ops = reinterpret_cast<const AflOp*>(ops_dev); P.n_ops = n_ops; P.pool_floats = pool_floats;
  P.w = w; P.m = m; P.v = v; P.gpart = gpart; P.d_pad = d_pad; P.data = data; P.rec = rec;
  P.idx = idx; P.idx_stride_c = idx_stride_c; P.idx_stride_e = idx_stride_e; P.n_data = n_data;
  P.client_ids = client_ids; P.n_clients = n_clients; P.G = G; P.M = M; P.S = S; P.batch = batch; P.epochs = epochs;
  P.opt_kind = opt_kind; P.lr = lr; P.beta1 = beta1; P.beta2 = beta2; P.eps = eps; P.momentum = momentum;
  P.clip_proper = clip_proper; P.clip = clip; P.seed_lo = (uint32_t)seed; P.seed_hi = (uint32_t)(seed >> 32);
  P.dropout = dropout; P.loss_out = loss_out; P.steps_out = steps_out; P.bar = bar; P.gnorm = gnorm;
  P.chain = chain ? 1 : 0;
  cudaStream_t st = (cudaStream_t)stream;
  AFL_CUDA(cudaMemsetAsync(bar, 0, sizeof(uint32_t) * (size_t)n_clients, st));
  AFL_CUDA(cudaMemsetAsync(gnorm, 0, sizeof(float) * 2 * (size_t)n_clients, st));
  AFL_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float) * 2 * (size_t)n_clients, st));
  void* args[] = {&P};
  static int cluster_ok = -1;                 // -1 unknown, 0 unavailable (fall back to the software barrier), 1 ok
  // Measured on B200: the cluster launch runs the same step 2.2x SLOWER (193 vs 87 us) than the cooperative launch with
  // the software barrier, so it is opt-in (ATTACKFL_CLUSTER=1) and off by default.
  if (cluster_ok < 0) cluster_ok = getenv("ATTACKFL_CLUSTER") ? 1 : 0;
  P.use_cluster = 0;
  if (cluster_ok && G > 1 && G <= 16 && (G & (G - 1)) == 0) {
    // the G blocks of a client = one thread-block cluster: hardware co-scheduling + barrier.cluster (G = 16 is a
    // non-portable cluster size: 16 SMs of one GPC)
    cudaError_t e = cudaSuccess;
    if (G > 8) e = cudaFuncSetAttribute(train_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e == cudaSuccess) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(n_clients * G); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = G; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int max_clusters = 0;
      e = cudaOccupancyMaxActiveClusters(&max_clusters, train_kernel, &cfg);
      if (e == cudaSuccess && max_clusters >= 1) {
        P.use_cluster = 1;
        e = cudaLaunchKernelEx(&cfg, train_kernel, P);
        if (e == cudaSuccess) { afl_rt::count_launch(); return afl_rt::check_last("afl_train_round"); }
      }
    }
    (void)cudaGetLastError();                  // clusters not usable here: remember and use the software barrier
    cluster_ok = 0;
    P.use_cluster = 0;
  }
  // cooperative launch: the driver guarantees all blocks are co-resident, which the software barrier needs
  AFL_CUDA(cudaLaunchCooperativeKernel((void*)train_kernel, dim3(n_clients * G), dim3(kThreads), args, smem, st));
  afl_rt::count_launch();
  return afl_rt::check_last("afl_train_round");
}

int afl_infer(const int32_t* ops_dev, int n_ops, int pool_floats, const float* w, int64_t d_pad, int n_models,
              const float* data, int rec, int64_t n_rows, float* probs, int S, void* stream) {
  const size_t smem = train_smem(pool_floats, n_ops);
  AFL_CHECK_ARG((int)smem <= afl_rt::max_smem_optin(), "afl_infer: pool too large for shared memory");
  AFL_CUDA(cudaFuncSetAttribute(infer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  InferParams P;
  P.ops = reinterpret_cast<const AflOp*>(ops_dev); P.n_ops = n_ops; P.pool_floats = pool_floats; P.w = w;
  P.d_pad = d_pad; P.n_models = n_models; P.data = data; P.rec = rec; P.n_rows = n_rows; P.probs = probs; P.S = S;
  P.tiles_per_model = (int)((n_rows + S - 1) / S);
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, infer_kernel, kThreads, smem);
  if (per_sm < 1) per_sm = 1;
  int64_t grid = (int64_t)n_models * P.tiles_per_model;
  const int64_t cap = (int64_t)per_sm * afl_rt::num_sms();
  if (grid > cap) grid = cap;
  infer_kernel<<<(int)grid, kThreads, smem, (cudaStream_t)stream>>>(P);
  afl_rt::count_launch();
  return afl_rt::check_last("afl_infer");
}

}  // extern "C"
