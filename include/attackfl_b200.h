/* attackfl_b200 native library — C ABI.
 *
 * Two shared objects export this interface:
 *   attackfl_b200/lib/libattackfl_b200.so   CUDA build, sm_100a (pointers are device pointers, `stream` is a
 *                                           cudaStream_t; every call is asynchronous on that stream)
 *   attackfl_b200/lib/libattackfl_cpu.so    CPU twin compiled from the same kernel sources (pointers are host
 *                                           pointers, `stream` is ignored); backs `--device cpu` and the tests
 * All functions return 0 on success; otherwise afl_last_error() describes the failure.
 * Citations are relative to the reference repository filrg/attackFL.
 */
#ifndef ATTACKFL_B200_H_
#define ATTACKFL_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* afl_last_error(void);
long long afl_launch_count(void);          /* kernels launched by this library since the last reset */
void afl_reset_launch_count(void);
int afl_is_cuda_build(void);
int afl_num_sms(void);

/* T8 — flat optimisers over n = rows*d_pad coordinates (client.py:78,107: Adam(lr), re-created per round;
 * `momentum` is plumbed but unused in the reference, honoured here by afl_sgdm_flat).  gscale (device scalar,
 * nullable) multiplies the gradient (used for gradient clipping, client.py:104-105 / server.py:667-668). */
int afl_adam_flat(float* w, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
                  int step, const float* gscale, void* stream);
int afl_sgdm_flat(float* w, const float* g, float* buf, int64_t n, float lr, float momentum, int first_step,
                  const float* gscale, void* stream);

/* K1 + K2 — sample-weighted FedAvg reduce fused with the broadcast of the new global row to `Cb` client rows
 * (server.py:751-775 + the START fan-out at :588-614) and, in the same pass over the rows, the per-attacker
 * shifted moments -> mean / unbiased std / LIE row mu + z*sd (src/Utils.py:83-98,207-214).
 * rows [C][d_pad]; weights [C]; member [A][C] 0/1; kcount,z [A]; mu,sd,lie [A][d_pad]; A <= 4 per call. */
int afl_fedavg(const float* rows, const float* weights, int C, int64_t d_pad, float* out, float* bcast, int Cb,
               int A, const float* ref, const float* member, const float* kcount, const float* z, float* mu,
               float* sd, float* lie, int strict_k1, void* stream);

/* K1-fused — multi-GPU FedAvg over NVLink peer memory in one kernel per rank (server.py:751-775 reduce +
 * :588-614 broadcast).  Rank r owns coordinate slice r; the slice is cut into up to 32 chunks and every
 * block walks a three-stage software pipeline over them: local pre-reduce of this rank's rows (local_rows [CL][d_pad],
 * weights/member for the LOCAL clients) into its symmetric partial buffer, owner sum of the G partials through
 * peer-mapped pointers (part [world], each [1+2A][d_pad]) with the result (and the LIE rows of up to 4 attackers, lie
 * [world], [A][d_pad]) stored into every rank's buffers (glob [world]), fan-out to local_rows_out.  No grid-wide or
 * cross-GPU barrier: per-chunk flags in sig [world] (symmetric uint32 buffers of >= 1024 words, zero-initialised once;
 * st.release.sys / ld.acquire.sys, bounded spins).  ctrl: 65 zero-initialised local uint32 {[2][32] phase counters, error};
 * epoch increases by 1 per call.  part_mc / glob_mc / lie_mc: NVLS multicast addresses of the same buffers (null: P2P);
 * when given, the slice reduce is multimem.ld_reduce (in-switch sum) and the broadcast is multimem.st. */
int afl_fedavg_fused(const float* local_rows, const float* weights, int CL, int64_t d_pad, int world, int rank,
                     float* const* part, float* const* glob, float* const* lie, uint32_t* const* sig, uint32_t epoch,
                     int A, const float* ref, const float* member, const float* kcount, const float* z, int strict_k1,
                     float* local_rows_out, uint32_t* ctrl, float* part_mc, float* glob_mc, float* lie_mc, void* stream);

/* K3 multi-GPU — slice ownership of the Gram GEMM (src/Utils.py:83-98 mean/std + :108-113,142-147,176-185 pairwise
 * distances): afl_slice_moments pulls the k genuine rows of the owned K-slice [lo,hi) through peer-mapped pointers
 * (row_ptrs_dev [k], device array), forms mu / sigma / sign(mu) and writes the centred operand rows [k+2][ld] that
 * feed afl_gram_tf32 on the slice; the (k+2)^2 partial Grams are summed across ranks; afl_craft_scatter then stores
 * mu - gamma*p of the slice into every rank's crafted row (dst [world], peer-mapped). */
int afl_slice_moments(const float* const* row_ptrs_dev, int k, const float* center, int64_t lo, int64_t hi, int64_t ld,
                      int sign_mode, float* out, float* mu_out, float* p_out, void* stream);
int afl_craft_scatter(float* const* dst, int world, const float* mu, const float* p, const float* gamma_dev, int64_t lo,
                      int64_t hi, void* stream);

/* K4 — gamma bisection on the (k+2)x(k+2) Gram matrix of [g_1-c..g_k-c, mu-c, p] (src/Utils.py:115-130;
 * rule 0 = Min-Max / Opt-Fang, 1 = Min-Sum; exclude_first reproduces the aliasing of genuine model 0,
 * paper=1 returns gamma_succ).  out4 = {gamma, gamma_succ, threshold, probes} in device memory. */
int afl_gamma_search_dev(const float* G, int k, int rule, int exclude_first, int paper, float* out4, void* stream);
/* out = mu - gamma*p (sign_mode=1: p = sign(mu), Opt-Fang); gamma is read from device memory. */
int afl_craft_axpy(float* out, const float* mu, const float* p, const float* gamma_dev, int sign_mode, int64_t n,
                   void* stream);
/* Random attack: out = (w + N(0,1)*perturbation) * valid   (src/Utils.py:52-57), Philox + Box-Muller. */
int afl_craft_random(float* out, const float* w, const float* valid, float perturbation, uint64_t seed,
                     uint32_t stream_id, int64_t n, void* stream);

/* K6 — per-coordinate selection over C rows: mode 0 lower median (src/Utils.py:344-357), mode 1 trimmed mean
 * dropping `trim` values per side (src/Utils.py:267-302). */
int afl_coord_select(const float* rows, int C, int64_t d_pad, float* out, int mode, int trim, void* stream);
/* K6 slice-owned: the same selection on coordinates [lo, hi) with the C rows given as a device array of base pointers (rows
 * on other GPUs are read in place through peer-mapped pointers) and the result stored into every buffer of dst_dev [n_dst]
 * (every rank's copy of the global model): src/Utils.py:267-302,344-357 without gathering any row. */
int afl_coord_select_ptrs(const float* const* row_ptrs_dev, int C, int64_t lo, int64_t hi, float* const* dst_dev, int n_dst,
                          int mode, int trim, void* stream);

/* HAR linear layers on tcgen05 (kind::tf32 on fp32 data, fp32 accumulate): nn.Linear inside nn.TransformerEncoderLayer
 * and its autograd (src/Model.py:435-458, client.py:114-131).  All matrices row-major and contiguous, N and K multiples
 * of 32 in [32, 256], pointers 16-byte aligned.
 *   fwd : Y[M,N] = dropout(relu?(X[M,K] . W[N,K]^T + bias)); dropout keep decision = afl_dropout_keep's Philox stream
 *         (seed, rng_client, rng_step, sample = row, rng_id, elem = column).
 *   dx  : dX[M,K] (+)= (dY[M,N] . W[N,K]) * (mask_src[M,K] > 0 ? mask_scale : 0)   (mask_src NULL = no mask)
 *   dw  : dW[N,K] (+)= dY^T . X, db[N] (+)= column sums of dY (db NULL = skip); partials: scratch of
 *         afl_tc_linear_dw_workspace floats. */
int afl_tc_linear_fwd(const float* X, int64_t M, int K, const float* W, int N, const float* bias, int relu, float drop_p,
                      uint64_t seed, uint32_t rng_client, uint32_t rng_step, uint32_t rng_id, float* Y, void* stream);
/* fwd with the residual + dropout + LayerNorm fused into the epilogue (N = 64): Y = LN(resid + dropout(X W^T + bias)) * gamma
 * + beta, xhat [M,64] and rstd [M] saved for afl_har_ln_bwd; same dropout stream as afl_har_add_ln_fwd. */
int afl_tc_linear_ln_fwd(const float* X, int64_t M, int K, const float* W, const float* bias, const float* resid, float drop_p,
                         uint64_t seed, uint32_t rng_client, uint32_t rng_step, uint32_t rng_id, const float* gamma,
                         const float* beta, float* Y, float* xhat, float* rstd, void* stream);
int afl_tc_linear_dx(const float* dY, int64_t M, int N, const float* W, int K, const float* mask_src, float mask_scale,
                     int accumulate, float* dX, void* stream);
int afl_tc_linear_dw_workspace(int N, int K, int64_t* partial_floats);
int afl_tc_linear_dw(const float* dY, const float* X, int64_t M, int N, int K, float* partials, float* dW, float* db,
                     int accumulate, void* stream);

/* HAR TransformerClassifier training kernels (src/Model.py:420-458, client.py:114-131); activations are [B*T, 64]
 * row-major with row = b*T + t; d_model 64, 4 heads of 16.  Dropout masks are Philox functions of
 * (seed, client, step, row, site, column) and are recomputed in the backward kernels.
 *   embed_fwd : out = Conv1d(1->64, k3, p1)(x) + pe[t]          embed_bwd : dconv_w, dconv_b += ...
 *   add_ln_fwd: y = LayerNorm(x + dropout(r)); saves xhat, rstd   (r NULL: plain LayerNorm)
 *   ln_bwd    : dres = d/d(x + dropout(r)); dr = dres * mask (NULL: skip); dgamma, dbeta += ...;
 *               dy_bcast [B,64] with bscale: dy[row] = dy_bcast[row / T] * bscale (mean-pool backward fused)
 *   attn_fwd  : softmax(q k^T / 4) with attention-weight dropout, times v, per (sample, head); lse in base-2 units
 *   attn_bwd  : dqkv [B*T,192] from dout (two kernels: dQ per query, dK/dV per key)
 *   pool_fwd  : pooled[b] = mean_t h[b,t]
 *   head      : Linear(64,64)-ReLU-Dropout-Linear(64,n_cls) + mean cross-entropy, forward and backward in one
 *               kernel; out3 += {sum of per-sample losses, NaN/Inf count, correct predictions}. */
int afl_har_embed_fwd(const float* x, const float* conv_w, const float* conv_b, const float* pe, int B, int T, float* out,
                      void* stream);
int afl_har_embed_bwd(const float* dh, const float* x, int B, int T, float* dconv_w, float* dconv_b, void* stream);
int afl_har_add_ln_fwd(const float* x, const float* r, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                       uint32_t site, const float* gamma, const float* beta, int64_t M, float* y, float* xhat, float* rstd,
                       void* stream);
int afl_har_ln_bwd(const float* dy, const float* dy_bcast, int T, float bscale, const float* xhat, const float* rstd,
                   const float* gamma, float drop_p, uint64_t seed, uint32_t client, uint32_t step, uint32_t site, int64_t M,
                   float* dres, float* dr, float* dgamma, float* dbeta, void* stream);
int afl_har_attn_fwd(const float* qkv, int B, int T, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                     uint32_t site, float* out, float* lse, void* stream);
int afl_har_attn_bwd(const float* qkv, const float* out, const float* dout, const float* lse, int B, int T, float drop_p,
                     uint64_t seed, uint32_t client, uint32_t step, uint32_t site, float* dqkv, void* stream);
/* attn_fwd on the tensor cores (csrc/attn_tcgen05.cu): same contract and same dropout stream as afl_har_attn_fwd, scores
 * and probabilities as TF32 tcgen05 operands; afl_har_attn_tc_supported(T) says whether K/V of one (sample, head pair)
 * fit in shared memory (T <= 568 with the 227 KB limit). */
int afl_har_attn_tc_supported(int T);
int afl_har_attn_fwd_tc(const float* qkv, int B, int T, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                        uint32_t site, float* out, float* lse, uint32_t* keep_mask, void* stream);
/* lse (and the backward's delta) rows are lse_stride(T) = T rounded up to 64 floats long: [B][4][lse_stride(T)].
 * keep_mask (NULL = not wanted): [B][4][afl_har_attn_mask_words(T)][lse_stride(T)] uint32, bit (key & 31) of word
 * [key >> 5][query] = that attention weight survived the dropout; consumed by afl_har_attn_bwd_tc. */
int afl_har_lse_stride(int T);
int afl_har_attn_mask_words(int T);
/* attention backward on the tensor cores (csrc/attn_bwd_tcgen05.cu): dqkv [B*T,192] from dout; needs lse and
 * (drop_p > 0) keep_mask written by afl_har_attn_fwd_tc; delta: scratch [B][4][lse_stride(T)]. */
int afl_har_attn_bwd_tc(const float* qkv, const float* out, const float* dout, const float* lse, const uint32_t* keep_mask,
                        int B, int T, float drop_p, float* delta, float* dqkv, void* stream);
int afl_har_pool_fwd(const float* h, int B, int T, float* pooled, void* stream);
int afl_har_head(const float* pooled, const float* w1, const float* b1, const float* w2, const float* b2,
                 const int64_t* labels, int B, int n_cls, float drop_p, uint64_t seed, uint32_t client, uint32_t step,
                 uint32_t site, int train, float* logits, float* out3, float* dpooled, float* dw1, float* db1, float* dw2,
                 float* db2, void* stream);

/* ScionFL building blocks (server.py:436-492, src/Utils.py:372-387) on rows of the flat arena.  `valid` is the 0/1
 * padding mask of the layout (NULL = every coordinate counts); words = d_pad / 32.
 *   afl_scion_minmax   : minmax[c] = {min, max} of row c over valid coordinates.
 *   afl_scion_quantize : bits[c][w] packs sigma ~ Bernoulli((x - min) / (max - min + 1e-6)) (Philox keyed by seed, c,
 *                        coordinate); counts[c] = number of set bits.
 *   afl_scion_scores   : from the packed bits and the (possibly L2-clipped) levels lohi[c] = {lo, hi}:
 *                        out[c] = sum_j sigma_cj * agg_j, out[C] = sum_j agg_j, out[C+1] = sum_j agg_j^2, where
 *                        agg_j = mean_c (lo_c + sigma_cj (hi_c - lo_c)); the dequantised rows are never built. */
int afl_scion_minmax(const float* rows, const float* valid, int C, int64_t d_pad, float* minmax, void* stream);
int afl_scion_quantize(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                       uint32_t* bits, int* counts, void* stream);
/* Tie-aware ROC-AUC of n <= 8192 scores (= sklearn auc(roc_curve), src/Validation.py:113-120) in one launch:
 * out4 = {auc (NaN if one class is empty or a score is NaN), number of NaN scores, positives, negatives}. */
int afl_roc_auc(const float* scores, const float* labels, int64_t n, float* out4, void* stream);
/* Per-round sample indices of C clients (RpcClient.py:166,169): idx [C][E][mx] int32, entry (c, e, i) = the record client
 * client_ids[c] visits at position i of epoch e, -1 for i >= sizes[c].  Subset and per-epoch order are keyed permutations
 * (Feistel network + cycle walking) of (seed, client, epoch): no sort, no state, identical on the CPU twin and the GPU. */
int afl_round_indices(int32_t* idx, const int32_t* sizes, const int32_t* client_ids, int C, int E, int mx, int n_rows,
                      uint64_t seed, void* stream);
/* N-GPU forms (CUDA build): _quantize_rows keys the Philox stream of local row c by row0 + c (its position in the round's
 * reporting list), so every rank draws the bits the one-GPU pass would; _scores_ptrs takes the packed rows by address
 * (bits_ptrs_dev [C], peer-mapped allowed) and visits the words [w_lo, w_hi) only: partial sums of an owned slice. */
int afl_scion_quantize_rows(const float* rows, const float* valid, const float* minmax, int C, int64_t d_pad, uint64_t seed,
                            int row0, uint32_t* bits, int* counts, void* stream);
int afl_scion_scores_ptrs(const uint32_t* const* bits_ptrs_dev, const float* valid, const float* lohi, int C, int64_t d_pad,
                          int64_t w_lo, int64_t w_hi, float* out, void* stream);
int afl_scion_scores(const uint32_t* bits, const float* valid, const float* lohi, int C, int64_t d_pad, float* out,
                     void* stream);

/* K5a — hyper-network emit: theta[c] = W_h . MLP(E[clients[c]]) + b_h for n clients (src/Model.py:285-304,
 * server.py:589-590,674-678).  small = params|m|v of {emb, mlp}; acts [E+5H] (n==1 only); f_scratch [n*H]. */
int afl_hyper_emit(const float* head_w, const float* head_b, float* small, int n_nodes, int E, int H,
                   const int* clients_dev, int n, float* f_scratch, float* acts, float* theta, int64_t theta_ld,
                   int64_t d_pad, void* stream);
/* K5b — one sequential pFedHN step for `client` (server.py:644-670): delta = h(client) - theta_client, closed-form
 * rank-1 head gradient, MLP/embedding backward, global-norm clip, dense Adam.  red [H+3] scratch:
 * red[H+1] = clip coefficient, red[H+2] = pre-clip gradient norm. */
int afl_hyper_step(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, float* small,
                   int n_nodes, int E, int H, int client, const int* client_dev, const float* theta_client,
                   float* delta, float* f_scratch, float* acts, float* red, int64_t d_pad, float lr, float b1,
                   float b2, float eps, int step, float clip, void* stream);

/* K3 — split-K Gram matrix G [R][R] = A A^T on tcgen05 tensor cores (TMA -> 128B-swizzled smem -> tcgen05.mma ->
 * TMEM -> tcgen05.ld), R <= 256 (src/Utils.py:30-49,108-113,142-147,176-185 pairwise distances; :326-342 Krum;
 * server.py:310-350 ShieldFL; :703-743 FLTrust).  afl_gram_tf32 reads the fp32 rows directly as TF32 operands (no
 * cast pass, TMA zero-fills the padding); afl_gram_bf16 first casts to a zero-padded bf16 copy (a16).  partials:
 * split-K scratch, sizes from the *_workspace calls. */
int afl_gram_workspace(int R, int64_t D, int64_t* bf16_elems, int64_t* partial_floats);
int afl_gram_bf16(const float* A, int R, int64_t D, int64_t lda, void* a16, float* partials, float* G, void* stream);
int afl_gram_tf32_workspace(int R, int64_t D, int64_t* partial_floats);
int afl_gram_tf32(const float* A, int R, int64_t D, int64_t lda, float* partials, float* G, void* stream);

/* One Gram for all attackers (src/Utils.py:83-98,108-113,142-147,176-185; the per-attacker samples of server.py:596-600 all
 * come from the same n genuine rows).  afl_multi_moments: one pass over the genuine rows of slice [lo, hi) (row_ptrs_dev [n],
 * peer-mapped pointers allowed) -> X [n + A][ld] = centred genuine rows followed by the direction rows p_a, mu_out / p_out
 * [A][ld]; member_dev [A][n] 0/1, kcount_dev [A], sign_mode_dev [A] (1: p = sign(mu), Opt-Fang).  afl_multi_gamma: G [R][R]
 * (R = n + A) -> per attacker the (k+2)^2 sub-Gram of its sample (subsets_dev [A][kmax]: positions in the genuine list, in
 * sample order; k_dev [A]; rule_dev [A]: 0 max / 1 sum) and the reference's bisection, out [A][4] as afl_gamma_search_dev.
 * afl_multi_craft: mu_a - gamma_a p_a of slice [lo, hi) stored to dst_dev [A][world] row pointers (null entries skipped). */
int afl_multi_moments(const float* const* row_ptrs_dev, int n, const float* center, int64_t lo, int64_t hi, int64_t ld,
                      const float* member_dev, const float* kcount_dev, const int* sign_mode_dev, int A, float* X,
                      float* mu_out, float* p_out, void* stream);
int afl_multi_gamma(const float* G, int R, int n, const int* subsets_dev, const int* k_dev, int kmax, const int* rule_dev,
                    int A, int paper, float* out, void* stream);
int afl_multi_craft(float* const* dst_dev, int world, const float* mu, const float* p, int64_t ld, const float* gammas,
                    int A, int64_t lo, int64_t hi, void* stream);

/* K5 persistent + row-sharded: one cooperative launch per rank walks all reporting clients of a round in order
 * (server.py:637-678).  head_w / head_b / mw / vw / mb / vb are THIS rank's row shard (row 0 = global row lo, `rows` rows);
 * `small` = params | m | v of the replicated MLP + embedding table, `small_alt` [tot] a second copy of the parameters
 * (double buffer inside the launch).  order_dev [n]: client ids; theta_dev [n]: device array of the clients' row base
 * pointers (may be peer-mapped: the kernel reads rows [lo, lo + rows) where they live).  Multi-GPU: xchg / sig are host
 * arrays of `world` peer-mapped symmetric buffers (afl_hyper_round_xchg_floats / _sig_words for n_max clients), epoch
 * increases by 1 per call; world = 1: both may be null.  delta [rows], red [2][H+1], bar [2] are local scratch (bar[1] != 0
 * after the call: a spin timed out).  step0: Adam steps taken before this call.  norms [n]: pre-clip gradient norms. */
int64_t afl_hyper_round_xchg_floats(int n_max, int world, int H);
int64_t afl_hyper_round_sig_words(int n_max, int world);
int afl_hyper_round(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, int64_t rows, int64_t lo,
                    float* small, float* small_alt, int n_nodes, int E, int H, const int* order_dev,
                    const float* const* theta_dev, int n, float* delta, float* red, float* const* xchg,
                    uint32_t* const* sig, int world, int rank, int n_max, uint32_t epoch, uint32_t* bar, int step0, float lr,
                    float b1, float b2, float eps, float clip, float* norms, void* stream);
/* K5a scatter: theta_j[lo + d] = W_h[d] . MLP(E[clients[j]]) + b_h[d] on the owned rows, stored through the pointer table
 * theta_dev [n] (peer-mapped pointers scatter the personalised models to their owner ranks, server.py:589-590,674-678). */
int afl_hyper_emit_rows(const float* head_w, const float* head_b, int64_t rows, int64_t lo, const float* small, int n_nodes,
                        int E, int H, const int* clients_dev, int n, float* f_scratch, float* const* theta_dev,
                        void* stream);

/* Allocation-free variant of afl_hyper_step (no counterpart in the reference: it exists for callers of THIS
 * library that may not allocate device memory -- pre-reserved real-time loops, plugins inside another runtime's
 * memory budget, host callbacks that must not trigger a device synchronisation).  Identical arithmetic and
 * outputs; the scratch (delta, features, activations, reduction slots) lives in a file-scope device array inside
 * the library: limits d_pad <= 262144, H <= 128, E <= 64 (1.006 MiB per GPU, held until process exit); larger sizes
 * are rejected with an error.  The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress) and should therefore be made before any stream capture.  All calls share the array:
 * calls that overlap on the GPU race.  coef_norm_out (device, [2], nullable) receives {clip coefficient, pre-clip
 * gradient norm}. */
int afl_hyper_step_noalloc(float* head_w, float* head_b, float* mw, float* vw, float* mb, float* vb, float* small,
                           int n_nodes, int E, int H, int client, const int* client_dev, const float* theta_client,
                           int64_t d_pad, float lr, float b1, float b2, float eps, int step, float clip,
                           float* coef_norm_out, void* stream);

/* T1-T7 — local training of a chunk of clients for one round as ONE persistent kernel (client.py:75-111;
 * models src/Model.py:27-246).  ops: int32 program [n_ops][32] (attackfl_b200/models/programs.py);
 * w,m,v [C][d_pad]; gpart [C][G][d_pad]; data [rows][rec]; idx [C][epochs][>=n] int32; G blocks per client,
 * M tiles of S samples per block (G*M*S >= batch).  loss_out [C][2] = {sum of per-sample losses, #NaN losses},
 * steps_out [C]; bar [C] and gnorm [C][2] are scratch.  opt_kind 0 Adam / 1 SGD-momentum.  chain=1 runs each sample's
 * forward/backward chain on its own warp team with a deferred parameter-gradient pass (program must be planned without
 * gradient-buffer reuse; chain_S = S for afl_train_capacity, 0 otherwise). */
int afl_train_capacity(int pool_floats, int n_ops, int chain_S);
int afl_train_round(const int32_t* ops_dev, int n_ops, int pool_floats, float* w, float* m, float* v, float* gpart,
                    int64_t d_pad, const float* data, int rec, const int32_t* idx, int64_t idx_stride_c,
                    int64_t idx_stride_e, const int32_t* n_data, const int32_t* client_ids, int n_clients, int G,
                    int M, int S, int batch, int epochs, int opt_kind, float lr, float beta1, float beta2, float eps,
                    float momentum, int clip_proper, float clip, uint64_t seed, int dropout, float* loss_out,
                    int32_t* steps_out, uint32_t* bar, float* gnorm, int chain, void* stream);
/* Tensor-core trainer of the ICU TransformerModel (client.py:75-111, src/Model.py:184-246): one cluster of CTAs per client,
 * the whole batch (<= 128 samples) is one tcgen05 M tile, every contraction a tf32 MMA with fp32 TMEM accumulators.  `lay`
 * is the TcLay word list built by attackfl_b200/ops/train_tc.py:build_layout; `ranges` [2 * n_ranges] the merged parameter
 * ranges the optimiser walks; gw [n_clients][d_pad] and stash [n_clients][afl_tc_stash_floats_per_client()] are scratch
 * (stash zero-initialised once by the caller).  loss_out [n_clients][2] = (sum of batch losses, NaN count).  The CTAs of a
 * cluster beyond the first share the optimiser pass; afl_tc_last_cluster() reports the size the last call chose
 * (ATTACKFL_TC_CLUSTER=1|2|4|8 forces one). */
int afl_tc_train_round(const int32_t* lay, int lay_words, float* w, float* m, float* v, float* gw, float* stash,
                       int64_t d_pad, const float* data, int rec, const int32_t* idx, int64_t idx_stride_c,
                       int64_t idx_stride_e, const int32_t* n_data, const int32_t* client_ids, int n_clients, int batch,
                       int epochs, int opt_kind, float lr, float beta1, float beta2, float eps, float momentum,
                       int clip_proper, float clip, uint64_t seed, int dropout, float p1, float p3, const int32_t* ranges,
                       int n_ranges, float* loss_out, int32_t* steps_out, void* stream);
int afl_tc_stash_floats_per_client(void);
int afl_tc_last_cluster(void);
void afl_tc_set_debug(long long* dev_buf);
/* Batched inference of n_models rows over the same records (src/Validation.py:92-122,178-213); probs [n_models][n_rows]. */
int afl_infer(const int32_t* ops_dev, int n_ops, int pool_floats, const float* w, int64_t d_pad, int n_models,
              const float* data, int rec, int64_t n_rows, float* probs, int S, void* stream);

/* Test hooks (host-side in both builds): one forward+backward of a single tile on host memory through the
 * same op implementations, and the dropout keep decision for (seed, client, step, sample, rng, elem, p). */
int afl_debug_tile(const int32_t* ops, int n_ops, int pool_floats, const float* w, float* gw, const float* data,
                   int rec, const int32_t* idx, int S, int n_valid, float inv_b, uint64_t seed, uint32_t client,
                   uint32_t step, int sample_base, int train, float* pool_out, float* loss_out2);
int afl_dropout_keep(uint64_t seed, uint32_t client, uint32_t step, uint32_t sample, uint32_t rng, uint32_t elem,
                     float p);

#ifdef __cplusplus
}
#endif
#endif /* ATTACKFL_B200_H_ */
