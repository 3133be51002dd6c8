/*
 * rq_engine.h — C ABI of the B200-native residual-quantization engine.
 *
 * Drop-in boundary for the quantization hot path of AdamLTy/RQ-VAE-Recommender.  The
 * reference is pure Python with no FFI of its own; these entry points
 * are what a binding for that path attaches to.  Each one names the reference code it
 * replaces (paths relative to the reference root).  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless the function name ends in `_host`.
 *   - Row-major, contiguous.  fp32 values, int64 indices (the reference's dtypes).
 *   - `stream` is a cudaStream_t (0 = legacy default stream).  Nothing here synchronises
 *     the host except the `_host` functions, and nothing allocates except those and
 *     functions that take an explicit workspace.  All launches are CUDA-graph capturable.
 *   - Return value: 0 on success; RQ_ERR_* (<0) for argument errors; a positive
 *     cudaError_t for CUDA failures.  rq_last_error() returns a thread-local message.
 *   - There is no CPU implementation behind any of these.
 */
#ifndef RQ_ENGINE_H_
#define RQ_ENGINE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

typedef void* rq_stream_t; /* cudaStream_t */

#define RQ_ABI_VERSION 1

/* modules/quantize.py:16-20 (QuantizeForwardMode); 0 = the `not self.training` branch :148-150 */
#define RQ_MODE_EVAL 0
#define RQ_MODE_GUMBEL_SOFTMAX 1
#define RQ_MODE_STE 2
#define RQ_MODE_ROTATION_TRICK 3
/* modules/quantize.py:23-25 (QuantizeDistance) */
#define RQ_DIST_L2 1
#define RQ_DIST_COSINE 2
/* kernel family selection */
#define RQ_IMPL_AUTO 0    /* tcgen05 kernel when the shape is supported, else exact SIMT */
#define RQ_IMPL_SIMT 1    /* fp32 CUDA-core kernel (exact expanded form), any shape */
#define RQ_IMPL_TCGEN05 2 /* tensor-core candidate search + fp32 rescoring; error if unsupported */

#define RQ_ERR_INVALID_ARGUMENT (-1)
#define RQ_ERR_UNSUPPORTED_SHAPE (-2)
#define RQ_ERR_NULL_POINTER (-3)
#define RQ_ERR_WORKSPACE_TOO_SMALL (-4)
#define RQ_ERR_NO_DEVICE (-5)

int rq_abi_version(void);
const char* rq_last_error(void);
/* Number of kernel launches issued through this library by the calling process
 * (bench.py reports it as `gpu_launches`). */
uint64_t rq_launch_count(void);
/* Returns 1 if (D, K, L, mode, distance) is served by the tcgen05 kernel. */
int rq_tcgen05_supported(int D, int K, int L, int mode, int distance);

/*
 * Fused L-level residual quantisation, forward.
 * Replaces the level loop of RqVae.get_semantic_ids (modules/rqvae.py:121-131) together with
 * Quantize.forward (modules/quantize.py:105-150), QuantizeLoss (modules/loss.py:39-42),
 * efficient_rotation_trick_transform (modules/quantize.py:34-45) and gumbel_softmax_sample
 * (distributions/gumbel.py:7-17).
 *
 *   x           [B, D]     latents (output of RqVae.encode)
 *   codebooks   [L, K, D]  EFFECTIVE codebooks, i.e. out_proj(embedding.weight) (quantize.py:105)
 *   gumbel_u    [L, B, K]  the uniform draws of sample_gumbel (gumbel.py:9) or NULL; when NULL
 *                          and mode == GUMBEL the kernel draws Philox4x32-10 noise from
 *                          (seed, offset).  Ignored by the other modes.
 *   emb_out     [L, B, D]  `quantized.embeddings` per level, or NULL (ids-only)
 *   residuals   [L, B, D]  input of each level (rqvae.py:125), or NULL
 *   ids         [L, B]     argmin indices, first index wins ties (quantize.py:121); required
 *   loss        [B]        sum over levels of emb_loss + beta*query_loss, or NULL
 */
int rq_quantize_forward(const float* x, const float* codebooks, int64_t B, int D, int K, int L,
                        int mode, int distance, float beta, float temperature,
                        const float* gumbel_u, uint64_t seed, uint64_t offset,
                        float* emb_out, float* residuals, int64_t* ids, float* loss,
                        int impl, rq_stream_t stream);

/*
 * The same with one more output for the Gumbel-softmax pair on the tensor cores: gumbel_weights [L, B, Kp], Kp = K
 * rounded up to a multiple of 4 — the softmax weights W of every level (what autograd keeps alive in the reference:
 * the output of F.softmax, quantize.py:127).  Handing the buffer back to rq_quantize_backward_ws saves the backward the
 * recomputation of the logits product and of the softmax (a fifth of the step at 4 x 1024 codes).  NULL, another mode,
 * or a launch that the thread-per-row kernels serve (see rq_gumbel_weights_saved) = rq_quantize_forward.
 */
int rq_quantize_forward_ws(const float* x, const float* codebooks, int64_t B, int D, int K, int L,
                           int mode, int distance, float beta, float temperature,
                           const float* gumbel_u, uint64_t seed, uint64_t offset,
                           float* emb_out, float* residuals, int64_t* ids, float* loss,
                           float* gumbel_weights, int impl, rq_stream_t stream);
/* 1 if rq_quantize_forward_ws with these arguments fills gumbel_weights (the tensor-core Gumbel path is taken). */
int rq_gumbel_weights_saved(int64_t B, int D, int K, int L, int mode, int distance, int impl);

/*
 * Fused backward of the above (the reference relies on autograd).
 *   g_emb [L,B,D], g_res [L,B,D], g_loss [B]  upstream gradients (any may be NULL = zero)
 *   residuals, ids                             saved from the forward
 *   g_x [B, D]          gradient w.r.t. the latents (written)
 *   g_codebooks [L,K,D] gradient w.r.t. the effective codebooks (overwritten, not accumulated)
 * distance must be RQ_DIST_L2 for GUMBEL_SOFTMAX (the only mode whose gradient sees `dist`).
 */
int rq_quantize_backward(const float* g_emb, const float* g_res, const float* g_loss,
                         const float* residuals, const int64_t* ids, const float* codebooks,
                         int64_t B, int D, int K, int L, int mode, float beta, float temperature,
                         const float* gumbel_u, uint64_t seed, uint64_t offset,
                         float* g_x, float* g_codebooks, rq_stream_t stream);

/* The same with flags.  RQ_BWD_ACCUMULATE: the codebook gradient is ADDED to what g_codebooks holds (gradient
 * accumulation; or zeros the caller wrote before the forward, which keeps the forward and this kernel adjacent in
 * the stream so that programmatic dependent launch can overlap this kernel's launch with the forward's tail).
 * Honoured by the EVAL / STE / ROTATION_TRICK modes; GUMBEL_SOFTMAX overwrites. */
#define RQ_BWD_ACCUMULATE 1
#define RQ_BWD_THREAD_PER_ROW 2   /* Gumbel-softmax: keep the CUDA-core kernel whatever the batch size (A/B, tests) */
int rq_quantize_backward_ex(const float* g_emb, const float* g_res, const float* g_loss,
                            const float* residuals, const int64_t* ids, const float* codebooks,
                            int64_t B, int D, int K, int L, int mode, float beta, float temperature,
                            const float* gumbel_u, uint64_t seed, uint64_t offset,
                            float* g_x, float* g_codebooks, int flags, rq_stream_t stream);

/*
 * One k-means assignment + accumulation pass.  Replaces init/kmeans.py:41-57 (the [B,K,D]
 * broadcast, argmin, and the per-cluster masked means' numerators).
 *   x [B,D], centroids [K,D]
 *   assign [B] int64     in: previous assignment when has_prev != 0; out: new assignment
 *   sums [K,D] fp32, counts [K] int64   (overwritten)
 *   n_changed [1] int64  number of rows whose assignment changed (B when !has_prev)
 */
int rq_kmeans_assign_accumulate(const float* x, const float* centroids, int64_t B, int D, int K,
                                int64_t* assign, int has_prev, float* sums, int64_t* counts,
                                int64_t* n_changed, int impl, rq_stream_t stream);

/*
 * Centroid update from (all-reduced) sums/counts: c_k = sums_k / counts_k for non-empty
 * clusters (kmeans.py:57); empty clusters keep their old value and are reported so that the
 * host re-seeds them (kmeans.py:51-53).
 *   max_shift [1] fp32   max_k ||c_new - c_old||_2   (kmeans.py:67)
 *   n_empty [1] int64, empty_ids [K] int64 (first n_empty entries valid, ascending)
 */
int rq_kmeans_update(const float* sums, const int64_t* counts, float* centroids, int D, int K,
                     float* max_shift, int64_t* n_empty, int64_t* empty_ids, rq_stream_t stream);

/*
 * Dedup column of the corpus id table.  Replaces the O(N^2) hit matrices of
 * SemanticIdTokenizer.precompute_corpus_ids (modules/tokenizer/semids.py:86-99):
 * dedup[i] = number of j < i with sem_ids[j,:] == sem_ids[i,:].
 *   sem_ids [N, L] int64 with arbitrary element strides (the [L,B]-memory view is accepted)
 *   table   [N, L+1] int64 contiguous: [sem_ids | dedup]
 *   base_counts/out_counts: optional (may be NULL) — multi-rank use, see DESIGN.md.
 */
int64_t rq_corpus_table_workspace_bytes(int64_t N, int L);
int rq_corpus_table(const int64_t* sem_ids, int64_t stride_n, int64_t stride_l, int64_t N, int L,
                    int64_t codebook_size, int64_t* table, void* workspace, int64_t workspace_bytes,
                    rq_stream_t stream);

/*
 * Prefix index over the cached id table.  Replaces SemanticIdTokenizer.exists_prefix
 * (semids.py:103-120): out[q] = any_i(table[i, :p] == queries[q, :p]).
 *   rq_prefix_index_build: sorted_keys [N] uint64 <- packed table[:, :p] (radix codebook_size+? see .cu)
 *   rq_prefix_exists: binary search, out [Q] uint8 (bool)
 */
int64_t rq_prefix_index_workspace_bytes(int64_t N);
int rq_prefix_index_build(const int64_t* table, int64_t N, int row_stride, int p, int64_t radix,
                          uint64_t* sorted_keys, void* workspace, int64_t workspace_bytes,
                          rq_stream_t stream);
int rq_prefix_exists(const uint64_t* sorted_keys, int64_t N, const int64_t* queries, int64_t Q, int p,
                     int64_t radix, uint8_t* out, rq_stream_t stream);

/*
 * Sequence tokenisation by table gather.  Replaces _tokenize_seq_batch_from_cached and the
 * padding fill (semids.py:122-123, 137-141): out[b, n*W + w] = mask[b,n] ? table[ids[b,n], w] : -1.
 *   table is [n_rows, W]; negative item ids wrap (torch indexing); seq_mask [n_tokens] may be
 *   NULL (no padding fill).
 */
int rq_gather_tokens(const int64_t* table, int64_t n_rows, int W, const int64_t* item_ids,
                     const uint8_t* seq_mask, int64_t n_tokens, int64_t* out, rq_stream_t stream);

/*
 * Id-diversity metrics of a finished corpus table.  Replaces the three torch.unique calls of the reference's
 * eval block (train_rqvae.py:227-237):
 *   entropy  = -sum_t p_t log p_t over the distinct L-tuples t, p_t = count_t / N      (rqvae_entropy)
 *   used[l]  = number of distinct ids in column l            (codebook_usage_l = used[l] / codebook_size)
 *   max_dup  = max of the dedup column                        (max_id_duplicates = max_dup / N)
 * table [N, L+1] int64 contiguous = [ids | dedup] as written by rq_corpus_table; entropy [1] double, used [L]
 * int64, max_dup [1] int64 on the device.  One pass, no sort: the entropy telescopes over the dedup column.
 * workspace: rq_id_diversity_workspace_bytes(L, codebook_size) bytes (a bitmap; zeroed by the call).
 */
int64_t rq_id_diversity_workspace_bytes(int L, int64_t codebook_size);
int rq_id_diversity(const int64_t* table, int64_t N, int L, int64_t codebook_size, double* entropy, int64_t* used,
                    int64_t* max_dup, void* workspace, int64_t workspace_bytes, rq_stream_t stream);

/*
 * rq_id_diversity without a workspace and without any allocation (no cudaMalloc / cudaFree / allocating library
 * call on any path), for callers that cannot allocate: code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it.  No counterpart in the reference.
 * Inputs, outputs and results are those of rq_id_diversity; the (level, id) bitmap lives in a file-scope device
 * array of rq_id_diversity_static_max_bytes() = 65,536 bytes that every GPU loading the library holds until exit.
 * Limit: L * ceil(codebook_size / 32) * 4 <= 65,536 (e.g. L = 4 levels of up to 131,072 ids); beyond it the call
 * returns RQ_ERR_WORKSPACE_TOO_SMALL and launches nothing.  The array is shared by all calls on a device and is
 * cleared at the start of each: calls that overlap on the GPU (different streams, no ordering) race on it.
 */
int64_t rq_id_diversity_static_max_bytes(void);
int rq_id_diversity_static(const int64_t* table, int64_t N, int L, int64_t codebook_size, double* entropy,
                           int64_t* used, int64_t* max_dup, rq_stream_t stream);

/*
 * One-shot all-reduce (sum * scale) of a small fp32 region over peer-mapped symmetric memory (NVLink loads, no
 * NCCL): the codebook-gradient exchange of a training step, which the reference leaves to DDP
 * (train_rqvae.py:115-117,153), and the k-means sums / counts exchange.  Capturable in a CUDA graph; every rank
 * launches it with the same n_floats in the same order.
 *   peer_bufs_dev   device array [world] of the ranks' allocation base pointers (rank order)
 *   peer_pads_dev   device array [world] of the ranks' signal pads (zero-initialised uint32 words, at least
 *                   rq_peer_allreduce_pad_bytes() each; words from 2048 on are used)
 *   offset_floats, n_floats   the reduced region inside every allocation (multiples of 4, n_floats <=
 *                   rq_peer_allreduce_max_floats())
 *   out             local destination, 16-byte aligned; may be this rank's own region (written after every rank
 *                   has finished reading)
 *   epoch_dev       device-local zero-initialised uint32 [rq_peer_allreduce_state_words()], one per allocation
 * Ranks sum in rank order, so all ranks end with bit-identical values.  torch.distributed._symmetric_memory
 * provides the pointer tables (rq_vae_recommender_b200/parallel.py::PeerAllReduce).
 */
int64_t rq_peer_allreduce_max_floats(void);
int rq_peer_allreduce_state_words(void);
int rq_peer_allreduce_pad_bytes(void);
int rq_peer_allreduce_f32(const void* peer_bufs_dev, const void* peer_pads_dev, int rank, int world,
                          int64_t offset_floats, int64_t n_floats, float* out, float scale, uint32_t* epoch_dev,
                          rq_stream_t stream);
int rq_debug_peer_timeout_flag(void);   /* 1 once a rank failed to arrive at a barrier of the kernel above */

/*
 * Host-buffer convenience entry (the call a non-torch integrator makes): copies `x_host`
 * (pageable or pinned) to the device, runs rq_quantize_forward in RQ_MODE_EVAL, copies the ids
 * [L,B] (and loss [B] when non-NULL) back and returns when they are complete.  Rows are processed
 * in chunks (262,144 rows; env RQ_HOST_CHUNK_ROWS) as a three-stage pipeline -- H2D of chunk i+1,
 * compute of chunk i on `stream`, D2H of chunk i-1 -- so device scratch is two chunk-sized buffer
 * pairs from the stream-ordered allocator whatever B is.  `codebooks` is a DEVICE pointer.
 */
int rq_semantic_ids_host(const float* x_host, const float* codebooks, int64_t B, int D, int K, int L,
                         int distance, int64_t* ids_host, float* loss_host, int impl,
                         rq_stream_t stream);

/* rq_quantize_backward_ex with the softmax weights rq_quantize_forward_ws saved (NULL = recompute them). */
int rq_quantize_backward_ws(const float* g_emb, const float* g_res, const float* g_loss,
                            const float* residuals, const int64_t* ids, const float* codebooks,
                            int64_t B, int D, int K, int L, int mode, float beta, float temperature,
                            const float* gumbel_u, uint64_t seed, uint64_t offset,
                            float* g_x, float* g_codebooks, const float* gumbel_weights, int flags, rq_stream_t stream);

/* Test hooks of the tcgen05 family: 1 if any mbarrier wait timed out since load (synchronises the
 * device); and a [128][256] fp32 device buffer that receives tile 0 / level 0's raw accumulator
 * (||c||^2 - 2 x.c) on subsequent launches (NULL disables). */
int rq_debug_tc_timeout_flag(void);
int rq_debug_tc_set_accumulator_dump(float* device_buf_128x256);
/* Profiling hook: a [32 warps][64 marks] int64 device buffer that receives clock64() stamps of CTA 0's
 * phases (prologue end, then per level: operand written, MMAs issued, accumulator ready, scan done,
 * level done) on subsequent launches (NULL disables). */
int rq_debug_tc_set_timeline(long long* device_buf_32x64);
/* Test hook: the fp32-grade tensor-core GEMM of the Gumbel-softmax path on its own (csrc/rq_gemm_split.cu):
 * C[M, N] (+)= alpha * A[M, Kd] * B[N, Kd]^T, fp32 row-major device buffers with row strides lda / ldb / ldc (floats;
 * lda, ldb multiples of 4, A and B 16-byte aligned).  atomic != 0 or k_splits > 1 adds into C with float atomics.
 * Stands where the reference calls torch.matmul inside the Gumbel branch (modules/quantize.py:125-128 and its autograd). */
int rq_debug_split_gemm(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd, int64_t lda,
                        int64_t ldb, int64_t ldc, float alpha, int atomic, int k_splits, rq_stream_t stream);
/* The same product with both operands given transposed, At[Kd, M] and Bt[Kd, N] (row strides ldat / ldbt), N <= 64: the
 * shape of the Gumbel backward's codebook gradients (W^T de, dd^T x over the batch), staged MN-major so that no transposed
 * copy of the [B, K] matrices is written. */
int rq_debug_split_gemm_tn(const float* At, const float* Bt, float* C, int64_t M, int64_t N, int64_t Kd, int64_t ldat,
                           int64_t ldbt, int64_t ldc, float alpha, int atomic, int k_splits, rq_stream_t stream);

/* ---- encoder layer ---------------------------------------------------------------------------
 * y[rows, out] = act(x[rows, in] * weight[out, in]^T): one bias-free Linear of the reference's MLP
 * (modules/encoder.py:17-32, reached from RqVae.encode, modules/rqvae.py:117-118) with its activation fused into the
 * GEMM epilogue.  fp32 row-major contiguous device buffers, 16-byte aligned; in / out multiples of 4.  The product is
 * computed in TF32 on the tensor cores - the precision the reference selects for these matmuls
 * (torch.set_float32_matmul_precision('high'), modules/rqvae.py:19).
 *   RQ_ACT_NONE    plain Linear (the last layer when normalize=False)
 *   RQ_ACT_SILU    Linear + nn.SiLU (encoder.py:22-23)
 *   RQ_ACT_L2NORM  Linear + L2NormalizationLayer (encoder.py:27, modules/normalize.py:9-13: x / max(|x|, 1e-12)); out <= 64
 * OR-ing RQ_ACT_TF32_OUT into act (NONE / SILU only) rounds every output to the nearest value with a 10-bit significand:
 * the next layer's tensor-core product then reads its operand without the truncation it would otherwise apply, which
 * makes a chain of layers unbiased (round-to-nearest) where back-to-back TF32 GEMMs truncate.  Use it for hidden
 * activations that only the next rq_encoder_layer call consumes.
 * rq_encoder_layer_supported returns 1 when rq_encoder_layer accepts the shape. */
#define RQ_ACT_NONE 0
#define RQ_ACT_SILU 1
#define RQ_ACT_L2NORM 2
#define RQ_ACT_TF32_OUT 0x10
/* Run the layer on CTA pairs (clusters of two, tcgen05 cta_group::2, each CTA staging half of the weight tile) when
 * out is a multiple of 256 >= 512 and rows >= 512; ignored otherwise.  Same results; measured no faster than the
 * default single-CTA schedule on B200, kept selectable for other shapes. */
#define RQ_ACT_CTA_PAIR 0x20
int rq_encoder_layer_supported(int64_t rows, int in_features, int out_features, int act);
int rq_encoder_layer(const float* x, const float* weight, float* y, int64_t rows, int in_features, int out_features,
                     int act, rq_stream_t stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* RQ_ENGINE_H_ */
