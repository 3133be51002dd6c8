// Parameter blocks and host-side launchers of the kernel families (internal).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rq {

struct FwdParams {
    const float* x;
    const float* cb;
    int64_t B;
    int K, L, mode, dist;
    float beta, temp;
    const float* gu;
    uint64_t seed, offset;
    float* emb;
    float* res;
    int64_t* ids;
    float* loss;
    int KC;  // codes per shared-memory chunk (multiple of 4)
    // k-means use of the forward (L = 1, ids = assignment): count rows whose id differs from the value already
    // stored at ids[] (all rows when !has_prev).  nullptr = off.
    unsigned long long* n_changed = nullptr;
    int has_prev = 0;
    // k-means accumulation fused into the assignment (d = 32 tcgen05 kernel, L = 1, ids only): per-cluster sums [K, D]
    // and counts [K] of the rows the launch serves (whole 128-row tiles), added into caller-zeroed buffers.  nullptr = off.
    float* km_sums = nullptr;
    unsigned long long* km_counts = nullptr;
    // optional [B]: exact fp32 distance (SIMT arithmetic) of the id chosen at the LAST level — lets a caller merge
    // the winners of several codebook chunks (rq_quantize_bigk.cu).  nullptr = off.  tcgen05 d=64 family only.
    float* best_dist = nullptr;
    // > 1: the launch covers `n_chunks` consecutive 256-code chunks of ONE codebook of `K` codes (L must be 1):
    // CTA b serves chunk b % n_chunks, ids / best_dist of chunk c are written at offset c * B.
    int n_chunks = 1;
    // first row served by the launch (SIMT family only): the tcgen05 d=32 kernel handles whole 128-row tiles and
    // hands the last B % 128 rows to the exact family; level strides stay B
    int64_t row0 = 0;
};

struct BwdParams {
    const float* g_emb;
    const float* g_res;
    const float* g_loss;
    const float* res;
    const int64_t* ids;
    const float* cb;
    int64_t B;
    int K, L, mode;
    float beta;
    float* g_x;
    float* g_cb;
    int accumulate = 0;      // 1: add into g_cb as it is (no zeroing)
};


struct KmeansParams {
    const float* x;
    const float* centroids;
    int64_t B;
    int K;
    int64_t* assign;
    int has_prev;
    float* sums;
    unsigned long long* counts;
    unsigned long long* n_changed;
    int KC;
};

int quantize_forward_simt(const FwdParams& p, int D, cudaStream_t st);
int quantize_backward_simt(const BwdParams& p, int D, cudaStream_t st);
int quantize_backward_gumbel(const BwdParams& p, int D, float temperature, const float* gu, uint64_t seed,
                             uint64_t offset, cudaStream_t st);
int kmeans_assign_accumulate_simt(const KmeansParams& p, int D, cudaStream_t st);
int kmeans_assign_accumulate_tc(const KmeansParams& p, int D, cudaStream_t st);
int kmeans_update(const float* sums, const int64_t* counts, float* centroids, int D, int K, float* max_shift,
                  int64_t* n_empty, int64_t* empty_ids, cudaStream_t st);

// tcgen05 family: D = 32 in rq_quantize_tc.cu (bf16 x3 split), D = 64 in rq_quantize_tc64.cu (scaled fp16).
// tc_supported / quantize_forward_tc route by D.
bool tc_supported(int D, int K, int L, int mode, int distance);
int quantize_forward_tc(const FwdParams& p, int D, cudaStream_t st);
bool tc32_supported(int D, int K, int L, int mode, int distance);
int quantize_forward_tc32(const FwdParams& p, cudaStream_t st);
bool tc64_supported(int D, int K, int L, int mode, int distance);
int quantize_forward_tc64(const FwdParams& p, cudaStream_t st);
int tc64_timeout_flag();
int tc64_bigk_timeout_flag();
// K > 256 at D = 64: level-major composition of the d=64 kernel over 256-code chunks (rq_quantize_bigk.cu)
bool tc64_bigk_supported(int D, int K, int L, int mode, int distance);
// Gumbel-softmax on the tensor cores (rq_gumbel_tc.cu)
bool gumbel_tc_supported(int D, int K, int L, int distance);
int quantize_forward_gumbel_tc(const FwdParams& p, int D, cudaStream_t st, float* weights_out = nullptr);
int quantize_backward_gumbel_tc(const BwdParams& p, int D, float temperature, const float* gu, uint64_t seed,
                                uint64_t offset, cudaStream_t st, const float* weights = nullptr);
int quantize_forward_tc64_bigk(const FwdParams& p, cudaStream_t st);
void* engine_pool_alloc(size_t bytes, cudaStream_t st);     // library-owned stream-ordered pool (rq_capi.cu)

// rq_linear_tf32.cu: one encoder layer, Y = act(X W^T) in TF32 with the activation fused (act: RQ_ACT_*)
bool linear_tf32_supported(int64_t M, int N, int Kd, int act);
int linear_act_tf32(const float* X, const float* W, float* Y, int64_t M, int N, int Kd, int act, cudaStream_t st);
int linear_tf32_timeout_flag();

// rq_linear_tf32.cu: cuTensorMapEncodeTiled for a row-major fp32 matrix (boxes of box_rows x 32 floats, 128-byte swizzle);
// `map` is a CUtensorMap*
bool make_tensor_map_f32(void* map, const float* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);

}  // namespace rq
