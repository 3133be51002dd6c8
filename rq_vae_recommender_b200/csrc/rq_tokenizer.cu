// Integer paths of the semantic-ID tokenizer (modules/tokenizer/semids.py of the reference).
//
//  * corpus table  — precompute_corpus_ids' dedup column (:86-99).  The reference compares every
//    batch of 512 against everything cached so far (O(N^2 L) booleans).  Here the L-tuple is
//    packed into one 64-bit key, (key, item) pairs are stably radix-sorted, and the dedup value
//    is the item's rank inside its run of equal keys: O(N) work, HBM-bound.
//  * prefix index  — exists_prefix (:103-120): sorted packed prefixes + binary search instead
//    of a [16, n, N, p] broadcast compare.
//  * token gather  — _tokenize_seq_batch_from_cached + padding fill (:122-123, :137-141).
//  * id diversity  — the eval-time metrics train_rqvae.py:227-237 takes from three torch.unique calls
//    (tuple entropy, per-level codebook usage, max duplicates): one pass over the finished table.
//
// The radix sort is cub::DeviceRadixSort (ships with the CUDA toolkit); everything around it
// is hand-written.
#include <cub/cub.cuh>

#include <mutex>

#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

__global__ void pack_keys_kernel(const int64_t* __restrict__ ids, int64_t stride_n, int64_t stride_l, int64_t N,
                                 int L, uint64_t radix, uint64_t* __restrict__ keys, uint32_t* __restrict__ idx) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    uint64_t k = 0;
    for (int l = 0; l < L; ++l) k = k * radix + (uint64_t)ids[i * stride_n + l * stride_l];
    keys[i] = k;
    if (idx) idx[i] = (uint32_t)i;
}

// start[i] = i if key[i] starts a run else 0; an inclusive max-scan turns it into the run start.
__global__ void run_heads_kernel(const uint64_t* __restrict__ keys, int64_t N, uint32_t* __restrict__ start) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    start[i] = (i > 0 && keys[i] != keys[i - 1]) ? (uint32_t)i : 0u;
}

__global__ void write_table_kernel(const int64_t* __restrict__ ids, int64_t stride_n, int64_t stride_l, int64_t N,
                                   int L, const uint32_t* __restrict__ sorted_idx,
                                   const uint32_t* __restrict__ run_start, int64_t* __restrict__ table) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    // position i of the sorted order holds item `it`; stable sort => rank in run = #earlier equal items
    uint32_t it = sorted_idx[i];
    int64_t* row = table + (int64_t)it * (L + 1);
    for (int l = 0; l < L; ++l) row[l] = ids[(int64_t)it * stride_n + l * stride_l];
    row[L] = (int64_t)(i - run_start[i]);
}

struct CorpusWs {
    size_t keys_a, keys_b, idx_a, idx_b, start, cub, total, cub_bytes;
};

static int key_bits(int64_t radix, int L) {
    double b = log2((double)radix) * L;
    int bits = (int)ceil(b - 1e-9);
    return bits < 1 ? 1 : bits;
}

static CorpusWs corpus_layout(int64_t N) {
    CorpusWs w;
    size_t off = 0;
    w.keys_a = off; off += align256(N * 8);
    w.keys_b = off; off += align256(N * 8);
    w.idx_a = off; off += align256(N * 4);
    w.idx_b = off; off += align256(N * 4);
    w.start = off; off += align256(N * 4);
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                    (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)N, 0, 64);
    cub::DeviceScan::InclusiveScan(nullptr, scan_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, cub::Max(),
                                   (int64_t)N);
    w.cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    w.cub = off; off += align256(w.cub_bytes);
    w.total = off;
    return w;
}

}  // namespace rq

using namespace rq;

extern "C" int64_t rq_corpus_table_workspace_bytes(int64_t N, int L) {
    (void)L;
    if (N <= 0) return 256;
    return (int64_t)corpus_layout(N).total;
}

extern "C" int rq_corpus_table(const int64_t* sem_ids, int64_t stride_n, int64_t stride_l, int64_t N, int L,
                               int64_t codebook_size, int64_t* table, void* workspace, int64_t workspace_bytes,
                               rq_stream_t stream) {
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(N >= 0 && L >= 1 && codebook_size >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_corpus_table: bad N/L/codebook_size");
    if (N == 0) return 0;
    RQ_CHECK_ARG(sem_ids && table && workspace, RQ_ERR_NULL_POINTER, "rq_corpus_table: null pointer");
    RQ_CHECK_ARG(N < (int64_t)4294967295LL, RQ_ERR_UNSUPPORTED_SHAPE, "rq_corpus_table: N must be < 2^32");
    RQ_CHECK_ARG(log2((double)codebook_size) * L <= 63.0, RQ_ERR_UNSUPPORTED_SHAPE,
                 "rq_corpus_table: codebook_size^L does not fit 63 bits");
    CorpusWs w = corpus_layout(N);
    RQ_CHECK_ARG((int64_t)w.total <= workspace_bytes, RQ_ERR_WORKSPACE_TOO_SMALL,
                 "rq_corpus_table: workspace %lld < %lld bytes", (long long)workspace_bytes, (long long)w.total);
    char* base = (char*)workspace;
    uint64_t* keys_a = (uint64_t*)(base + w.keys_a);
    uint64_t* keys_b = (uint64_t*)(base + w.keys_b);
    uint32_t* idx_a = (uint32_t*)(base + w.idx_a);
    uint32_t* idx_b = (uint32_t*)(base + w.idx_b);
    uint32_t* start = (uint32_t*)(base + w.start);
    const int T = 256;
    const unsigned G = (unsigned)((N + T - 1) / T);
    pack_keys_kernel<<<G, T, 0, st>>>(sem_ids, stride_n, stride_l, N, L, (uint64_t)codebook_size, keys_a, idx_a);
    RQ_LAUNCHED();
    size_t cb = w.cub_bytes;
    RQ_CUDA(cub::DeviceRadixSort::SortPairs(base + w.cub, cb, keys_a, keys_b, idx_a, idx_b, (int64_t)N, 0,
                                            key_bits(codebook_size, L), st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    run_heads_kernel<<<G, T, 0, st>>>(keys_b, N, start);
    RQ_LAUNCHED();
    cb = w.cub_bytes;
    RQ_CUDA(cub::DeviceScan::InclusiveScan(base + w.cub, cb, start, start, cub::Max(), (int64_t)N, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    write_table_kernel<<<G, T, 0, st>>>(sem_ids, stride_n, stride_l, N, L, idx_b, start, table);
    RQ_LAUNCHED();
    return 0;
}

// ---- prefix index ----------------------------------------------------------------------------
extern "C" int64_t rq_prefix_index_workspace_bytes(int64_t N) {
    if (N <= 0) return 256;
    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int64_t)N, 0, 64);
    return (int64_t)(align256(N * 8) + align256(sort_bytes));
}

extern "C" int rq_prefix_index_build(const int64_t* table, int64_t N, int row_stride, int p, int64_t radix,
                                     uint64_t* sorted_keys, void* workspace, int64_t workspace_bytes,
                                     rq_stream_t stream) {
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(N >= 0 && p >= 1 && p <= row_stride && radix >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_prefix_index_build: bad arguments");
    if (N == 0) return 0;
    RQ_CHECK_ARG(table && sorted_keys && workspace, RQ_ERR_NULL_POINTER, "rq_prefix_index_build: null pointer");
    RQ_CHECK_ARG(log2((double)radix) * p <= 63.0, RQ_ERR_UNSUPPORTED_SHAPE, "rq_prefix_index_build: radix^p does not fit 63 bits");
    RQ_CHECK_ARG(rq_prefix_index_workspace_bytes(N) <= workspace_bytes, RQ_ERR_WORKSPACE_TOO_SMALL,
                 "rq_prefix_index_build: workspace too small");
    char* base = (char*)workspace;
    uint64_t* keys = (uint64_t*)base;
    size_t sort_bytes = (size_t)workspace_bytes - align256(N * 8);
    const int T = 256;
    pack_keys_kernel<<<(unsigned)((N + T - 1) / T), T, 0, st>>>(table, row_stride, 1, N, p, (uint64_t)radix, keys, nullptr);
    RQ_LAUNCHED();
    RQ_CUDA(cub::DeviceRadixSort::SortKeys(base + align256(N * 8), sort_bytes, keys, sorted_keys, (int64_t)N, 0,
                                           key_bits(radix, p), st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

__global__ void prefix_exists_kernel(const uint64_t* __restrict__ sorted, int64_t N, const int64_t* __restrict__ q,
                                     int64_t Q, int p, uint64_t radix, uint8_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Q) return;
    uint64_t key = 0;
    bool in_range = true;
    for (int l = 0; l < p; ++l) {
        int64_t v = q[i * p + l];
        in_range &= (v >= 0 && (uint64_t)v < radix);
        key = key * radix + (uint64_t)v;
    }
    bool hit = false;
    if (in_range) {
        int64_t lo = 0, hi = N;  // lower bound
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (sorted[mid] < key) lo = mid + 1; else hi = mid;
        }
        hit = lo < N && sorted[lo] == key;
    }
    out[i] = hit ? 1 : 0;
}

extern "C" int rq_prefix_exists(const uint64_t* sorted_keys, int64_t N, const int64_t* queries, int64_t Q, int p,
                                int64_t radix, uint8_t* out, rq_stream_t stream) {
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(Q >= 0 && N >= 0 && p >= 1 && radix >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_prefix_exists: bad arguments");
    if (Q == 0) return 0;
    RQ_CHECK_ARG(queries && out && (sorted_keys || N == 0), RQ_ERR_NULL_POINTER, "rq_prefix_exists: null pointer");
    const int T = 256;
    prefix_exists_kernel<<<(unsigned)((Q + T - 1) / T), T, 0, st>>>(sorted_keys, N, queries, Q, p, (uint64_t)radix, out);
    RQ_LAUNCHED();
    return 0;
}

// ---- token gather ----------------------------------------------------------------------------
__global__ void gather_tokens_kernel(const int64_t* __restrict__ table, int64_t n_rows, int W,
                                     const int64_t* __restrict__ item_ids, const uint8_t* __restrict__ mask,
                                     int64_t n_tokens, int64_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tokens * W) return;
    int64_t t = i / W;
    int w = (int)(i % W);
    bool keep = mask ? mask[t] != 0 : true;
    // The reference gathers first and overwrites padding with -1 afterwards (semids.py:139-141);
    // negative ids wrap like torch indexing does (ids_fut padding is -1).
    int64_t id = item_ids[t];
    if (id < 0) id += n_rows;
    // ids outside [-n_rows, n_rows) raise IndexError in the reference; the Python mirror checks the range before the
    // launch (SemanticIdTokenizer.forward).  The kernel itself never reads out of bounds: such a token becomes -1.
    const bool in_range = id >= 0 && id < n_rows;
    out[i] = (keep && in_range) ? table[id * W + w] : -1;
}

extern "C" int rq_gather_tokens(const int64_t* table, int64_t n_rows, int W, const int64_t* item_ids,
                                const uint8_t* seq_mask, int64_t n_tokens, int64_t* out, rq_stream_t stream) {
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(n_tokens >= 0 && W >= 1 && n_rows >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_gather_tokens: bad arguments");
    if (n_tokens == 0) return 0;
    RQ_CHECK_ARG(table && item_ids && out, RQ_ERR_NULL_POINTER, "rq_gather_tokens: null pointer");
    const int T = 256;
    int64_t total = n_tokens * W;
    gather_tokens_kernel<<<(unsigned)((total + T - 1) / T), T, 0, st>>>(table, n_rows, W, item_ids, seq_mask, n_tokens, out);
    RQ_LAUNCHED();
    return 0;
}

// ---- id diversity ----------------------------------------------------------------------------
// entropy = -sum_t p_t log p_t over the distinct L-tuples t (p_t = n_t / N) needs no sort once the dedup column
// exists: the items of a tuple carry dedup = 0 .. n_t - 1, and with f(n) = (n/N) log(n/N), f(0) = 0,
//   f(n_t) = sum_{d=0}^{n_t-1} [f(d+1) - f(d)]   =>   entropy = -sum_items [f(dedup_i + 1) - f(dedup_i)].
// usage: one bit per (level, id) in a CTA-private shared-memory bitmap, OR-ed into the global bitmap once per CTA.
__global__ void id_diversity_kernel(const int64_t* __restrict__ table, int64_t N, int L, int64_t K, int words,
                                    bool use_smem, double* __restrict__ entropy, unsigned long long* __restrict__ max_dup,
                                    uint32_t* __restrict__ bitmap) {
    extern __shared__ uint32_t s_bits[];
    if (use_smem) {
        for (int i = threadIdx.x; i < L * words; i += blockDim.x) s_bits[i] = 0u;
        __syncthreads();
    }
    const double invN = 1.0 / (double)N;
    double acc = 0.0;
    unsigned long long mx = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t* row = table + i * (L + 1);
        for (int l = 0; l < L; ++l) {
            const int64_t v = row[l];
            if (v >= 0 && v < K) {
                const uint32_t bit = 1u << (v & 31);
                uint32_t* w = (use_smem ? s_bits : bitmap) + l * words + (int)(v >> 5);
                if (!(*w & bit)) atomicOr(w, bit);
            }
        }
        const int64_t d = row[L];
        const double a = (double)(d + 1) * invN, b = (double)d * invN;
        acc += a * log(a) - (d > 0 ? b * log(b) : 0.0);
        mx = (unsigned long long)d > mx ? (unsigned long long)d : mx;
    }
    // block reduction (double sum, max)
    __shared__ double s_acc[32];
    __shared__ unsigned long long s_mx[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, mx, o);
        mx = t > mx ? t : mx;
    }
    if ((threadIdx.x & 31) == 0) { s_acc[threadIdx.x >> 5] = acc; s_mx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        unsigned long long m = 0;
        for (int w = 0; w < (int)(blockDim.x + 31) / 32; ++w) { a += s_acc[w]; m = s_mx[w] > m ? s_mx[w] : m; }
        atomicAdd(entropy, -a);
        atomicMax(max_dup, m);
    }
    if (use_smem) {
        for (int i = threadIdx.x; i < L * words; i += blockDim.x)
            if (s_bits[i]) atomicOr(bitmap + i, s_bits[i]);
    }
}

__global__ void id_usage_kernel(const uint32_t* __restrict__ bitmap, int L, int words, int64_t* __restrict__ used) {
    const int l = blockIdx.x;
    int c = 0;
    for (int i = threadIdx.x; i < words; i += blockDim.x) c += __popc(bitmap[l * words + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    __shared__ int s_c[32];
    if ((threadIdx.x & 31) == 0) s_c[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < (int)(blockDim.x + 31) / 32; ++w) t += s_c[w];
        used[l] = t;
    }
}


extern "C" int64_t rq_id_diversity_workspace_bytes(int L, int64_t codebook_size) {
    if (L < 1 || codebook_size < 1) return 0;
    return (int64_t)L * ((codebook_size + 31) / 32) * 4;
}

extern "C" int rq_id_diversity(const int64_t* table, int64_t N, int L, int64_t codebook_size, double* entropy,
                               int64_t* used, int64_t* max_dup, void* workspace, int64_t workspace_bytes,
                               rq_stream_t stream) {
    using namespace rq;
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(N >= 1 && L >= 1 && codebook_size >= 1, RQ_ERR_INVALID_ARGUMENT,
                 "rq_id_diversity: need N >= 1, L >= 1, codebook_size >= 1");
    RQ_CHECK_ARG(table && entropy && used && max_dup && workspace, RQ_ERR_NULL_POINTER, "rq_id_diversity: null pointer");
    const int64_t need = rq_id_diversity_workspace_bytes(L, codebook_size);
    RQ_CHECK_ARG(workspace_bytes >= need, RQ_ERR_INVALID_ARGUMENT, "rq_id_diversity: workspace too small (%lld < %lld)",
                 (long long)workspace_bytes, (long long)need);
    const int words = (int)((codebook_size + 31) / 32);
    RQ_CUDA(cudaMemsetAsync(workspace, 0, (size_t)need, st));
    RQ_CUDA(cudaMemsetAsync(entropy, 0, sizeof(double), st));
    RQ_CUDA(cudaMemsetAsync(max_dup, 0, sizeof(int64_t), st));
    const bool use_smem = need <= 32 * 1024;
    const int T = 256;
    int64_t blocks = (N + T - 1) / T;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    id_diversity_kernel<<<(unsigned)blocks, T, use_smem ? (size_t)need : 0, st>>>(
        table, N, L, codebook_size, words, use_smem, entropy, reinterpret_cast<unsigned long long*>(max_dup),
        reinterpret_cast<uint32_t*>(workspace));
    RQ_LAUNCHED();
    id_usage_kernel<<<L, 256, 0, st>>>(reinterpret_cast<const uint32_t*>(workspace), L, words, used);
    RQ_LAUNCHED();
    return 0;
}

// ---- id diversity without any allocation -----------------------------------------------------
// Same two kernels, with the (level, id) bitmap in a file-scope device array instead of a caller workspace, for
// callers that may not allocate (graph capture, serving loops).  The array is part of the module image: every GPU
// that loads the library holds kIdDiversityStaticBytes of it until exit.  In host code `g_id_bitmap_static` is only
// a stand-in for the symbol; the device address comes from cudaGetSymbolAddress, looked up once per device with
// that device current.  All calls on a device share the array: it is cleared at the start of every call, and two
// calls that overlap on the GPU (different streams) race on it.
namespace rq {
constexpr int64_t kIdDiversityStaticBytes = 64 * 1024;
static __device__ __align__(256) unsigned char g_id_bitmap_static[kIdDiversityStaticBytes];

static cudaError_t id_bitmap_static_ptr(uint32_t** out) {
    static std::mutex mu;
    static uint32_t* per_dev[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!per_dev[dev]) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_id_bitmap_static);
        if (e != cudaSuccess) return e;
        per_dev[dev] = static_cast<uint32_t*>(ptr);
    }
    *out = per_dev[dev];
    return cudaSuccess;
}
}  // namespace rq

extern "C" int64_t rq_id_diversity_static_max_bytes(void) { return rq::kIdDiversityStaticBytes; }

extern "C" int rq_id_diversity_static(const int64_t* table, int64_t N, int L, int64_t codebook_size, double* entropy,
                                      int64_t* used, int64_t* max_dup, rq_stream_t stream) {
    using namespace rq;
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(N >= 1 && L >= 1 && codebook_size >= 1, RQ_ERR_INVALID_ARGUMENT,
                 "rq_id_diversity_static: need N >= 1, L >= 1, codebook_size >= 1");
    RQ_CHECK_ARG(table && entropy && used && max_dup, RQ_ERR_NULL_POINTER, "rq_id_diversity_static: null pointer");
    const int64_t words64 = (codebook_size + 31) / 32;
    RQ_CHECK_ARG(words64 <= kIdDiversityStaticBytes / 4 && (int64_t)L * words64 * 4 <= kIdDiversityStaticBytes,
                 RQ_ERR_WORKSPACE_TOO_SMALL,
                 "rq_id_diversity_static: L * ceil(codebook_size / 32) * 4 = %lld bytes exceeds the built-in %lld-byte "
                 "bitmap; use rq_id_diversity with a caller workspace",
                 (long long)L * (long long)words64 * 4, (long long)kIdDiversityStaticBytes);
    const int64_t need = (int64_t)L * words64 * 4;
    const int words = (int)words64;
    uint32_t* bitmap = nullptr;
    RQ_CUDA(id_bitmap_static_ptr(&bitmap));
    RQ_CUDA(cudaMemsetAsync(bitmap, 0, (size_t)need, st));
    RQ_CUDA(cudaMemsetAsync(entropy, 0, sizeof(double), st));
    RQ_CUDA(cudaMemsetAsync(max_dup, 0, sizeof(int64_t), st));
    const bool use_smem = need <= 32 * 1024;
    const int T = 256;
    int64_t blocks = (N + T - 1) / T;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    id_diversity_kernel<<<(unsigned)blocks, T, use_smem ? (size_t)need : 0, st>>>(
        table, N, L, codebook_size, words, use_smem, entropy, reinterpret_cast<unsigned long long*>(max_dup), bitmap);
    RQ_LAUNCHED();
    id_usage_kernel<<<L, 256, 0, st>>>(bitmap, L, words, used);
    RQ_LAUNCHED();
    return 0;
}
