// extern "C" entry points of librqengine.so (see include/rq_engine.h for the contract).
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <initializer_list>
#include <mutex>
#include <vector>

#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// SM count of the current device minus the SMs reserved for concurrently running collectives (RQ_SM_RESERVE is read
// once per process: NCCL kernels need a few CTAs' worth of SM resources and a persistent grid that fills every SM
// would make one of the two wait for the other).  Cached per device.
int sm_count() {
    static std::atomic<int> cached[64];
    static const int reserve = [] {
        const char* e = getenv("RQ_SM_RESERVE");
        return e ? atoi(e) : 0;
    }();
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (!n) {
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
        if (reserve > 0 && reserve < n) n -= reserve;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

bool pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("RQ_NO_PDL");
        return !(e && atoi(e) != 0);
    }();
    return on;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per (function, device): set it once for each pair.
cudaError_t ensure_max_dynamic_smem(const void* fn, int bytes) {
    struct Entry { const void* fn; int dev; int bytes; };
    static std::mutex mu;
    static std::vector<Entry> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    for (const Entry& d : done)
        if (d.fn == fn && d.dev == dev && d.bytes >= bytes) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done.push_back({fn, dev, bytes});
    return e;
}

static bool pow2_dim(int D) { return D == 4 || D == 8 || D == 16 || D == 32 || D == 64 || D == 128; }

}  // namespace rq

using namespace rq;

extern "C" int rq_abi_version(void) { return RQ_ABI_VERSION; }
extern "C" const char* rq_last_error(void) { return g_err; }
extern "C" uint64_t rq_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int rq_tcgen05_supported(int D, int K, int L, int mode, int distance) {
    if (mode == RQ_MODE_GUMBEL_SOFTMAX) return gumbel_tc_supported(D, K, L, distance) ? 1 : 0;
    return (tc_supported(D, K, L, mode, distance) || tc64_bigk_supported(D, K, L, mode, distance)) ? 1 : 0;
}

// the Gumbel-softmax launch goes to the tensor-core path (rq_gumbel_tc.cu)
static bool gumbel_tc_taken(int64_t B, int D, int K, int L, int mode, int distance, int impl) {
    return mode == RQ_MODE_GUMBEL_SOFTMAX && gumbel_tc_supported(D, K, L, distance) &&
           (impl == RQ_IMPL_TCGEN05 || (impl == RQ_IMPL_AUTO && B >= 4096));
}
extern "C" int rq_gumbel_weights_saved(int64_t B, int D, int K, int L, int mode, int distance, int impl) {
    return gumbel_tc_taken(B, D, K, L, mode, distance, impl) ? 1 : 0;
}

extern "C" int rq_quantize_forward(const float* x, const float* codebooks, int64_t B, int D, int K, int L, int mode,
                                   int distance, float beta, float temperature, const float* gumbel_u,
                                   uint64_t seed, uint64_t offset, float* emb_out, float* residuals, int64_t* ids,
                                   float* loss, int impl, rq_stream_t stream) {
    return rq_quantize_forward_ws(x, codebooks, B, D, K, L, mode, distance, beta, temperature, gumbel_u, seed, offset,
                                  emb_out, residuals, ids, loss, nullptr, impl, stream);
}

extern "C" int rq_quantize_forward_ws(const float* x, const float* codebooks, int64_t B, int D, int K, int L, int mode,
                                      int distance, float beta, float temperature, const float* gumbel_u,
                                      uint64_t seed, uint64_t offset, float* emb_out, float* residuals, int64_t* ids,
                                      float* loss, float* gumbel_weights, int impl, rq_stream_t stream) {
    RQ_CHECK_ARG(B >= 0 && K >= 1 && L >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_quantize_forward: need B >= 0, K >= 1, L >= 1");
    RQ_CHECK_ARG(pow2_dim(D), RQ_ERR_UNSUPPORTED_SHAPE, "rq_quantize_forward: unsupported embed_dim D=%d (4..128, power of 2)", D);
    RQ_CHECK_ARG(mode >= RQ_MODE_EVAL && mode <= RQ_MODE_ROTATION_TRICK, RQ_ERR_INVALID_ARGUMENT,
                 "Unsupported Quantize forward mode.");
    RQ_CHECK_ARG(distance == RQ_DIST_L2 || distance == RQ_DIST_COSINE, RQ_ERR_INVALID_ARGUMENT,
                 "Unsupported Quantize distance mode.");
    RQ_CHECK_ARG(!(mode == RQ_MODE_GUMBEL_SOFTMAX && !(temperature > 0.f)), RQ_ERR_INVALID_ARGUMENT,
                 "rq_quantize_forward: temperature must be > 0");
    if (B == 0) return 0;
    RQ_CHECK_ARG(x && codebooks && ids, RQ_ERR_NULL_POINTER, "rq_quantize_forward: x, codebooks and ids are required");
    FwdParams p;
    p.x = x; p.cb = codebooks; p.B = B; p.K = K; p.L = L; p.mode = mode; p.dist = distance;
    p.beta = beta; p.temp = temperature; p.gu = gumbel_u; p.seed = seed; p.offset = offset;
    p.emb = emb_out; p.res = residuals; p.ids = ids; p.loss = loss; p.KC = 0;
    cudaStream_t st = (cudaStream_t)stream;
    // the tcgen05 family stages the codebooks with TMA bulk copies (16-byte aligned source rows)
    const bool aligned = (((uintptr_t)codebooks & 15) == 0) && (((uintptr_t)x & 31) == 0);
    const bool tc_ok = tc_supported(D, K, L, mode, distance) && aligned;
    // Gumbel-softmax: GEMM-shaped contractions on the tensor cores (rq_gumbel_tc.cu) once the batch is worth the
    // [B, K] scratch; the thread-per-row family below serves small batches and the shapes outside D in {32, 64}
    if (gumbel_tc_taken(B, D, K, L, mode, distance, impl)) {
        RQ_CHECK_ARG((((uintptr_t)x & 15) == 0) && (((uintptr_t)codebooks & 15) == 0) && (((uintptr_t)gumbel_weights & 15) == 0),
                     RQ_ERR_INVALID_ARGUMENT, "rq_quantize_forward: Gumbel-softmax on the tensor cores needs 16-byte aligned x, codebooks and weights");
        return quantize_forward_gumbel_tc(p, D, st, gumbel_weights);
    }
    if (tc64_bigk_supported(D, K, L, mode, distance) && aligned && (impl == RQ_IMPL_TCGEN05 || (impl == RQ_IMPL_AUTO && B >= 2048)))
        return quantize_forward_tc64_bigk(p, st);
    if (impl == RQ_IMPL_TCGEN05) {
        RQ_CHECK_ARG(tc_ok, RQ_ERR_UNSUPPORTED_SHAPE,
                     "rq_quantize_forward: (D=%d,K=%d,L=%d,mode=%d,dist=%d) is not served by the tcgen05 family", D, K, L,
                     mode, distance);
        return quantize_forward_tc(p, D, st);
    }
    if (impl == RQ_IMPL_AUTO && tc_ok) return quantize_forward_tc(p, D, st);
    RQ_CHECK_ARG(impl == RQ_IMPL_AUTO || impl == RQ_IMPL_SIMT, RQ_ERR_INVALID_ARGUMENT, "rq_quantize_forward: bad impl %d", impl);
    return quantize_forward_simt(p, D, st);
}

extern "C" int rq_quantize_backward(const float* g_emb, const float* g_res, const float* g_loss,
                                    const float* residuals, const int64_t* ids, const float* codebooks, int64_t B,
                                    int D, int K, int L, int mode, float beta, float temperature,
                                    const float* gumbel_u, uint64_t seed, uint64_t offset, float* g_x,
                                    float* g_codebooks, rq_stream_t stream) {
    return rq_quantize_backward_ex(g_emb, g_res, g_loss, residuals, ids, codebooks, B, D, K, L, mode, beta, temperature,
                                   gumbel_u, seed, offset, g_x, g_codebooks, 0, stream);
}

extern "C" int rq_quantize_backward_ex(const float* g_emb, const float* g_res, const float* g_loss,
                                       const float* residuals, const int64_t* ids, const float* codebooks, int64_t B,
                                       int D, int K, int L, int mode, float beta, float temperature,
                                       const float* gumbel_u, uint64_t seed, uint64_t offset, float* g_x,
                                       float* g_codebooks, int flags, rq_stream_t stream) {
    return rq_quantize_backward_ws(g_emb, g_res, g_loss, residuals, ids, codebooks, B, D, K, L, mode, beta, temperature,
                                   gumbel_u, seed, offset, g_x, g_codebooks, nullptr, flags, stream);
}

extern "C" int rq_quantize_backward_ws(const float* g_emb, const float* g_res, const float* g_loss,
                                       const float* residuals, const int64_t* ids, const float* codebooks, int64_t B,
                                       int D, int K, int L, int mode, float beta, float temperature,
                                       const float* gumbel_u, uint64_t seed, uint64_t offset, float* g_x,
                                       float* g_codebooks, const float* gumbel_weights, int flags, rq_stream_t stream) {
    RQ_CHECK_ARG(B >= 0 && K >= 1 && L >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_quantize_backward: need B >= 0, K >= 1, L >= 1");
    RQ_CHECK_ARG(pow2_dim(D), RQ_ERR_UNSUPPORTED_SHAPE, "rq_quantize_backward: unsupported embed_dim D=%d", D);
    RQ_CHECK_ARG(mode >= RQ_MODE_EVAL && mode <= RQ_MODE_ROTATION_TRICK, RQ_ERR_INVALID_ARGUMENT,
                 "Unsupported Quantize forward mode.");
    RQ_CHECK_ARG(g_codebooks && (g_x || B == 0), RQ_ERR_NULL_POINTER, "rq_quantize_backward: g_x and g_codebooks are required");
    RQ_CHECK_ARG(B == 0 || (residuals && ids && codebooks), RQ_ERR_NULL_POINTER,
                 "rq_quantize_backward: residuals, ids and codebooks are required");
    BwdParams p;
    p.g_emb = g_emb; p.g_res = g_res; p.g_loss = g_loss; p.res = residuals; p.ids = ids; p.cb = codebooks;
    p.B = B; p.K = K; p.L = L; p.mode = mode; p.beta = beta; p.g_x = g_x; p.g_cb = g_codebooks;
    p.accumulate = (flags & RQ_BWD_ACCUMULATE) ? 1 : 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == RQ_MODE_GUMBEL_SOFTMAX) {
        RQ_CHECK_ARG(temperature > 0.f, RQ_ERR_INVALID_ARGUMENT, "rq_quantize_backward: temperature must be > 0");
        // the tensor-core path once the batch is worth its [B, K] scratch (as in the forward); RQ_BWD_THREAD_PER_ROW
        // keeps the CUDA-core kernel (tests compare the two)
        const bool ptr_ok = (((uintptr_t)residuals & 15) == 0) && (((uintptr_t)codebooks & 15) == 0) &&
                            (!g_emb || ((uintptr_t)g_emb & 15) == 0) && (!g_res || ((uintptr_t)g_res & 15) == 0) &&
                            (((uintptr_t)g_x & 15) == 0);
        // saved weights exist only when the forward took the tensor-core path, whatever the batch size
        if (!(flags & RQ_BWD_THREAD_PER_ROW) && (B >= 4096 || gumbel_weights) && gumbel_tc_supported(D, K, L, RQ_DIST_L2) &&
            ptr_ok && (((uintptr_t)gumbel_weights & 15) == 0))
            return quantize_backward_gumbel_tc(p, D, temperature, gumbel_u, seed, offset, st, gumbel_weights);
        RQ_CHECK_ARG(!gumbel_weights, RQ_ERR_INVALID_ARGUMENT,
                     "rq_quantize_backward: saved Gumbel weights are only consumed by the tensor-core path");
        return quantize_backward_gumbel(p, D, temperature, gumbel_u, seed, offset, st);
    }
    return quantize_backward_simt(p, D, st);
}

extern "C" int rq_kmeans_assign_accumulate(const float* x, const float* centroids, int64_t B, int D, int K,
                                           int64_t* assign, int has_prev, float* sums, int64_t* counts,
                                           int64_t* n_changed, int impl, rq_stream_t stream) {
    RQ_CHECK_ARG(B >= 0 && K >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_kmeans_assign_accumulate: need B >= 0, K >= 1");
    RQ_CHECK_ARG(pow2_dim(D), RQ_ERR_UNSUPPORTED_SHAPE, "rq_kmeans_assign_accumulate: unsupported D=%d", D);
    RQ_CHECK_ARG(sums && counts && n_changed && centroids, RQ_ERR_NULL_POINTER, "rq_kmeans_assign_accumulate: null pointer");
    RQ_CHECK_ARG(B == 0 || (x && assign), RQ_ERR_NULL_POINTER, "rq_kmeans_assign_accumulate: null pointer");
    KmeansParams p;
    p.x = x; p.centroids = centroids; p.B = B; p.K = K; p.assign = assign; p.has_prev = has_prev;
    p.sums = sums; p.counts = (unsigned long long*)counts; p.n_changed = (unsigned long long*)n_changed; p.KC = 0;
    // AUTO: the tcgen05 distance kernel when it serves the shape (it computes the expanded form ||c||^2 - 2x.c
    // with fp32 re-scoring; assignments can differ from the direct form only inside the fp32 tie band)
    const bool tc_ok = tc_supported(D, K, 1, RQ_MODE_EVAL, RQ_DIST_L2) && (((uintptr_t)centroids & 15) == 0) &&
                       (((uintptr_t)x & 31) == 0);
    if (impl == RQ_IMPL_TCGEN05) {
        RQ_CHECK_ARG(tc_ok, RQ_ERR_UNSUPPORTED_SHAPE, "rq_kmeans_assign_accumulate: (D=%d,K=%d) is not served by the tcgen05 family", D, K);
        return kmeans_assign_accumulate_tc(p, D, (cudaStream_t)stream);
    }
    if (impl == RQ_IMPL_AUTO && tc_ok && B >= 4096) return kmeans_assign_accumulate_tc(p, D, (cudaStream_t)stream);
    return kmeans_assign_accumulate_simt(p, D, (cudaStream_t)stream);
}

extern "C" int rq_kmeans_update(const float* sums, const int64_t* counts, float* centroids, int D, int K,
                                float* max_shift, int64_t* n_empty, int64_t* empty_ids, rq_stream_t stream) {
    RQ_CHECK_ARG(K >= 1 && D >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_kmeans_update: need K >= 1, D >= 1");
    RQ_CHECK_ARG(sums && counts && centroids && max_shift && n_empty && empty_ids, RQ_ERR_NULL_POINTER,
                 "rq_kmeans_update: null pointer");
    return kmeans_update(sums, counts, centroids, D, K, max_shift, n_empty, empty_ids, (cudaStream_t)stream);
}

// Host-buffer tokenisation as a three-stage pipeline over row chunks: H2D of chunk i+1, the fused
// forward of chunk i and the D2H of chunk i-1 run on three streams (PCIe is full duplex), device
// memory is two chunk-sized buffer pairs regardless of B.  `stream` carries the compute stage and
// orders the call after the caller's earlier work; the call returns when ids_host is complete.
static int64_t host_chunk_rows() {
    const char* e = getenv("RQ_HOST_CHUNK_ROWS");
    const int64_t rows = e ? atoll(e) : 0;
    return rows > 0 ? rows : 262144;
}

// Scratch of the host entry comes from a library-owned stream-ordered pool that keeps freed blocks (the default
// pool's release threshold is 0: it would return the chunk buffers to the driver at every synchronisation and
// re-map them on the next call, which costs milliseconds).
static cudaMemPool_t host_entry_pool() {
    static cudaMemPool_t pools[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (!pools[dev]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        // up to 4 GiB stays cached between calls: the Gumbel-softmax backward's [B, K] scratch is 1.1 GB at 65,536 x 1,024
        // and would otherwise be handed back to the driver at every synchronisation
        uint64_t keep = 4ull << 30;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        pools[dev] = pool;
    }
    return pools[dev];
}
namespace rq {
void* engine_pool_alloc(size_t bytes, cudaStream_t st) {
    void* ptr = nullptr;
    cudaMemPool_t pool = host_entry_pool();
    cudaError_t e = pool ? cudaMallocFromPoolAsync(&ptr, bytes, pool, st) : cudaMallocAsync(&ptr, bytes, st);
    if (e != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return ptr;
}
}  // namespace rq

static cudaError_t host_entry_alloc(void** p, size_t bytes, cudaStream_t st) {
    cudaMemPool_t pool = host_entry_pool();
    return pool ? cudaMallocFromPoolAsync(p, bytes, pool, st) : cudaMallocAsync(p, bytes, st);
}

// Side streams and events of the host entry: created once per device and reused (a call used to create and destroy
// 2 streams + 7 events).  A second caller on the same device while the set is in use gets a fresh, temporary one.
struct HostPlumbing {
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[2] = {}, ev_comp[2] = {}, ev_out[2] = {}, ev_start = nullptr;
    bool cached = false;
    cudaError_t create() {
        cudaError_t e = cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming);
        for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
            e = cudaEventCreateWithFlags(&ev_in[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_comp[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_out[b], cudaEventDisableTiming);
        }
        return e;
    }
    void destroy() {
        for (int b = 0; b < 2; ++b) {
            if (ev_in[b]) cudaEventDestroy(ev_in[b]);
            if (ev_comp[b]) cudaEventDestroy(ev_comp[b]);
            if (ev_out[b]) cudaEventDestroy(ev_out[b]);
            ev_in[b] = ev_comp[b] = ev_out[b] = nullptr;
        }
        if (ev_start) cudaEventDestroy(ev_start);
        if (s_in) cudaStreamDestroy(s_in);
        if (s_out) cudaStreamDestroy(s_out);
        ev_start = nullptr; s_in = s_out = nullptr;
    }
};
static std::mutex g_plumbing_mu;
static HostPlumbing g_plumbing[64];
static bool g_plumbing_busy[64] = {};
// returns the device's cached set when it is free (marks it busy), otherwise a temporary set; `e` reports creation errors
static HostPlumbing acquire_plumbing(int dev, cudaError_t& e) {
    e = cudaSuccess;
    if (dev >= 0 && dev < 64) {
        std::lock_guard<std::mutex> lk(g_plumbing_mu);
        if (!g_plumbing_busy[dev]) {
            if (!g_plumbing[dev].s_in) {
                e = g_plumbing[dev].create();
                if (e != cudaSuccess) { g_plumbing[dev].destroy(); HostPlumbing t; return t; }
                g_plumbing[dev].cached = true;
            }
            g_plumbing_busy[dev] = true;
            return g_plumbing[dev];
        }
    }
    HostPlumbing t;
    e = t.create();
    return t;
}
static void release_plumbing(int dev, HostPlumbing& h) {
    if (h.cached) {
        std::lock_guard<std::mutex> lk(g_plumbing_mu);
        g_plumbing_busy[dev] = false;
    } else {
        h.destroy();
    }
}

extern "C" int rq_semantic_ids_host(const float* x_host, const float* codebooks, int64_t B, int D, int K, int L,
                                    int distance, int64_t* ids_host, float* loss_host, int impl,
                                    rq_stream_t stream) {
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(B >= 0, RQ_ERR_INVALID_ARGUMENT, "rq_semantic_ids_host: B < 0");
    if (B == 0) return 0;
    RQ_CHECK_ARG(x_host && codebooks && ids_host, RQ_ERR_NULL_POINTER, "rq_semantic_ids_host: null pointer");
    RQ_CHECK_ARG(L >= 1 && D >= 1, RQ_ERR_INVALID_ARGUMENT, "rq_semantic_ids_host: need L >= 1, D >= 1");
    const int64_t chunk = host_chunk_rows();
    const int64_t C = B < chunk ? B : chunk;
    const int64_t n_chunks = (B + C - 1) / C;
    const int NB = n_chunks > 1 ? 2 : 1;
    float* dx[2] = {nullptr, nullptr};
    int64_t* dids[2] = {nullptr, nullptr};
    float* dloss[2] = {nullptr, nullptr};
    int rc = 0, dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    cudaError_t e_pl = cudaSuccess;
    HostPlumbing pl = acquire_plumbing(dev, e_pl);
    if (e == cudaSuccess) e = e_pl;
    cudaStream_t s_in = pl.s_in, s_out = pl.s_out;
    cudaEvent_t* ev_in = pl.ev_in;
    cudaEvent_t* ev_comp = pl.ev_comp;
    cudaEvent_t* ev_out = pl.ev_out;
    cudaEvent_t ev_start = pl.ev_start;
#define HOST_TRY(call)                                                                        \
    do { if (rc == 0 && e == cudaSuccess) { e = (call); } } while (0)
    for (int b = 0; b < NB; ++b) {
        HOST_TRY(host_entry_alloc((void**)&dx[b], (size_t)C * D * sizeof(float), st));
        HOST_TRY(host_entry_alloc((void**)&dids[b], (size_t)L * C * sizeof(int64_t), st));
        if (loss_host) HOST_TRY(host_entry_alloc((void**)&dloss[b], (size_t)C * sizeof(float), st));
    }
    // the side streams start after the allocations (and whatever the caller queued before us)
    HOST_TRY(cudaEventRecord(ev_start, st));
    HOST_TRY(cudaStreamWaitEvent(s_in, ev_start, 0));
    HOST_TRY(cudaStreamWaitEvent(s_out, ev_start, 0));
    for (int64_t i = 0; i < n_chunks && rc == 0 && e == cudaSuccess; ++i) {
        const int b = (int)(i & 1);
        const int64_t lo = i * C, n = (B - lo) < C ? (B - lo) : C;
        if (i >= 2) HOST_TRY(cudaStreamWaitEvent(s_in, ev_comp[b], 0));     // x buffer b consumed
        HOST_TRY(cudaMemcpyAsync(dx[b], x_host + (size_t)lo * D, (size_t)n * D * sizeof(float),
                                 cudaMemcpyHostToDevice, s_in));
        HOST_TRY(cudaEventRecord(ev_in[b], s_in));
        HOST_TRY(cudaStreamWaitEvent(st, ev_in[b], 0));
        if (i >= 2) HOST_TRY(cudaStreamWaitEvent(st, ev_out[b], 0));        // ids buffer b drained
        if (rc == 0 && e == cudaSuccess)
            rc = rq_quantize_forward(dx[b], codebooks, n, D, K, L, RQ_MODE_EVAL, distance, 0.25f, 1.0f, nullptr, 0,
                                     0, nullptr, nullptr, dids[b], dloss[b], impl, stream);
        HOST_TRY(cudaEventRecord(ev_comp[b], st));
        HOST_TRY(cudaStreamWaitEvent(s_out, ev_comp[b], 0));
        // device ids are [L, n] for this chunk; the host table is [L, B]
        HOST_TRY(cudaMemcpy2DAsync(ids_host + lo, (size_t)B * sizeof(int64_t), dids[b], (size_t)n * sizeof(int64_t),
                                   (size_t)n * sizeof(int64_t), (size_t)L, cudaMemcpyDeviceToHost, s_out));
        if (loss_host)
            HOST_TRY(cudaMemcpyAsync(loss_host + lo, dloss[b], (size_t)n * sizeof(float), cudaMemcpyDeviceToHost,
                                     s_out));
        HOST_TRY(cudaEventRecord(ev_out[b], s_out));
    }
#undef HOST_TRY
    // drain every stage before the buffers go back to the pool, whatever happened above
    cudaError_t e1 = s_in ? cudaStreamSynchronize(s_in) : cudaSuccess;
    cudaError_t e2 = s_out ? cudaStreamSynchronize(s_out) : cudaSuccess;
    for (int b = 0; b < NB; ++b) {
        if (dx[b]) cudaFreeAsync(dx[b], st);
        if (dids[b]) cudaFreeAsync(dids[b], st);
        if (dloss[b]) cudaFreeAsync(dloss[b], st);
    }
    cudaError_t e3 = cudaStreamSynchronize(st);
    release_plumbing(dev, pl);
    if (rc != 0) return rc;
    for (cudaError_t x : {e, e1, e2, e3})
        if (x != cudaSuccess) {
            set_error("rq_semantic_ids_host: %s", cudaGetErrorString(x));
            return (int)x;
        }
    return 0;
}
