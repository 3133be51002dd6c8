// Codebooks larger than one UMMA tile (K > 256) at D = 64 — BASELINE config 5: 4 levels x 1024 codes.
//
// The operand tile of 1024 x 64 codes (128 KB in fp16) leaves no room for more than one level per SM, so
// the levels run one after the other (level-major) and every level is a composition of launches:
//
//   rq_fwd_tc64_kernel with L = 1, chunked: CTA b scores every row of the current residual against the
//       256-code chunk b % n_chunks of the level's codebook        -> per chunk, winner id and its exact
//                                                                          fp32 distance per row (best_dist)
//   rq_bigk_merge_kernel: the chunk whose winner has the smallest exact distance (lowest chunk on a tie,
//       i.e. the lowest code index), codeword gather, estimator, loss, outputs, next residual.
//
// Each chunk winner is the exact fp32 argmin of its chunk (rq_quantize_tc64.cu) and its distance is computed
// with the SIMT family's arithmetic, so the merged id is the SIMT family's id bit for bit; the row epilogue is
// the shared level_epilogue<D>.  The residual travels through HBM between levels (it is an output of the
// training forward anyway); scratch comes from the library's stream-ordered pool.
#include <stdlib.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"

namespace rq {

constexpr int kBigkChunk = 256;
constexpr int kBigkThreads = 128;

struct BigkMergeParams {
    const float* r_in;        // [B, D] residual entering the level
    const float* cb;          // [K, D] this level's codebook
    const int64_t* cand_ids;  // [n_chunks][B] winner inside each chunk
    const float* cand_d;      // [n_chunks][B] its exact distance
    int64_t B;
    int K, n_chunks, mode, first;
    float beta;
    float* emb;               // [B, D] or nullptr
    float* res;               // [B, D] or nullptr
    int64_t* ids;             // [B]
    float* loss;              // [B] or nullptr (accumulated over the levels)
    float* r_out;             // [B, D] or nullptr (last level)
};

template <int D>
__global__ void __launch_bounds__(kBigkThreads) rq_bigk_merge_kernel(BigkMergeParams m) {
    const int64_t row = (int64_t)blockIdx.x * kBigkThreads + threadIdx.x;
    if (row >= m.B) return;
    int bc = 0;
    float bd = __ldg(m.cand_d + row);
    for (int c = 1; c < m.n_chunks; ++c) {
        const float d = __ldg(m.cand_d + (int64_t)c * m.B + row);
        if (d < bd) { bd = d; bc = c; }                       // strict: the lowest chunk wins a tie
    }
    int k = bc * kBigkChunk + (int)__ldg(m.cand_ids + (int64_t)bc * m.B + row);
    if (k >= m.K) k = m.K - 1;
    float r[D], e[D];
    load_row<D>(m.r_in + row * D, r);
    load_row<D>(m.cb + (int64_t)k * D, e);
    const float xx = dot_self<D>(r);
    if (m.res) store_row<D>(m.res + row * D, r);
    const float lvl = level_epilogue<D>(m.mode, m.beta, r, e, xx, e);     // estimator value in place
    if (m.emb) store_row<D>(m.emb + row * D, e);
    m.ids[row] = k;
    if (m.loss) m.loss[row] = m.first ? lvl : m.loss[row] + lvl;
    if (m.r_out) {
        residual_update<D>(r, e);                                          // rqvae.py:129
        store_row<D>(m.r_out + row * D, r);
    }
}

bool tc64_bigk_supported(int D, int K, int L, int mode, int distance) {
    // also serves K <= 256 with more levels than the fused d=64 kernel keeps resident (L > 3)
    return D == 64 && K >= 1 && L >= 1 && (K > kBigkChunk || L > 3) && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

int quantize_forward_tc64_bigk(const FwdParams& p, cudaStream_t st) {
    constexpr int D = 64;
    if (p.B == 0) return 0;
    const int n_chunks = (p.K + kBigkChunk - 1) / kBigkChunk;
    int64_t* cand_ids = (int64_t*)engine_pool_alloc((size_t)n_chunks * p.B * sizeof(int64_t), st);
    float* cand_d = (float*)engine_pool_alloc((size_t)n_chunks * p.B * sizeof(float), st);
    float* rbuf[2] = {nullptr, nullptr};
    if (p.L > 1) {
        rbuf[0] = (float*)engine_pool_alloc((size_t)p.B * D * sizeof(float), st);
        if (p.L > 2) rbuf[1] = (float*)engine_pool_alloc((size_t)p.B * D * sizeof(float), st);
    }
    int rc = 0;
    if (!cand_ids || !cand_d || (p.L > 1 && !rbuf[0]) || (p.L > 2 && !rbuf[1])) {
        set_error("rq_quantize_forward: out of device memory for the K > 256 scratch buffers");
        rc = (int)cudaErrorMemoryAllocation;
    }
    const float* r_in = p.x;
    for (int l = 0; l < p.L && rc == 0; ++l) {
        const float* cbl = p.cb + (int64_t)l * p.K * D;
        {
            // one launch scores every row against all chunks: CTA b serves chunk b % n_chunks
            FwdParams q;
            q.x = r_in; q.cb = cbl; q.B = p.B; q.K = p.K;
            q.L = 1; q.mode = RQ_MODE_EVAL; q.dist = RQ_DIST_L2; q.beta = p.beta; q.temp = p.temp;
            q.gu = nullptr; q.seed = 0; q.offset = 0;
            q.emb = nullptr; q.res = nullptr; q.loss = nullptr; q.KC = 0;
            q.ids = cand_ids;
            q.best_dist = cand_d;
            q.n_chunks = n_chunks;
            rc = quantize_forward_tc64(q, st);
        }
        if (rc) break;
        BigkMergeParams m;
        m.r_in = r_in; m.cb = cbl; m.cand_ids = cand_ids; m.cand_d = cand_d; m.B = p.B; m.K = p.K;
        m.n_chunks = n_chunks; m.mode = p.mode; m.first = (l == 0); m.beta = p.beta;
        m.emb = p.emb ? p.emb + (int64_t)l * p.B * D : nullptr;
        m.res = p.res ? p.res + (int64_t)l * p.B * D : nullptr;
        m.ids = p.ids + (int64_t)l * p.B;
        m.loss = p.loss;
        m.r_out = (l + 1 < p.L) ? rbuf[l & 1] : nullptr;
        const unsigned grid = (unsigned)((p.B + kBigkThreads - 1) / kBigkThreads);
        rq_bigk_merge_kernel<D><<<grid, kBigkThreads, 0, st>>>(m);
        RQ_LAUNCHED();
        r_in = m.r_out;
    }
    if (cand_ids) cudaFreeAsync(cand_ids, st);
    if (cand_d) cudaFreeAsync(cand_d, st);
    if (rbuf[0]) cudaFreeAsync(rbuf[0], st);
    if (rbuf[1]) cudaFreeAsync(rbuf[1], st);
    return rc;
}

}  // namespace rq
