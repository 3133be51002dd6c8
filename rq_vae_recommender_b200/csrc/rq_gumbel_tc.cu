// Gumbel-softmax estimator on the tensor cores (modules/quantize.py:124-129 + distributions/gumbel.py:7-17 of the
// reference, and their autograd): level by level, the two GEMM-shaped contractions of the forward — x C^T and W C —
// and the five of the backward run as fp32-grade tcgen05 products (rq_gemm_split.cu: three bf16 planes per value), the
// [B, K] logits / weights live in one scratch matrix that the row kernels below rewrite in place, and the ids come
// from the exact distance kernels of the other modes (one single-level ids-only launch per level).  The CUDA-core
// family (rq_quantize_simt.cu / rq_gumbel_bwd.cu: one thread per row, 4 K D FMAs per row and level) stays as the
// small-batch path and as the yardstick of the tests.
//
//   forward, per level     ids   = argmin dist(r, C)                         exact kernel, L = 1
//                          S     = r C^T                                     GEMM  [B, D] x [K, D]^T
//                          W     = softmax((-(|r|^2 + |c|^2 - 2 S) + g) / T) rows, in place; g from the given uniforms
//                                                                            or the SIMT family's Philox counters
//                          emb   = W C                                       GEMM  [B, K] x [D, K]^T
//                          loss += (1 + beta) |r - emb|^2,  r <- r - emb      rows
//   backward, per level (top down, G = d objective / d r_{l+1})
//                          S, W, emb as above (recomputed from the saved residual)
//                          de    = g_emb - G + 2 (emb - x) g_loss,  t = de . emb
//                          dW    = de C^T                                    GEMM  [B, D] x [K, D]^T
//                          dd    = -W (dW - t) / T   (in place), and W^T, dd^T for the codebook gradient
//                          v     = dd C                                      GEMM  [B, K] x [D, K]^T
//                          g_C  += W^T de - 2 dd^T x + 2 C * colsum(dd)       GEMMs [K, B] x [D, B]^T, split over B
//                          G     = G + 2 beta (x - emb) g_loss + g_res - 2 v  (the 2 x sum_k dd_k term is zero: the rows
//                                                                            of a softmax Jacobian sum to nothing)
#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"

namespace rq {

int split_gemm(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd, int64_t lda, int64_t ldb,
               int64_t ldc, float alpha, int atomic, int k_splits, cudaStream_t st, int terms = 6);
int split_gemm_tn(const float* At, const float* Bt, float* C, int64_t M, int64_t N, int64_t Kd, int64_t ldat, int64_t ldbt,
                  int64_t ldc, float alpha, int atomic, int k_splits, cudaStream_t st, int terms = 6);
int split_gemm_softmax_jacobian(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd, int64_t lda,
                                int64_t ldb, int64_t ldc, const float* Wm, const float* t, float scale, float* csum,
                                cudaStream_t st);

namespace gtc {

// ||c||^2 in the canonical order of code_sqnorm, and C^T ([D, Kp], zero padded) for the GEMMs whose contraction runs
// over the codes.  One thread per code.
template <int D>
__global__ void __launch_bounds__(256) codebook_prep_kernel(const float* __restrict__ cb, int K, int Kp, float* __restrict__ cc,
                                                            float* __restrict__ ct) {
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= Kp) return;
    if (k < K) {
        cc[k] = code_sqnorm<D>(cb + (int64_t)k * D);
#pragma unroll 4
        for (int j = 0; j < D; ++j) ct[(int64_t)j * Kp + k] = __ldg(cb + (int64_t)k * D + j);
    } else {
        cc[k] = 0.f;
        for (int j = 0; j < D; ++j) ct[(int64_t)j * Kp + k] = 0.f;
    }
}

struct RowParams {
    float* S;            // [B, Kp]: in: r C^T, out: softmax weights
    const float* r;      // [B, D] this level's residual
    const float* cc;     // [Kp]
    const float* gu;     // [B, K] this level's uniforms or nullptr (Philox)
    int64_t B;
    int K, Kp, level;
    float inv_t;
    uint64_t seed, offset;
};

// One warp per row, a lane owns the codes 4 (lane + 32 i) .. + 3: float4 accesses (512 contiguous bytes per warp
// instruction) and one Philox call per four codes, with the SIMT family's counters (row, code / 4, level).
template <int D>
__global__ void __launch_bounds__(256) gumbel_softmax_rows_kernel(RowParams q) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= q.B) return;
    // |r|^2, canonical chunk order (rq_rowmath.cuh): D / 8 chunks over the first lanes, combined by a balanced tree
    float xx;
    {
        float part = 0.f;
        if (lane < D / 8) {
            float v[8];
            const float4 a = __ldg(reinterpret_cast<const float4*>(q.r + row * D + lane * 8));
            const float4 b = __ldg(reinterpret_cast<const float4*>(q.r + row * D + lane * 8 + 4));
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            part = chunk_dot<8>(v, v);
        }
#pragma unroll
        for (int o = 1; o < D / 8; o <<= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        xx = __shfl_sync(0xffffffffu, part, 0);
    }
    float* Srow = q.S + row * q.Kp;
    const float* gurow = q.gu ? q.gu + row * q.K : nullptr;
    // logits of the four codes k .. k + 3 from the products s (quantize.py:108-112, 125-126)
    auto logits4 = [&](int k, const float4& s, float (&y)[4]) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(q.cc + k));
        float u[4];
        if (gurow) {
#pragma unroll
            for (int t = 0; t < 4; ++t) u[t] = (k + t < q.K) ? __ldg(gurow + k + t) : 0.5f;
        } else {
            const uint4 rnd = philox4x32_10(make_uint4((uint32_t)row, (uint32_t)(row >> 32), (uint32_t)(k >> 2), (uint32_t)q.level),
                                            make_uint2((uint32_t)(q.seed ^ q.offset), (uint32_t)((q.seed >> 32) ^ (q.offset >> 32))));
            u[0] = u01(rnd.x); u[1] = u01(rnd.y); u[2] = u01(rnd.z); u[3] = u01(rnd.w);
        }
        const float sv[4] = {s.x, s.y, s.z, s.w}, cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const float d = (xx + cv[t]) - 2.f * sv[t];
            // gumbel.py:9-10; the outer log in its fast form: its argument lies in [1e-20, 47] and what matters is the
            // ABSOLUTE error of g (2e-7 for MUFU.LG2 near 1), the inner one stays accurate (u -> 1 makes -log u tiny)
            const float g = -__logf(-logf(u[t] + 1e-20f) + 1e-20f);
            y[t] = (k + t < q.K) ? (-d + g) * q.inv_t : -INFINITY;
        }
    };
    if (q.Kp <= 1024) {
        // the row fits the warp's registers (8 float4 per lane): one read and one write of the [B, K] matrix
        float y[8][4];
        float m = -INFINITY;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int k = lane * 4 + 128 * i;
            if (k < q.Kp) {
                logits4(k, *reinterpret_cast<const float4*>(Srow + k), y[i]);
                m = fmaxf(fmaxf(m, fmaxf(y[i][0], y[i][1])), fmaxf(y[i][2], y[i][3]));
            } else {
                y[i][0] = y[i][1] = y[i][2] = y[i][3] = -INFINITY;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
#pragma unroll
            for (int t = 0; t < 4; ++t) y[i][t] = __expf(y[i][t] - m);
            sum += (y[i][0] + y[i][1]) + (y[i][2] + y[i][3]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const float inv = 1.0f / sum;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int k = lane * 4 + 128 * i;
            if (k < q.Kp) *reinterpret_cast<float4*>(Srow + k) = make_float4(y[i][0] * inv, y[i][1] * inv, y[i][2] * inv, y[i][3] * inv);
        }
        return;
    }
    // longer rows: three passes in place (the row is L1 / L2 resident after the first)
    float m = -INFINITY;
    for (int k = lane * 4; k < q.Kp; k += 128) {
        float y[4];
        logits4(k, *reinterpret_cast<const float4*>(Srow + k), y);
        m = fmaxf(fmaxf(m, fmaxf(y[0], y[1])), fmaxf(y[2], y[3]));
        *reinterpret_cast<float4*>(Srow + k) = make_float4(y[0], y[1], y[2], y[3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float sum = 0.f;
    for (int k = lane * 4; k < q.Kp; k += 128) {
        float4 y = *reinterpret_cast<const float4*>(Srow + k);
        y.x = __expf(y.x - m); y.y = __expf(y.y - m); y.z = __expf(y.z - m); y.w = __expf(y.w - m);
        sum += (y.x + y.y) + (y.z + y.w);
        *reinterpret_cast<float4*>(Srow + k) = y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.0f / sum;
    for (int k = lane * 4; k < q.Kp; k += 128) {
        float4 y = *reinterpret_cast<const float4*>(Srow + k);
        y.x *= inv; y.y *= inv; y.z *= inv; y.w *= inv;
        *reinterpret_cast<float4*>(Srow + k) = y;
    }
}

struct TailParams {
    const float* r;      // [B, D]
    const float* emb;    // [B, D]
    float* res_out;      // [B, D] or nullptr: this level's residual output
    float* r_next;       // [B, D] or nullptr
    float* loss;         // [B] or nullptr
    int64_t B;
    float beta;
    int first;
};

// D / 4 lanes per row, float4 each: the level's residual output, loss term and next residual (rqvae.py:126-129)
template <int D>
__global__ void __launch_bounds__(256) gumbel_level_tail_kernel(TailParams q) {
    constexpr int LPR = D / 4;
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int64_t row = i / LPR;
    const bool active = row < q.B;
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f), e = r;
    if (active) {
        r = __ldg(reinterpret_cast<const float4*>(q.r) + i);
        e = __ldg(reinterpret_cast<const float4*>(q.emb) + i);
    }
    const float t0 = r.x - e.x, t1 = r.y - e.y, t2 = r.z - e.z, t3 = r.w - e.w;
    float dsq = fmaf(t0, t0, fmaf(t1, t1, fmaf(t2, t2, t3 * t3)));
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) dsq += __shfl_xor_sync(0xffffffffu, dsq, o);
    if (!active) return;
    if (q.res_out) reinterpret_cast<float4*>(q.res_out)[i] = r;
    if (q.r_next) reinterpret_cast<float4*>(q.r_next)[i] = make_float4(t0, t1, t2, t3);
    if (q.loss && (i % LPR) == 0) {
        const float lvl = dsq + q.beta * dsq;                                    // loss.py:40-42
        q.loss[row] = q.first ? lvl : q.loss[row] + lvl;
    }
}


// ---- backward row kernels ---------------------------------------------------------------------------------------------
struct DeParams {
    const float* x;       // [B, D] saved residual of the level
    const float* emb;     // [B, D] W C, or nullptr: emb = x - x_next (the saved residuals of two consecutive levels)
    const float* x_next;  // [B, D] residual of level l + 1 (used when emb is nullptr)
    const float* g_emb;   // [B, D] or nullptr
    const float* g_res;   // [B, D] or nullptr
    const float* g_loss;  // [B] or nullptr
    float* G;             // [B, D] in: d objective / d r_{l+1}; out: G + 2 beta (x - emb) g_loss + g_res
    float* de;            // [B, D] out
    float* t;             // [B] out: de . emb
    int64_t B;
    float beta;
};
template <int D>
__global__ void __launch_bounds__(256) gumbel_de_kernel(DeParams q) {
    constexpr int LPR = D / 4;
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int64_t row = i / LPR;
    const bool active = row < q.B;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f), e = x, ge = x, gr = x, G = x;
    float gl = 0.f;
    if (active) {
        x = __ldg(reinterpret_cast<const float4*>(q.x) + i);
        if (q.emb) e = __ldg(reinterpret_cast<const float4*>(q.emb) + i);
        else {
            const float4 n = __ldg(reinterpret_cast<const float4*>(q.x_next) + i);          // rqvae.py:129: r_next = r - emb
            e = make_float4(x.x - n.x, x.y - n.y, x.z - n.z, x.w - n.w);
        }
        if (q.g_emb) ge = __ldg(reinterpret_cast<const float4*>(q.g_emb) + i);
        if (q.g_res) gr = __ldg(reinterpret_cast<const float4*>(q.g_res) + i);
        G = reinterpret_cast<const float4*>(q.G)[i];
        if (q.g_loss) gl = __ldg(q.g_loss + row);
    }
    float4 de;
    de.x = (ge.x - G.x) + 2.f * (e.x - x.x) * gl; de.y = (ge.y - G.y) + 2.f * (e.y - x.y) * gl;
    de.z = (ge.z - G.z) + 2.f * (e.z - x.z) * gl; de.w = (ge.w - G.w) + 2.f * (e.w - x.w) * gl;
    float t = fmaf(de.x, e.x, fmaf(de.y, e.y, fmaf(de.z, e.z, de.w * e.w)));
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (!active) return;
    const float b2 = 2.f * q.beta * gl;
    G.x += b2 * (x.x - e.x) + gr.x; G.y += b2 * (x.y - e.y) + gr.y;
    G.z += b2 * (x.z - e.z) + gr.z; G.w += b2 * (x.w - e.w) + gr.w;
    reinterpret_cast<float4*>(q.G)[i] = G;
    reinterpret_cast<float4*>(q.de)[i] = de;
    if ((i % LPR) == 0) q.t[row] = t;
}

// dd = -W (dW - t) / T in place over dW, and colsum(dd) (one atomic per block and column).  Thread = four columns,
// a block walks 256 rows; the codebook-gradient GEMMs read W and dd as they lie (split_gemm_tn), so nothing is transposed.
__global__ void __launch_bounds__(256) gumbel_dd_kernel(const float* __restrict__ W, float* __restrict__ dW,
                                                        const float* __restrict__ t, int64_t B, int K, int Kp, float inv_t,
                                                        float* __restrict__ csum) {
    const int c = (blockIdx.y * 256 + threadIdx.x) * 4;
    if (c >= Kp) return;
    const int64_t r0 = (int64_t)blockIdx.x * 256;
    const int64_t r1 = r0 + 256 < B ? r0 + 256 : B;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    auto one = [&](int64_t r, const float4& w, const float4& d, float tr) {
        float4 o;
        o.x = (c + 0 < K) ? -(w.x * (d.x - tr)) * inv_t : 0.f;
        o.y = (c + 1 < K) ? -(w.y * (d.y - tr)) * inv_t : 0.f;
        o.z = (c + 2 < K) ? -(w.z * (d.z - tr)) * inv_t : 0.f;
        o.w = (c + 3 < K) ? -(w.w * (d.w - tr)) * inv_t : 0.f;
        *reinterpret_cast<float4*>(dW + r * Kp + c) = o;
        acc[0] += o.x; acc[1] += o.y; acc[2] += o.z; acc[3] += o.w;
    };
    int64_t r = r0;
    for (; r + 4 <= r1; r += 4) {
        float4 w[4], d[4];
        float tr[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            w[u] = __ldg(reinterpret_cast<const float4*>(W + (r + u) * Kp + c));
            d[u] = *reinterpret_cast<const float4*>(dW + (r + u) * Kp + c);
            tr[u] = __ldg(t + r + u);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) one(r + u, w[u], d[u], tr[u]);
    }
    for (; r < r1; ++r)
        one(r, __ldg(reinterpret_cast<const float4*>(W + r * Kp + c)), *reinterpret_cast<const float4*>(dW + r * Kp + c),
            __ldg(t + r));
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (c + i < K) atomicAdd(csum + c + i, acc[i]);
}

// g_C[k, :] += 2 C[k, :] * colsum(dd)[k]
__global__ void __launch_bounds__(256) gumbel_colsum_apply_kernel(const float* __restrict__ csum, const float* __restrict__ cb,
                                                                  float* __restrict__ g_cb, int K, int D) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= K * D) return;
    atomicAdd(g_cb + i, 2.f * __ldg(cb + i) * __ldg(csum + i / D));
}

// G <- G - 2 v (and g_x on the last level processed)
__global__ void __launch_bounds__(256) gumbel_g_update_kernel(float* __restrict__ G, const float* __restrict__ v, int64_t n4,
                                                              float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n4) return;
    float4 g = reinterpret_cast<float4*>(G)[i];
    const float4 w = __ldg(reinterpret_cast<const float4*>(v) + i);
    g.x -= 2.f * w.x; g.y -= 2.f * w.y; g.z -= 2.f * w.z; g.w -= 2.f * w.w;
    reinterpret_cast<float4*>(G)[i] = g;
    if (out) reinterpret_cast<float4*>(out)[i] = g;
}

}  // namespace gtc

bool gumbel_tc_supported(int D, int K, int L, int distance) {
    return (D == 32 || D == 64) && K >= 4 && L >= 1 && distance == RQ_DIST_L2;
}

template <int D>
static int gumbel_forward_tc_impl(const FwdParams& p, cudaStream_t st, float* weights_out) {
    using namespace gtc;
    const int64_t B = p.B;
    const int K = p.K, Kp = (K + 3) & ~3;
    const size_t bytes_S = (size_t)B * Kp * 4, bytes_row = (size_t)B * D * 4;
    const size_t bytes_ct = (size_t)D * Kp * 4, bytes_cc = (size_t)Kp * 4;
    // scratch: S | r ping | r pong | emb (when the caller wants none) | C^T | cc
    // (the [B, Kp] matrix of a level is the caller's weights buffer when there is one: the backward gets W for free)
    const size_t total = bytes_S + 2 * bytes_row + (p.emb ? 0 : bytes_row) + bytes_ct + bytes_cc + 256;
    uint8_t* scratch = (uint8_t*)engine_pool_alloc(total, st);
    if (!scratch) {
        set_error("rq_quantize_forward: out of device memory for the Gumbel-softmax scratch (%zu bytes)", total);
        return (int)cudaErrorMemoryAllocation;
    }
    float* S0 = (float*)scratch;
    float* rbuf[2] = {(float*)(scratch + bytes_S), (float*)(scratch + bytes_S + bytes_row)};
    uint8_t* cur = scratch + bytes_S + 2 * bytes_row;
    float* emb_scratch = nullptr;
    if (!p.emb) { emb_scratch = (float*)cur; cur += bytes_row; }
    float* ct = (float*)cur; cur += bytes_ct;
    float* cc = (float*)cur;
    int rc = 0;
    const float* r_in = p.x;
    for (int l = 0; l < p.L && rc == 0; ++l) {
        const float* cbl = p.cb + (int64_t)l * K * D;
        float* emb_l = p.emb ? p.emb + (int64_t)l * B * D : emb_scratch;
        float* S = weights_out ? weights_out + (int64_t)l * B * Kp : S0;
        {   // ids: the exact single-level distance kernel on this level's residual (quantize.py:121)
            FwdParams q;
            q.x = r_in; q.cb = cbl; q.B = B; q.K = K; q.L = 1; q.mode = RQ_MODE_EVAL; q.dist = RQ_DIST_L2;
            q.beta = p.beta; q.temp = 1.f; q.gu = nullptr; q.seed = 0; q.offset = 0;
            q.emb = nullptr; q.res = nullptr; q.loss = nullptr; q.KC = 0;
            q.ids = p.ids + (int64_t)l * B;
            const bool aligned = (((uintptr_t)cbl & 15) == 0) && (((uintptr_t)r_in & 31) == 0);
            if (aligned && (tc_supported(D, K, 1, RQ_MODE_EVAL, RQ_DIST_L2) || tc64_bigk_supported(D, K, 1, RQ_MODE_EVAL, RQ_DIST_L2)))
                rc = quantize_forward_tc(q, D, st);
            else
                rc = quantize_forward_simt(q, D, st);
            if (rc) break;
        }
        codebook_prep_kernel<D><<<(Kp + 255) / 256, 256, 0, st>>>(cbl, K, Kp, cc, ct);
        RQ_LAUNCHED();
        rc = split_gemm(r_in, cbl, S, B, K, D, D, D, Kp, 1.f, 0, 1, st);
        if (rc) break;
        RowParams q;
        q.S = S; q.r = r_in; q.cc = cc; q.gu = p.gu ? p.gu + (int64_t)l * B * K : nullptr;
        q.B = B; q.K = K; q.Kp = Kp; q.level = l; q.inv_t = 1.0f / p.temp; q.seed = p.seed; q.offset = p.offset;
        gumbel_softmax_rows_kernel<D><<<(unsigned)((B + 7) / 8), 256, 0, st>>>(q);
        RQ_LAUNCHED();
        rc = split_gemm(S, ct, emb_l, B, D, K, Kp, Kp, D, 1.f, 0, 1, st, 6);
        if (rc) break;
        TailParams t;
        t.r = r_in; t.emb = emb_l; t.res_out = p.res ? p.res + (int64_t)l * B * D : nullptr;
        t.r_next = (l + 1 < p.L) ? rbuf[l & 1] : nullptr;
        t.loss = p.loss; t.B = B; t.beta = p.beta; t.first = (l == 0);
        gumbel_level_tail_kernel<D><<<(unsigned)((B * (D / 4) + 255) / 256), 256, 0, st>>>(t);
        RQ_LAUNCHED();
        r_in = t.r_next;
    }
    cudaFreeAsync(scratch, st);
    return rc;
}

int quantize_forward_gumbel_tc(const FwdParams& p, int D, cudaStream_t st, float* weights_out) {
    if (p.B == 0) return 0;
    return D == 64 ? gumbel_forward_tc_impl<64>(p, st, weights_out) : gumbel_forward_tc_impl<32>(p, st, weights_out);
}

template <int D>
static int gumbel_backward_tc_impl(const BwdParams& p, float temperature, const float* gu, uint64_t seed, uint64_t offset,
                                   cudaStream_t st, const float* weights) {
    using namespace gtc;
    const int64_t B = p.B, Bp = (B + 3) & ~(int64_t)3;
    const int K = p.K, Kp = (K + 3) & ~3;
    const size_t bytes_S = (size_t)B * Kp * 4, bytes_row = (size_t)B * D * 4;
    const size_t bytes_ct = (size_t)D * Kp * 4, bytes_cc = (size_t)Kp * 4;
    // scratch: W | dW/dd | emb | de | v | G | C^T | cc | t | colsum
    const size_t total = 2 * bytes_S + 4 * bytes_row + bytes_ct + 2 * bytes_cc + (size_t)Bp * 4 + 256;
    uint8_t* scratch = (uint8_t*)engine_pool_alloc(total, st);
    if (!scratch) {
        set_error("rq_quantize_backward: out of device memory for the Gumbel-softmax scratch (%zu bytes)", total);
        return (int)cudaErrorMemoryAllocation;
    }
    uint8_t* cur = scratch;
    auto take = [&](size_t n) { float* q = (float*)cur; cur += n; return q; };
    float* Wbuf = take(bytes_S);
    float* dd = take(bytes_S);
    float* emb = take(bytes_row);
    float* de = take(bytes_row);
    float* v = take(bytes_row);
    float* G = take(bytes_row);
    float* ct = take(bytes_ct);
    float* cc = take(bytes_cc);
    float* t = take((size_t)Bp * 4);
    float* csum = take(bytes_cc);
    int rc = 0;
#define GTC_CUDA(call)                                                                   \
    do { if (rc == 0) { cudaError_t e_ = (call); if (e_ != cudaSuccess) { set_error("rq_quantize_backward: %s", cudaGetErrorString(e_)); rc = (int)e_; } } } while (0)
    GTC_CUDA(cudaMemsetAsync(G, 0, bytes_row, st));
    GTC_CUDA(cudaMemsetAsync(p.g_cb, 0, (size_t)p.L * K * D * sizeof(float), st));      // this mode overwrites (rq_engine.h)
    const unsigned rows8 = (unsigned)((B + 7) / 8), rowsq = (unsigned)((B * (D / 4) + 255) / 256);
    const int64_t n_kb = (B + 63) / 64;
    int splits = (int)((2 * (int64_t)sm_count() * 128 + K - 1) / K);          // ~2 CTAs per SM over the [K, D] output tiles
    if (splits < 1) splits = 1;
    if (splits > n_kb) splits = (int)n_kb;
    for (int l = p.L - 1; l >= 0 && rc == 0; --l) {
        const float* cbl = p.cb + (int64_t)l * K * D;
        const float* x = p.res + (int64_t)l * B * D;
        float* gcb = p.g_cb + (int64_t)l * K * D;
        codebook_prep_kernel<D><<<(Kp + 255) / 256, 256, 0, st>>>(cbl, K, Kp, cc, ct);
        RQ_LAUNCHED();
        // W: saved by the forward, or recomputed (logits product + softmax rows)
        const float* W = weights ? weights + (int64_t)l * B * Kp : Wbuf;
        if (!weights) {
            if ((rc = split_gemm(x, cbl, Wbuf, B, K, D, D, D, Kp, 1.f, 0, 1, st))) break;
            RowParams q;
            q.S = Wbuf; q.r = x; q.cc = cc; q.gu = gu ? gu + (int64_t)l * B * K : nullptr;
            q.B = B; q.K = K; q.Kp = Kp; q.level = l; q.inv_t = 1.0f / temperature; q.seed = seed; q.offset = offset;
            gumbel_softmax_rows_kernel<D><<<rows8, 256, 0, st>>>(q);
            RQ_LAUNCHED();
        }
        // emb = W C is needed for de and t: below the last level it is the difference of two saved residuals
        const bool have_next = (l + 1 < p.L);
        if (!have_next && (rc = split_gemm(W, ct, emb, B, D, K, Kp, Kp, D, 1.f, 0, 1, st))) break;
        // de, t, and the part of G that does not depend on the codes
        DeParams dq;
        dq.x = x; dq.emb = have_next ? nullptr : emb;
        dq.x_next = have_next ? p.res + (int64_t)(l + 1) * B * D : nullptr;
        dq.g_emb = p.g_emb ? p.g_emb + (int64_t)l * B * D : nullptr;
        dq.g_res = p.g_res ? p.g_res + (int64_t)l * B * D : nullptr;
        dq.g_loss = p.g_loss; dq.G = G; dq.de = de; dq.t = t; dq.B = B; dq.beta = p.beta;
        gumbel_de_kernel<D><<<rowsq, 256, 0, st>>>(dq);
        RQ_LAUNCHED();
        // dW = de C^T and dd = -W (dW - t) / T with its column sums: one launch when the B-resident GEMM applies (its
        // epilogue reads W and applies the softmax Jacobian), else the GEMM and a streaming pass over dW
        GTC_CUDA(cudaMemsetAsync(csum, 0, bytes_cc, st));
        if (rc) break;
        if (K > 64) {
            if ((rc = split_gemm_softmax_jacobian(de, cbl, dd, B, K, D, D, D, Kp, W, t, -1.0f / temperature, csum, st))) break;
        } else {
            if ((rc = split_gemm(de, cbl, dd, B, K, D, D, D, Kp, 1.f, 0, 1, st))) break;
            dim3 gT((unsigned)((B + 255) / 256), (unsigned)((Kp / 4 + 255) / 256));
            gumbel_dd_kernel<<<gT, 256, 0, st>>>(W, dd, t, B, K, Kp, 1.0f / temperature, csum);
            RQ_LAUNCHED();
        }
        // v = dd C
        if ((rc = split_gemm(dd, ct, v, B, D, K, Kp, Kp, D, 1.f, 0, 1, st, 3))) break;
        // codebook gradient: W^T de - 2 dd^T x + 2 C colsum(dd), contraction over the rows; W, dd, de and x are read as
        // they lie in memory (MN-major staging): no [K, B] copies
        if ((rc = split_gemm_tn(W, de, gcb, K, D, B, Kp, D, D, 1.f, 1, splits, st, 3))) break;
        if ((rc = split_gemm_tn(dd, x, gcb, K, D, B, Kp, D, D, -2.f, 1, splits, st, 3))) break;
        gumbel_colsum_apply_kernel<<<(K * D + 255) / 256, 256, 0, st>>>(csum, cbl, gcb, K, D);
        RQ_LAUNCHED();
        // G = G + ... - 2 v; the last level processed writes g_x
        gumbel_g_update_kernel<<<rowsq, 256, 0, st>>>(G, v, B * (D / 4), l == 0 ? p.g_x : nullptr);
        RQ_LAUNCHED();
    }
#undef GTC_CUDA
    cudaFreeAsync(scratch, st);
    return rc;
}

int quantize_backward_gumbel_tc(const BwdParams& p, int D, float temperature, const float* gu, uint64_t seed,
                                uint64_t offset, cudaStream_t st, const float* weights) {
    if (p.B == 0) {
        RQ_CUDA(cudaMemsetAsync(p.g_cb, 0, (size_t)p.L * p.K * D * sizeof(float), st));
        return 0;
    }
    return D == 64 ? gumbel_backward_tc_impl<64>(p, temperature, gu, seed, offset, st, weights)
                   : gumbel_backward_tc_impl<32>(p, temperature, gu, seed, offset, st, weights);
}

}  // namespace rq
