// Gumbel-softmax backward of the fused L-level quantiser (closed forms of SURVEY.md §8 a11; the reference
// obtains them from autograd through rand -> log -> softmax -> two GEMMs per level, quantize.py:124-129).
//
// One thread owns one item row and walks the levels top-down carrying G = d objective / d residual_{l+1}.
// Per level the softmax is recomputed from the saved residual (nothing of size [B, K] is ever stored):
//   pass 1  y_k = (-dist_k + gumbel_k) / T, online max / sum and emb = softmax(y) @ C      (as the forward)
//           de  = g_emb - G + 2 (emb - x) g_loss                      d objective / d emb_out
//   pass 2  W_k = softmax(y)_k, dW_k = de . c_k, dy_k = W_k (dW_k - de . emb), dd_k = -dy_k / T
//           g_x    += 2 x sum_k dd_k - 2 sum_k dd_k c_k + 2 beta (x - emb) g_loss
//           g_C[k] += W_k de + 2 dd_k (c_k - x)
// The per-code contribution vectors of a warp's 32 rows are summed with a transposing butterfly (D - 1
// shuffles for D values) so that one lane issues one atomic per (warp, code, element) instead of 32.
#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"

namespace rq {

#ifndef RQ_GB_THREADS
#define RQ_GB_THREADS 128
#endif
constexpr int kGbThreads = RQ_GB_THREADS;

// sum over the 32 lanes of v[0..D): on return lane j (and j + 32, ...) holds element j's total in v[0], v[1]...
// i.e. element (lane + 32 * i) is in out[i].  D in {4, 8, 16, 32, 64}.
template <int D>
__device__ __forceinline__ void warp_transpose_sum(float (&v)[D], float (&out)[(D + 31) / 32]) {
    const int lane = threadIdx.x & 31;
    if (D >= 32) {
#pragma unroll
        for (int blk = 0; blk < D / 32; ++blk) {
            float* w = v + 32 * blk;
            // halve the live set each step: keep the half selected by the lane bit, send the other half
#pragma unroll
            for (int half = 16; half >= 1; half >>= 1) {
                const bool up = (lane & half) != 0;
#pragma unroll
                for (int j = 0; j < half; ++j) {
                    const float keep = up ? w[j + half] : w[j];
                    const float send = up ? w[j] : w[j + half];
                    w[j] = keep + __shfl_xor_sync(0xffffffffu, send, half);
                }
            }
            out[blk] = w[0];      // element index = lane (bit pattern of the choices) within this block
        }
    } else {
        // D < 32: reduce each element fully (few elements)
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float x = v[j];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
            v[j] = x;
        }
        out[0] = 0.f;
#pragma unroll
        for (int j = 0; j < D; ++j)
            if (lane == j) out[0] = v[j];
    }
}

struct GumbelBwdParams {
    BwdParams b;
    float temp;
    const float* gu;
    uint64_t seed, offset;
    int KC;
};

template <int D>
__global__ void __launch_bounds__(kGbThreads) rq_bwd_gumbel_kernel(GumbelBwdParams P) {
    const BwdParams& p = P.b;
    extern __shared__ __align__(16) float smem[];
    float* sC = smem;                  // [KC][D]
    float* sN = smem + P.KC * D;       // [KC] ||c||^2
    // [D][threads]: the running gradient G waits here during the two passes over the codes (x, de, v and G together
    // are 4 D registers — at D = 64 one more than a thread has)
    float* sG = sN + ((P.KC + 3) & ~3) + threadIdx.x;
    const int tid = threadIdx.x, lane = tid & 31;
    const int64_t row = (int64_t)blockIdx.x * kGbThreads + tid;
    const bool active = row < p.B;
    const float invT = 1.0f / P.temp;
    const float gl = (active && p.g_loss) ? __ldg(p.g_loss + row) : 0.f;

    auto stage_chunk = [&](const float* cbl, int k0, int kc) {
        __syncthreads();
        for (int i = tid; i < kc * (D / 4); i += kGbThreads)
            reinterpret_cast<float4*>(sC)[i] = __ldg(reinterpret_cast<const float4*>(cbl + (int64_t)k0 * D) + i);
        __syncthreads();
        for (int k = tid; k < kc; k += kGbThreads) sN[k] = code_sqnorm<D>(sC + k * D);
        __syncthreads();
    };
    auto noise = [&](int l, int kk) -> float {              // the forward's draw for (row, level l, code kk)
        if (!active) return 0.5f;
        if (P.gu) return __ldg(P.gu + ((int64_t)l * p.B + row) * p.K + kk);
        uint4 rnd = philox4x32_10(make_uint4((uint32_t)row, (uint32_t)(row >> 32), (uint32_t)(kk >> 2), (uint32_t)l),
                                  make_uint2((uint32_t)(P.seed ^ P.offset), (uint32_t)((P.seed >> 32) ^ (P.offset >> 32))));
        const int q = kk & 3;
        return u01(q == 0 ? rnd.x : q == 1 ? rnd.y : q == 2 ? rnd.z : rnd.w);
    };

    float G[D];
#pragma unroll
    for (int j = 0; j < D; ++j) G[j] = 0.f;

    for (int l = p.L - 1; l >= 0; --l) {
        const float* cbl = p.cb + (int64_t)l * p.K * D;
        const int64_t off = ((int64_t)l * p.B + row) * D;
        float x[D], de[D];
        if (active) load_row<D>(p.res + off, x);
        else {
#pragma unroll
            for (int j = 0; j < D; ++j) x[j] = 0.f;
        }
        const float xx = dot_self<D>(x);
        // ---- pass 1: softmax statistics and emb --------------------------------------------------------
        float m = -INFINITY, s = 0.f;
        float emb[D];
#pragma unroll
        for (int j = 0; j < D; ++j) emb[j] = 0.f;
        for (int k0 = 0; k0 < p.K; k0 += P.KC) {
            const int kc = min(P.KC, p.K - k0);
            stage_chunk(cbl, k0, kc);
            for (int k = 0; k < kc; ++k) {
                const float* c = sC + k * D;
                float a = 0.f;
#pragma unroll
                for (int j = 0; j < D; ++j) a = fmaf(x[j], c[j], a);
                const float d = (xx + sN[k]) - 2.f * a;
                const float y = (-d + gumbel_from_u(noise(l, k0 + k))) * invT;
                const float mn = fmaxf(m, y);
                const float sc = __expf(m - mn), w = __expf(y - mn);
                s = s * sc + w;
#pragma unroll
                for (int j = 0; j < D; ++j) emb[j] = fmaf(w, c[j], emb[j] * sc);
                m = mn;
            }
        }
        const float inv_s = 1.0f / s;
        float t = 0.f;
        // emb is retired here (its share of d objective / d residual_l goes into G right away), so that pass 2 holds
        // x, de, v and G only: at D = 64 the extra 64 live registers spilled 1.7 KB per thread
#pragma unroll
        for (int j = 0; j < D; ++j) {
            const float ej = emb[j] * inv_s;
            const float ge = (active && p.g_emb) ? __ldg(p.g_emb + off + j) : 0.f;
            const float tg = (active && p.g_res) ? __ldg(p.g_res + off + j) : 0.f;
            de[j] = (ge - G[j]) + 2.f * (ej - x[j]) * gl;
            t = fmaf(de[j], ej, t);
            sG[j * kGbThreads] = G[j] + (2.f * p.beta * (x[j] - ej) * gl + tg);
        }
        // ---- pass 2: gradients ------------------------------------------------------------------------------
        float sdd = 0.f, v[D];
#pragma unroll
        for (int j = 0; j < D; ++j) v[j] = 0.f;
        for (int k0 = 0; k0 < p.K; k0 += P.KC) {
            const int kc = min(P.KC, p.K - k0);
            if (p.K > P.KC) stage_chunk(cbl, k0, kc);        // single chunk: still resident from pass 1
            for (int k = 0; k < kc; ++k) {
                const float* c = sC + k * D;
                float a = 0.f, dW = 0.f;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    a = fmaf(x[j], c[j], a);
                    dW = fmaf(de[j], c[j], dW);
                }
                const float d = (xx + sN[k]) - 2.f * a;
                const float y = (-d + gumbel_from_u(noise(l, k0 + k))) * invT;
                const float W = __expf(y - m) * inv_s;
                const float dd = -(W * (dW - t)) * invT;
                sdd += dd;
                float* dst = p.g_cb + ((int64_t)l * p.K + k0 + k) * D;
                constexpr int BLK = D >= 32 ? 32 : D;          // contributions are built and reduced 32 elements at a time
#pragma unroll
                for (int b = 0; b < D / BLK; ++b) {
                    float contrib[BLK];
#pragma unroll
                    for (int j = 0; j < BLK; ++j) {
                        const int jj = b * BLK + j;
                        v[jj] = fmaf(dd, c[jj], v[jj]);
                        contrib[j] = fmaf(W, de[jj], 2.f * dd * (c[jj] - x[jj]));
                    }
                    float red[1];
                    warp_transpose_sum<BLK>(contrib, red);
                    if (D >= 32) atomicAdd(dst + 32 * b + lane, red[0]);
                    else if (lane < D) atomicAdd(dst + lane, red[0]);
                }
            }
        }
        // ---- d objective / d residual_l ------------------------------------------------------------------------
#pragma unroll
        for (int j = 0; j < D; ++j) G[j] = sG[j * kGbThreads] + (2.f * x[j] * sdd - 2.f * v[j]);
    }
    if (active) store_row<D>(p.g_x + row * D, G);
}

template <int D>
static int launch_gumbel_bwd(const BwdParams& p, float temperature, const float* gu, uint64_t seed, uint64_t offset,
                             cudaStream_t st) {
    GumbelBwdParams P;
    P.b = p; P.temp = temperature; P.gu = gu; P.seed = seed; P.offset = offset;
    int kc = 16384 / D;
    if (kc > p.K) kc = p.K;
    P.KC = kc;
    const size_t smem = (size_t)kc * D * sizeof(float) + (size_t)((kc + 3) & ~3) * sizeof(float) +
                        (size_t)D * kGbThreads * sizeof(float);
    auto kern = rq_bwd_gumbel_kernel<D>;
    RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
    const unsigned grid = (unsigned)((p.B + kGbThreads - 1) / kGbThreads);
    kern<<<grid, kGbThreads, smem, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

int quantize_backward_gumbel(const BwdParams& p, int D, float temperature, const float* gu, uint64_t seed,
                             uint64_t offset, cudaStream_t st) {
    RQ_CUDA(cudaMemsetAsync(p.g_cb, 0, (size_t)p.L * p.K * D * sizeof(float), st));
    if (p.B == 0) return 0;
    switch (D) {
        case 4: return launch_gumbel_bwd<4>(p, temperature, gu, seed, offset, st);
        case 8: return launch_gumbel_bwd<8>(p, temperature, gu, seed, offset, st);
        case 16: return launch_gumbel_bwd<16>(p, temperature, gu, seed, offset, st);
        case 32: return launch_gumbel_bwd<32>(p, temperature, gu, seed, offset, st);
        case 64: return launch_gumbel_bwd<64>(p, temperature, gu, seed, offset, st);
        default:
            set_error("GUMBEL_SOFTMAX backward supports D in {4, 8, 16, 32, 64} (got %d)", D);
            return RQ_ERR_UNSUPPORTED_SHAPE;
    }
}

}  // namespace rq
