// Exact fp32 CUDA-core kernel for the fused L-level residual quantiser (forward).
//
// One thread owns one item row for all L levels: the residual never leaves registers
// between levels (modules/rqvae.py:124-131 in the reference round-trips it through HBM
// 2L times).  Codebooks are staged chunk-wise in shared memory and read with warp-broadcast
// LDS.128.  Distances use the reference's expanded form, (||x||^2 + ||c||^2) - (2x).c
// (modules/quantize.py:108-112), ties go to the lowest index (:121).
//
// This family serves every (D, K, L, mode, distance) and is the exactness yardstick for the
// tcgen05 family in rq_quantize_tc.cu.
#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"

namespace rq {

constexpr int kSimtThreads = 256;

template <int D, bool GUMBEL>
__global__ void __launch_bounds__(kSimtThreads) rq_fwd_simt_kernel(FwdParams p) {
    extern __shared__ __align__(16) float smem[];
    float* sC = smem;                 // [KC][D]
    float* sN = smem + p.KC * D;      // [KC]  ||c||^2 (L2) or ||c|| (COSINE)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t row = p.row0 + (int64_t)blockIdx.x * kSimtThreads + tid;
    const bool active = row < p.B;

    float r[D];
    if (active) load_row<D>(p.x + row * D, r);
    else {
#pragma unroll
        for (int j = 0; j < D; ++j) r[j] = 0.f;
    }
    float loss_acc = 0.f;

    for (int l = 0; l < p.L; ++l) {
        const float* cbl = p.cb + (int64_t)l * p.K * D;
        float xx = dot_self<D>(r);
        float xin = 0.f;  // 1/||x|| for COSINE
        if (p.dist == RQ_DIST_COSINE) xin = sqrtf(xx);
        float best = INFINITY;
        int bi = 0;
        // Gumbel-softmax running state (online softmax over K)
        float gm = -INFINITY, gs = 0.f;
        float acc[GUMBEL ? D : 1];
        if (GUMBEL) {
#pragma unroll
            for (int j = 0; j < (GUMBEL ? D : 1); ++j) acc[j] = 0.f;
        }

        for (int k0 = 0; k0 < p.K; k0 += p.KC) {
            const int kc = min(p.KC, p.K - k0);
            const int kc4 = (kc + 3) & ~3;
            __syncthreads();  // previous chunk fully consumed
            for (int i = tid; i < kc4 * (D / 4); i += kSimtThreads) {
                int k = i / (D / 4);
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k < kc) v = __ldg(reinterpret_cast<const float4*>(cbl + (int64_t)(k0 + k) * D) + (i % (D / 4)));
                reinterpret_cast<float4*>(sC)[i] = v;
            }
            __syncthreads();
            for (int k = tid; k < kc4; k += kSimtThreads) {
                float s = code_sqnorm<D>(sC + k * D);
                sN[k] = (p.dist == RQ_DIST_COSINE) ? sqrtf(s) : s;
            }
            __syncthreads();

            for (int k = 0; k < kc4; k += 4) {
                float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                const float4* c0 = reinterpret_cast<const float4*>(sC + (k + 0) * D);
                const float4* c1 = reinterpret_cast<const float4*>(sC + (k + 1) * D);
                const float4* c2 = reinterpret_cast<const float4*>(sC + (k + 2) * D);
                const float4* c3 = reinterpret_cast<const float4*>(sC + (k + 3) * D);
                if (p.dist == RQ_DIST_L2) {
#pragma unroll
                    for (int j = 0; j < D / 4; ++j) {
                        float4 v0 = c0[j], v1 = c1[j], v2 = c2[j], v3 = c3[j];
                        a0 = fmaf(r[4*j], v0.x, a0); a0 = fmaf(r[4*j+1], v0.y, a0); a0 = fmaf(r[4*j+2], v0.z, a0); a0 = fmaf(r[4*j+3], v0.w, a0);
                        a1 = fmaf(r[4*j], v1.x, a1); a1 = fmaf(r[4*j+1], v1.y, a1); a1 = fmaf(r[4*j+2], v1.z, a1); a1 = fmaf(r[4*j+3], v1.w, a1);
                        a2 = fmaf(r[4*j], v2.x, a2); a2 = fmaf(r[4*j+1], v2.y, a2); a2 = fmaf(r[4*j+2], v2.z, a2); a2 = fmaf(r[4*j+3], v2.w, a2);
                        a3 = fmaf(r[4*j], v3.x, a3); a3 = fmaf(r[4*j+1], v3.y, a3); a3 = fmaf(r[4*j+2], v3.z, a3); a3 = fmaf(r[4*j+3], v3.w, a3);
                    }
                } else {
                    // quantize.py:115: (x / ||x||) @ C^T — the division is elementwise on x
#pragma unroll
                    for (int j = 0; j < D / 4; ++j) {
                        float4 v0 = c0[j], v1 = c1[j], v2 = c2[j], v3 = c3[j];
                        float x0 = r[4*j] / xin, x1 = r[4*j+1] / xin, x2 = r[4*j+2] / xin, x3 = r[4*j+3] / xin;
                        a0 = fmaf(x0, v0.x, a0); a0 = fmaf(x1, v0.y, a0); a0 = fmaf(x2, v0.z, a0); a0 = fmaf(x3, v0.w, a0);
                        a1 = fmaf(x0, v1.x, a1); a1 = fmaf(x1, v1.y, a1); a1 = fmaf(x2, v1.z, a1); a1 = fmaf(x3, v1.w, a1);
                        a2 = fmaf(x0, v2.x, a2); a2 = fmaf(x1, v2.y, a2); a2 = fmaf(x2, v2.z, a2); a2 = fmaf(x3, v2.w, a2);
                        a3 = fmaf(x0, v3.x, a3); a3 = fmaf(x1, v3.y, a3); a3 = fmaf(x2, v3.z, a3); a3 = fmaf(x3, v3.w, a3);
                    }
                }
                float d[4];
                if (p.dist == RQ_DIST_L2) {
                    d[0] = (xx + sN[k]) - 2.f * a0;  d[1] = (xx + sN[k + 1]) - 2.f * a1;
                    d[2] = (xx + sN[k + 2]) - 2.f * a2;  d[3] = (xx + sN[k + 3]) - 2.f * a3;
                } else {
                    d[0] = -(a0 / sN[k]);  d[1] = -(a1 / sN[k + 1]);  d[2] = -(a2 / sN[k + 2]);  d[3] = -(a3 / sN[k + 3]);
                }
                float u4[4] = {0.f, 0.f, 0.f, 0.f};
                if (GUMBEL && active) {
                    if (p.gu) {
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            if (k + q < kc) u4[q] = __ldg(p.gu + ((int64_t)l * p.B + row) * p.K + k0 + k + q);
                    } else {
                        uint4 rnd = philox4x32_10(
                            make_uint4((uint32_t)row, (uint32_t)(row >> 32), (uint32_t)((k0 + k) >> 2), (uint32_t)l),
                            make_uint2((uint32_t)(p.seed ^ p.offset), (uint32_t)((p.seed >> 32) ^ (p.offset >> 32))));
                        u4[0] = u01(rnd.x); u4[1] = u01(rnd.y); u4[2] = u01(rnd.z); u4[3] = u01(rnd.w);
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (k + q >= kc) continue;
                    if (d[q] < best) { best = d[q]; bi = k0 + k + q; }
                    if (GUMBEL) {
                        // softmax((-dist + g) / T) @ C, online over k (quantize.py:125-128)
                        float y = (-d[q] + gumbel_from_u(u4[q])) / p.temp;
                        float mn = fmaxf(gm, y);
                        float sc = __expf(gm - mn), w = __expf(y - mn);
                        gs = gs * sc + w;
                        const float* cq = sC + (k + q) * D;
#pragma unroll
                        for (int j = 0; j < (GUMBEL ? D : 1); ++j) acc[j] = fmaf(w, cq[j], acc[j] * sc);
                        gm = mn;
                    }
                }
            }
        }

        if (active) {
            float e[D], o[D];
            float lvl;
            if (GUMBEL) {
                float inv = 1.0f / gs;
                float dsq = 0.f;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    e[j] = acc[GUMBEL ? j : 0] * inv;
                    o[j] = e[j];
                    float t = r[j] - e[j];
                    dsq = fmaf(t, t, dsq);
                }
                lvl = dsq + p.beta * dsq;
            } else {
                load_row<D>(cbl + (int64_t)bi * D, e);
                lvl = level_epilogue<D>(p.mode, p.beta, r, e, xx, o);
            }
            loss_acc += lvl;
            const int64_t off = ((int64_t)l * p.B + row) * D;
            if (p.res) store_row<D>(p.res + off, r);
            if (p.emb) store_row<D>(p.emb + off, o);
            int64_t* idp = p.ids + (int64_t)l * p.B + row;
            if (p.n_changed && (!p.has_prev || *idp != (int64_t)bi)) atomicAdd(p.n_changed, 1ull);   // k-means tail rows
            *idp = bi;
            residual_update<D>(r, o);                                      // rqvae.py:129
        }
    }
    if (active && p.loss) p.loss[row] = loss_acc;
}

// ---------------------------------------------------------------------------------------------
template <int D>
static int launch_fwd(const FwdParams& p0, cudaStream_t st) {
    FwdParams p = p0;
    // chunk of at most 64 KB of codewords, multiple of 4 codes
    int kc = (16384 / D);
    if (kc > ((p.K + 3) & ~3)) kc = (p.K + 3) & ~3;
    p.KC = kc;
    size_t smem = (size_t)kc * D * sizeof(float) + (size_t)kc * sizeof(float);
    dim3 grid((unsigned)((p.B - p.row0 + kSimtThreads - 1) / kSimtThreads));
    if (p.mode == RQ_MODE_GUMBEL_SOFTMAX) {
        if (D > 64) { set_error("GUMBEL_SOFTMAX supports D <= 64 in the SIMT family (got %d)", D); return RQ_ERR_UNSUPPORTED_SHAPE; }
        auto kern = rq_fwd_simt_kernel<(D > 64 ? 64 : D), true>;
        RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
        kern<<<grid, kSimtThreads, smem, st>>>(p);
    } else {
        auto kern = rq_fwd_simt_kernel<D, false>;
        RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
        kern<<<grid, kSimtThreads, smem, st>>>(p);
    }
    RQ_LAUNCHED();
    return 0;
}

int quantize_forward_simt(const FwdParams& p, int D, cudaStream_t st) {
    if (p.B - p.row0 <= 0) return 0;
    switch (D) {
        case 4: return launch_fwd<4>(p, st);
        case 8: return launch_fwd<8>(p, st);
        case 16: return launch_fwd<16>(p, st);
        case 32: return launch_fwd<32>(p, st);
        case 64: return launch_fwd<64>(p, st);
        case 128: return launch_fwd<128>(p, st);
        default:
            set_error("unsupported embed_dim D=%d (supported: 4, 8, 16, 32, 64, 128)", D);
            return RQ_ERR_UNSUPPORTED_SHAPE;
    }
}

}  // namespace rq
