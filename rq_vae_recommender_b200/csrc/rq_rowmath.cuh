// Per-row register math shared by the SIMT and tcgen05 kernel families.
#pragma once
#include "rq_common.cuh"

namespace rq {

template <int D>
__device__ __forceinline__ void load_row(const float* __restrict__ src, float (&r)[D]) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
    for (int j = 0; j < D / 4; ++j) {
        float4 v = __ldg(s4 + j);
        r[4 * j] = v.x; r[4 * j + 1] = v.y; r[4 * j + 2] = v.z; r[4 * j + 3] = v.w;
    }
}
template <int D>
__device__ __forceinline__ void store_row(float* __restrict__ dst, const float (&r)[D]) {
    float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
    for (int j = 0; j < D / 4; ++j) d4[j] = make_float4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
}
template <int D>
__device__ __forceinline__ float dot_self(const float (&r)[D]) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < D; ++j) s = fmaf(r[j], r[j], s);
    return s;
}

// Canonical ||c||^2 of a codeword, bit-identical in both kernel families: sequential fmaf over
// 8-element chunks, chunk partials combined as a balanced pairwise tree ((p0+p1)+(p2+p3) for D=32).
template <int D>
__device__ __forceinline__ float code_sqnorm(const float* __restrict__ c) {
    constexpr int W = D >= 8 ? 8 : D;
    constexpr int NP = D / W;
    float p[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < W; j += 4) {
            float4 v = *reinterpret_cast<const float4*>(c + q * W + j);
            s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
        }
        p[q] = s;
    }
#pragma unroll
    for (int s = 1; s < NP; s *= 2) {
#pragma unroll
        for (int q = 0; q + s < NP; q += 2 * s) p[q] += p[q + s];
    }
    return p[0];
}

// The per-mode epilogue of one level, shared by the SIMT and tcgen05 families.
// In: residual r, selected codeword e (EVAL/STE/ROT) .  Out: emb_out (o), returns the
// per-row loss of this level.  Operation order follows modules/quantize.py:130-150.
template <int D>
__device__ __forceinline__ float level_epilogue(int mode, float beta, const float (&r)[D],
                                                const float (&e)[D], float (&o)[D]) {
    float dsq = 0.f;
#pragma unroll
    for (int j = 0; j < D; ++j) {
        float t = r[j] - e[j];
        dsq = fmaf(t, t, dsq);
    }
    if (mode == RQ_MODE_STE) {
#pragma unroll
        for (int j = 0; j < D; ++j) o[j] = r[j] + (e[j] - r[j]);           // quantize.py:132
    } else if (mode == RQ_MODE_ROTATION_TRICK) {
        // quantize.py:135-142 with efficient_rotation_trick_transform (:34-45)
        float nx = sqrtf(dot_self<D>(r)), ne = sqrtf(dot_self<D>(e));
        float iu = 1.0f / (nx + 1e-8f), iq = 1.0f / (ne + 1e-8f);
        float ns = 0.f, xw = 0.f, xu = 0.f;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float s = r[j] * iu + e[j] * iq;
            ns = fmaf(s, s, ns);
        }
        float iw = 1.0f / fmaxf(sqrtf(ns), 1e-6f);                         // F.normalize(eps=1e-6)
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float u = r[j] * iu, q = e[j] * iq, w = (u + q) * iw;
            xw = fmaf(r[j], w, xw);
            xu = fmaf(r[j], u, xu);
        }
        float lam = ne / (nx + 1e-6f);
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float u = r[j] * iu, q = e[j] * iq, w = (u + q) * iw;
            o[j] = ((r[j] - 2.f * (xw * w)) + 2.f * (xu * q)) * lam;
        }
    } else {                                                               // EVAL: quantize.py:149
#pragma unroll
        for (int j = 0; j < D; ++j) o[j] = e[j];
    }
    return dsq + beta * dsq;                                               // loss.py:40-42
}

}  // namespace rq
