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

// The per-mode epilogue of one level, shared by the SIMT and tcgen05 families (bit-identical in both).
// In: residual r, its squared norm xx = dot_self(r), selected codeword e.  Out: emb_out (o); returns the
// per-row loss of this level (loss.py:40-42).
//
// ROTATION_TRICK (quantize.py:135-142 with efficient_rotation_trick_transform :34-45):
//   u = r/(|r|+1e-8), q = e/(|e|+1e-8), w = normalize(u+q, eps=1e-6), lam = |e|/(|r|+1e-6)
//   o = lam * (r - 2 (r.w) w + 2 (r.u) q)
// Every term is a multiple of r or e, so o = ca*r + ce*e with scalars built from three dot products
// (r.r, e.e, r.e): with s = u+q, |s|^2 = iu^2 r.r + 2 iu iq r.e + iq^2 e.e and r.s = iu r.r + iq r.e,
//   ca = lam (1 - 2 (r.s) iw^2 iu),   ce = 2 lam iq (iu r.r - (r.s) iw^2),   iw = 1/max(|s|, 1e-6).
// Two passes of 4 flops per element instead of the reference's ~12.  The expansion of |s|^2 cancels when
// u ~ -q (codeword antipodal to the residual); that case falls back to the direct sum.
template <int D>
__device__ __forceinline__ float level_epilogue(int mode, float beta, const float (&r)[D], const float (&e)[D],
                                                float xx, float (&o)[D]) {
    float dsq = 0.f;
    if (mode == RQ_MODE_ROTATION_TRICK) {
        float ee = 0.f, re = 0.f;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float t = r[j] - e[j];
            dsq = fmaf(t, t, dsq);
            ee = fmaf(e[j], e[j], ee);
            re = fmaf(r[j], e[j], re);
        }
        const float nx = fast_sqrt(xx), ne = fast_sqrt(ee);
        const float iu = fast_rcp(nx + 1e-8f), iq = fast_rcp(ne + 1e-8f);
        float ns = fmaf(iu * iu, xx, fmaf(2.f * iu * iq, re, iq * iq * ee));
        if (ns < 0.0625f) {                                   // |u+q| < 1/4: recompute without cancellation
            ns = 0.f;
#pragma unroll
            for (int j = 0; j < D; ++j) {
                float sj = fmaf(r[j], iu, e[j] * iq);
                ns = fmaf(sj, sj, ns);
            }
        }
        const float iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f)); // F.normalize(eps=1e-6)
        const float rs = fmaf(iu, xx, iq * re);               // r.(u+q)
        const float rsw = rs * iw * iw;                       // (r.w) * iw
        const float lam = ne * fast_rcp(nx + 1e-6f);
        const float ca = lam * (1.f - 2.f * rsw * iu);
        const float ce = 2.f * lam * iq * (iu * xx - rsw);
#pragma unroll
        for (int j = 0; j < D; ++j) o[j] = fmaf(ca, r[j], ce * e[j]);
    } else {
#pragma unroll
        for (int j = 0; j < D; ++j) {
            float t = r[j] - e[j];
            dsq = fmaf(t, t, dsq);
        }
        if (mode == RQ_MODE_STE) {
#pragma unroll
            for (int j = 0; j < D; ++j) o[j] = r[j] + (e[j] - r[j]);       // quantize.py:132
        } else {                                                           // EVAL: quantize.py:149
#pragma unroll
            for (int j = 0; j < D; ++j) o[j] = e[j];
        }
    }
    return dsq + beta * dsq;                                               // loss.py:40-42
}

}  // namespace rq
