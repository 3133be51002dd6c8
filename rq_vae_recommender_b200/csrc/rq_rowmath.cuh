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
// ---- canonical dot products ---------------------------------------------------------------------------------
// A row is cut into 8-element chunks.  A chunk's partial is one FFMA2 chain over its 4 element pairs (even and odd
// elements in the two lanes) followed by the sum of the two lanes; the chunk partials are combined by a balanced
// pairwise tree ((p0+p1)+(p2+p3) for D = 32).  A thread that owns a whole row and a group of D/8 lanes that own one
// chunk each (partials combined by shfl.xor 1, 2, 4) therefore produce the same bits — the SIMT family, the tcgen05
// families and the re-scoring paths all use these definitions.
constexpr int kRowChunk = 8;
template <int W>
__device__ __forceinline__ float chunk_dot(const float* a, const float* b) {
    f32x2 s = pack2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j + 1 < W; j += 2) s = fma2(pack2(a[j], a[j + 1]), pack2(b[j], b[j + 1]), s);
    return hsum2(s);
}
template <int W>
__device__ __forceinline__ float chunk_sqdiff(const float* a, const float* b) {      // sum (a - b)^2
    f32x2 s = pack2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j + 1 < W; j += 2) {
        const f32x2 t = sub2(pack2(a[j], a[j + 1]), pack2(b[j], b[j + 1]));
        s = fma2(t, t, s);
    }
    return hsum2(s);
}
template <int W>
__device__ __forceinline__ float chunk_sq_axpby(const float* a, const float* b, float ia, float ib) {   // sum (ia a + ib b)^2
    f32x2 s = pack2(0.f, 0.f);
    const f32x2 ia2 = pack2(ia, ia), ib2 = pack2(ib, ib);
#pragma unroll
    for (int j = 0; j + 1 < W; j += 2) {
        const f32x2 t = fma2(pack2(a[j], a[j + 1]), ia2, mul2(pack2(b[j], b[j + 1]), ib2));
        s = fma2(t, t, s);
    }
    return hsum2(s);
}
template <int NP>
__device__ __forceinline__ float tree_sum(float (&p)[NP]) {
#pragma unroll
    for (int s = 1; s < NP; s *= 2) {
#pragma unroll
        for (int q = 0; q + s < NP; q += 2 * s) p[q] += p[q + s];
    }
    return p[0];
}
template <int D>
__device__ __forceinline__ float dot_self(const float (&r)[D]) {
    constexpr int W = D >= kRowChunk ? kRowChunk : D;
    float p[D / W];
#pragma unroll
    for (int q = 0; q < D / W; ++q) p[q] = chunk_dot<W>(&r[q * W], &r[q * W]);
    return tree_sum(p);
}

// Scalars of the rotation trick (see level_epilogue) from the dot products xx = r.r, ee = e.e, re = r.e.
// rot_prepare returns |u+q|^2 by the expanded form; when it is < 1/16 the caller recomputes it directly
// (sum (iu r + iq e)^2, canonical order) before rot_finish.
// The part of the scalars that depends on the residual only (known before the codeword is): |r|, 1/(|r|+1e-8),
// 1/(|r|+1e-6).  The fused kernel computes it while the level's MMAs are in flight.
struct RotX {
    float nx, iu, il;
};
__device__ __forceinline__ RotX rot_x(float xx) {
    RotX X;
    X.nx = fast_sqrt(xx);
    X.iu = fast_rcp(X.nx + 1e-8f);
    X.il = fast_rcp(X.nx + 1e-6f);
    return X;
}
__device__ __forceinline__ float rot_prepare(const RotX& X, float xx, float ee, float re, float& ne, float& iq) {
    ne = fast_sqrt(ee);
    iq = fast_rcp(ne + 1e-8f);
    return fmaf(X.iu * X.iu, xx, fmaf(2.f * X.iu * iq, re, iq * iq * ee));
}
constexpr float kRotCancel = 0.0625f;                         // |u+q| < 1/4: recompute without cancellation
__device__ __forceinline__ void rot_finish(const RotX& X, float xx, float re, float ne, float iq, float ns, float& ca,
                                           float& ce) {
    const float iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f));   // F.normalize(eps=1e-6)
    const float rs = fmaf(X.iu, xx, iq * re);                 // r.(u+q)
    const float rsw = rs * iw * iw;                           // (r.w) * iw
    const float lam = ne * X.il;
    ca = lam * (1.f - 2.f * rsw * X.iu);
    ce = 2.f * lam * iq * (X.iu * xx - rsw);
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
    static_assert(D % 2 == 0, "rows are processed as fp32 pairs");
    constexpr int W = D >= kRowChunk ? kRowChunk : D;
    constexpr int NP = D / W;
    float pd[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) pd[q] = chunk_sqdiff<W>(&r[q * W], &e[q * W]);
    const float dsq = tree_sum(pd);
    if (mode == RQ_MODE_ROTATION_TRICK) {
        float pe[NP], pre[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            pe[q] = chunk_dot<W>(&e[q * W], &e[q * W]);
            pre[q] = chunk_dot<W>(&r[q * W], &e[q * W]);
        }
        const float ee = tree_sum(pe), re = tree_sum(pre);
        const RotX X = rot_x(xx);
        float ne, iq;
        float ns = rot_prepare(X, xx, ee, re, ne, iq);
        if (ns < kRotCancel) {
            float pn[NP];
#pragma unroll
            for (int q = 0; q < NP; ++q) pn[q] = chunk_sq_axpby<W>(&r[q * W], &e[q * W], X.iu, iq);
            ns = tree_sum(pn);
        }
        float ca, ce;
        rot_finish(X, xx, re, ne, iq, ns, ca, ce);
        const f32x2 ca2 = pack2(ca, ca), ce2 = pack2(ce, ce);
#pragma unroll
        for (int j = 0; j < D / 2; ++j)
            unpack2(fma2(ca2, pack2(r[2 * j], r[2 * j + 1]), mul2(ce2, pack2(e[2 * j], e[2 * j + 1]))), o[2 * j],
                    o[2 * j + 1]);
    } else if (mode == RQ_MODE_STE) {
#pragma unroll
        for (int j = 0; j < D / 2; ++j) {                                  // quantize.py:132: x + (emb - x)
            const f32x2 rp = pack2(r[2 * j], r[2 * j + 1]), ep = pack2(e[2 * j], e[2 * j + 1]);
            unpack2(add2(rp, sub2(ep, rp)), o[2 * j], o[2 * j + 1]);
        }
    } else {                                                               // EVAL: quantize.py:149
#pragma unroll
        for (int j = 0; j < D; ++j) o[j] = e[j];
    }
    return dsq + beta * dsq;                                               // loss.py:40-42
}

// r -= o (rqvae.py:129), pairwise
template <int D>
__device__ __forceinline__ void residual_update(float (&r)[D], const float (&o)[D]) {
#pragma unroll
    for (int j = 0; j + 1 < D; j += 2)
        unpack2(sub2(pack2(r[j], r[j + 1]), pack2(o[j], o[j + 1])), r[j], r[j + 1]);
}

}  // namespace rq
