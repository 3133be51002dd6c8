// Fused backward of the L-level residual quantiser for EVAL / STE / ROTATION_TRICK
// (closed forms of SURVEY.md §8 a11; the reference gets these from a ~45-node-per-level
// autograd graph plus an index_add-style scatter for the embedding gradient).
//
// HBM-streaming kernel: D/4 adjacent lanes share one item row (one float4 each), so every warp
// load/store instruction covers whole rows — 512 contiguous bytes for the saved residuals, the
// upstream gradients and g_x, and one 128-byte line per gathered codeword.  Per-row reductions
// (norms, dot products of the rotation trick) are log2(D/4) shuffle steps inside the lane group.
// A lane group walks the levels in reverse, carrying d objective / d residual in registers;
// codebook gradients leave as 16-byte L2 reductions (REDG.ADD.F32x4).
//
// Algorithmic traffic per item: L*(4D saved residual + 4D upstream gradient + 8 id) + 4 + 4D =
// 924 B at D=32, L=3 — the same as the forward.
#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {

constexpr int kBwdThreads = 256;

template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
    return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w)));
}
__device__ __forceinline__ float4 ldg_nc4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
// streaming (read-once) variant: bypass L1 allocation
__device__ __forceinline__ float4 ldg_stream4(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}

constexpr int kBwdMaxL = 4;          // levels whose loads are hoisted (all of them in flight at once)

// One level of the reverse walk.  G = d objective / d residual_{l+1} on entry, d/d residual_l on exit.
template <int LPR>
__device__ __forceinline__ void bwd_level(int mode, float beta, float gl, const float4& r, const float4& go_in,
                                          const float4& t, const float4& e, float4& G, float4& gc) {
    // res_{l+1} = r_l - out_l  =>  out_l receives -G, r_l receives +G
    const float4 go = make_float4(go_in.x - G.x, go_in.y - G.y, go_in.z - G.z, go_in.w - G.w);
    gc = make_float4(2.f * (e.x - r.x) * gl, 2.f * (e.y - r.y) * gl, 2.f * (e.z - r.z) * gl,
                     2.f * (e.w - r.w) * gl);                                  // d emb_loss / d e
    if (mode == RQ_MODE_STE) {
        G.x += go.x; G.y += go.y; G.z += go.z; G.w += go.w;                    // d out / d r = I
    } else if (mode == RQ_MODE_ROTATION_TRICK) {
        const float nx2 = group_sum<LPR>(dot4(r, r)), ne2 = group_sum<LPR>(dot4(e, e));
        const float nx = sqrtf(nx2), ne = sqrtf(ne2);
        const float iu = 1.0f / (nx + 1e-8f), iq = 1.0f / (ne + 1e-8f);
        const float lam = ne / (nx + 1e-6f);
        const float4 u = make_float4(r.x * iu, r.y * iu, r.z * iu, r.w * iu);
        const float4 q = make_float4(e.x * iq, e.y * iq, e.z * iq, e.w * iq);
        float4 w = make_float4(u.x + q.x, u.y + q.y, u.z + q.z, u.w + q.w);
        const float iw = 1.0f / fmaxf(sqrtf(group_sum<LPR>(dot4(w, w))), 1e-6f);
        w.x *= iw; w.y *= iw; w.z *= iw; w.w *= iw;
        const float4 gs = make_float4(lam * go.x, lam * go.y, lam * go.z, lam * go.w);
        const float gw = group_sum<LPR>(dot4(gs, w)), gq = group_sum<LPR>(dot4(gs, q));
        G.x += (gs.x - 2.f * gw * w.x) + 2.f * gq * u.x;
        G.y += (gs.y - 2.f * gw * w.y) + 2.f * gq * u.y;
        G.z += (gs.z - 2.f * gw * w.z) + 2.f * gq * u.z;
        G.w += (gs.w - 2.f * gw * w.w) + 2.f * gq * u.w;
    } else {                                                                   // EVAL: out = C[ids]
        gc.x += go.x; gc.y += go.y; gc.z += go.z; gc.w += go.w;
    }
    const float b2 = 2.f * beta * gl;                                          // d query_loss / d r
    G.x += b2 * (r.x - e.x) + t.x; G.y += b2 * (r.y - e.y) + t.y;
    G.z += b2 * (r.z - e.z) + t.z; G.w += b2 * (r.w - e.w) + t.w;
}

// HOIST = true (L <= kBwdMaxL): the ids of every level are loaded first, then every level's
// residual / upstream-gradient / codeword loads are issued before any arithmetic, so one lane
// group has 3L+ independent 16-byte loads in flight — the kernel is pure latency otherwise
// (ids -> codeword gather -> math, L times in series).
template <int D, bool HOIST>
__global__ void __launch_bounds__(kBwdThreads) rq_bwd_stream_kernel(BwdParams p, int64_t n_groups) {
    constexpr int LPR = D / 4;          // lanes per row
    constexpr int RPW = 32 / LPR;       // rows per warp instruction
    const int lane = threadIdx.x & 31;
    const int sub = lane % LPR, rsel = lane / LPR;
    const int64_t warp_global = ((int64_t)blockIdx.x * kBwdThreads + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * kBwdThreads) >> 5;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);

    for (int64_t g = warp_global; g < n_groups; g += n_warps) {
        const int64_t row = g * RPW + rsel;
        const bool valid = row < p.B;
        const int64_t rowc = valid ? row : (p.B - 1);      // clamp: tail lanes recompute the last row, never store
        const float gl = p.g_loss ? __ldg(p.g_loss + rowc) : 0.f;
        float4 G = z4;                                     // d objective / d residual_{l+1}
        if (HOIST) {
            int64_t id[kBwdMaxL];
            float4 r[kBwdMaxL], go[kBwdMaxL], t[kBwdMaxL], e[kBwdMaxL];
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l)
                if (l < p.L) id[l] = __ldg(p.ids + (int64_t)l * p.B + rowc);
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l) {
                if (l < p.L) {
                    const int64_t off = ((int64_t)l * p.B + rowc) * D + sub * 4;
                    r[l] = ldg_stream4(p.res + off);
                    go[l] = p.g_emb ? ldg_stream4(p.g_emb + off) : z4;
                    t[l] = p.g_res ? ldg_stream4(p.g_res + off) : z4;
                }
            }
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l)
                if (l < p.L) e[l] = ldg_nc4(p.cb + ((int64_t)l * p.K + id[l]) * D + sub * 4);
#pragma unroll
            for (int l = kBwdMaxL - 1; l >= 0; --l) {
                if (l < p.L) {
                    float4 gc;
                    bwd_level<LPR>(p.mode, p.beta, gl, r[l], go[l], t[l], e[l], G, gc);
                    if (valid) red_add_v4(p.g_cb + ((int64_t)l * p.K + id[l]) * D + sub * 4, gc.x, gc.y, gc.z, gc.w);
                }
            }
        } else {
            for (int l = p.L - 1; l >= 0; --l) {
                const int64_t off = ((int64_t)l * p.B + rowc) * D + sub * 4;
                const int64_t id = __ldg(p.ids + (int64_t)l * p.B + rowc);
                const float4 r = ldg_stream4(p.res + off);
                const float4 go_in = p.g_emb ? ldg_stream4(p.g_emb + off) : z4;
                const float4 t = p.g_res ? ldg_stream4(p.g_res + off) : z4;
                const int64_t coff = ((int64_t)l * p.K + id) * D + sub * 4;
                const float4 e = ldg_nc4(p.cb + coff);
                float4 gc;
                bwd_level<LPR>(p.mode, p.beta, gl, r, go_in, t, e, G, gc);
                if (valid) red_add_v4(p.g_cb + coff, gc.x, gc.y, gc.z, gc.w);
            }
        }
        if (valid) *reinterpret_cast<float4*>(p.g_x + row * D + sub * 4) = G;
    }
}

template <int D>
static int launch_bwd(const BwdParams& p, cudaStream_t st) {
    constexpr int RPW = 32 / (D / 4);
    const int64_t n_groups = (p.B + RPW - 1) / RPW;
    const int warps_per_block = kBwdThreads / 32;
    int64_t blocks = (n_groups + warps_per_block - 1) / warps_per_block;
    const int64_t cap = (int64_t)sm_count() * 32;          // a few waves; each warp walks <= a handful of row groups
    if (blocks > cap) blocks = cap;
    if (p.L <= kBwdMaxL) rq_bwd_stream_kernel<D, true><<<(unsigned)blocks, kBwdThreads, 0, st>>>(p, n_groups);
    else rq_bwd_stream_kernel<D, false><<<(unsigned)blocks, kBwdThreads, 0, st>>>(p, n_groups);
    RQ_LAUNCHED();
    return 0;
}

int quantize_backward_simt(const BwdParams& p, int D, cudaStream_t st) {
    RQ_CUDA(cudaMemsetAsync(p.g_cb, 0, (size_t)p.L * p.K * D * sizeof(float), st));
    if (p.B == 0) return 0;
    switch (D) {
        case 4: return launch_bwd<4>(p, st);
        case 8: return launch_bwd<8>(p, st);
        case 16: return launch_bwd<16>(p, st);
        case 32: return launch_bwd<32>(p, st);
        case 64: return launch_bwd<64>(p, st);
        case 128: return launch_bwd<128>(p, st);
        default:
            set_error("unsupported embed_dim D=%d (supported: 4, 8, 16, 32, 64, 128)", D);
            return RQ_ERR_UNSUPPORTED_SHAPE;
    }
}

}  // namespace rq
