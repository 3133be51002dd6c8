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
#include <stdlib.h>

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


// ---------------------------------------------------------------------------------------------------------
// TMA-fed persistent variant (the default): one CTA per SM walks 64-row tiles; the saved residuals and the
// upstream gradients of ALL levels of a tile arrive by cp.async.bulk (rows of a tile are contiguous in the
// [L, B, D] layout, one bulk copy per level and array) into a 2-stage shared-memory ring, so ~50-100 KB per
// SM are in flight without costing registers or load instructions.  When the codebooks fit (<= 96 KB) they
// are staged in shared memory once per CTA and the codeword gather is an LDS; ids of the next tile are
// prefetched into registers while the current one is computed.  Same lane-group-per-row arithmetic as above.
namespace bwdtma {
constexpr int THREADS = 512;
constexpr int STAGES = 2;
constexpr int kSmemBudget = 227 * 1024 - 1024;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 22); ++i) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity), "r"(200u)
            : "memory");
        if (ok) return true;
    }
    return false;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
}  // namespace bwdtma

// One level of the reverse walk in linear-combination form (every vector term of the rotation-trick Jacobian is
// a multiple of go, r or e; the scalars come from five lane-group dot products).
template <int LPR>
__device__ __forceinline__ void bwd_level_lc(int mode, float beta, float gl, const float4& r, const float4& go_in,
                                             const float4& t, const float4& e, float4& G, float4& gc) {
    const float4 go = make_float4(go_in.x - G.x, go_in.y - G.y, go_in.z - G.z, go_in.w - G.w);
    const float g2 = 2.f * gl;
    gc = make_float4(g2 * (e.x - r.x), g2 * (e.y - r.y), g2 * (e.z - r.z), g2 * (e.w - r.w));   // d emb_loss / d e
    float cg, cr, ce;                                   // G += cg*go + cr*r + ce*e (+ t)
    const float b2 = g2 * beta;                         // d query_loss / d r = b2 (r - e)
    if (mode == RQ_MODE_ROTATION_TRICK) {
        const float rr = group_sum<LPR>(dot4(r, r)), ee = group_sum<LPR>(dot4(e, e)), re = group_sum<LPR>(dot4(r, e));
        const float gr = group_sum<LPR>(dot4(go, r)), ge = group_sum<LPR>(dot4(go, e));
        const float nx = fast_sqrt(rr), ne = fast_sqrt(ee);
        const float iu = fast_rcp(nx + 1e-8f), iq = fast_rcp(ne + 1e-8f);
        const float lam = ne * fast_rcp(nx + 1e-6f);
        float ns = fmaf(iu * iu, rr, fmaf(2.f * iu * iq, re, iq * iq * ee));
        if (ns < 0.0625f) {                             // u ~ -q: the expansion cancels, sum directly (group-uniform)
            const float4 w = make_float4(fmaf(r.x, iu, e.x * iq), fmaf(r.y, iu, e.y * iq), fmaf(r.z, iu, e.z * iq),
                                         fmaf(r.w, iu, e.w * iq));
            float v = dot4(w, w);
            const unsigned gmask = (LPR == 32) ? 0xffffffffu : (((1u << LPR) - 1u) << ((threadIdx.x & 31) / LPR * LPR));
#pragma unroll
            for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gmask, v, o);
            ns = v;
        }
        const float iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f));
        const float gw = lam * iw * fmaf(iu, gr, iq * ge);      // gs.w, gs = lam*go, w = (u+q)*iw
        const float gq = lam * iq * ge;                          // gs.q
        cg = lam;
        cr = 2.f * iu * (gq - gw * iw) + b2;
        ce = -2.f * gw * iw * iq - b2;
    } else if (mode == RQ_MODE_STE) {
        cg = 1.f; cr = b2; ce = -b2;
    } else {                                                     // EVAL: out = C[ids]
        gc.x += go.x; gc.y += go.y; gc.z += go.z; gc.w += go.w;
        cg = 0.f; cr = b2; ce = -b2;
    }
    G.x += fmaf(cg, go.x, fmaf(cr, r.x, fmaf(ce, e.x, t.x)));
    G.y += fmaf(cg, go.y, fmaf(cr, r.y, fmaf(ce, e.y, t.y)));
    G.z += fmaf(cg, go.z, fmaf(cr, r.z, fmaf(ce, e.z, t.z)));
    G.w += fmaf(cg, go.w, fmaf(cr, r.w, fmaf(ce, e.w, t.w)));
}

struct BwdTmaCfg {
    int T;            // rows per tile
    int cb_smem;      // codebooks staged in shared memory
    int n_tiles;
    int stage_bytes, cb_bytes;
};

template <int D>
__global__ void __launch_bounds__(bwdtma::THREADS, 1) rq_bwd_tma_kernel(BwdParams p, BwdTmaCfg cfg) {
    using namespace bwdtma;
    constexpr int LPR = D / 4, RPW = 32 / LPR;
    constexpr int ROWS_PER_PASS = (THREADS / 32) * RPW;
    extern __shared__ __align__(128) uint8_t smem[];
    // [stage 0 | stage 1 | codebooks | barriers]; a stage = L x (res tile, g_emb tile), tile = T x D fp32
    uint8_t* sStage = smem;
    const float* sCb = reinterpret_cast<const float*>(smem + STAGES * cfg.stage_bytes);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + STAGES * cfg.stage_bytes + cfg.cb_bytes);   // [STAGES] + [1]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = lane % LPR, rsel = lane / LPR;
    const int T = cfg.T;
    const uint32_t tile_bytes = (uint32_t)T * D * 4;

    auto issue_tile = [&](int tile, int s) {             // one thread
        const int64_t row0 = (int64_t)tile * T;
        const int64_t rows = (p.B - row0 < T) ? (p.B - row0) : T;
        const uint32_t bytes = (uint32_t)rows * D * 4;
        mbar_expect_tx(&bar[s], bytes * (p.g_emb ? 2u : 1u) * p.L);
        for (int l = 0; l < p.L; ++l) {
            uint8_t* dst = sStage + s * cfg.stage_bytes + (size_t)l * 2 * tile_bytes;
            bulk_g2s(dst, p.res + ((int64_t)l * p.B + row0) * D, bytes, &bar[s]);
            if (p.g_emb) bulk_g2s(dst + tile_bytes, p.g_emb + ((int64_t)l * p.B + row0) * D, bytes, &bar[s]);
        }
    };
    if (tid == 0) {
        for (int i = 0; i <= STAGES; ++i) mbar_init(&bar[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (cfg.cb_smem) {
            mbar_expect_tx(&bar[STAGES], (uint32_t)cfg.cb_bytes);
            bulk_g2s(smem + STAGES * cfg.stage_bytes, p.cb, (uint32_t)cfg.cb_bytes, &bar[STAGES]);
        }
        for (int s = 0; s < STAGES; ++s) {
            const int tile = blockIdx.x + s * gridDim.x;
            if (tile < cfg.n_tiles) issue_tile(tile, s);
        }
    }
    __syncthreads();
    // ids of the first tile (this lane's row in every pass handled below is row = pass*ROWS_PER_PASS + warp*RPW + rsel)
    const int my_row = warp * RPW + rsel;
    constexpr int MAXP = 4;                               // passes per tile (T <= 4 * ROWS_PER_PASS)
    int32_t ids_next[MAXP][kBwdMaxL];
    auto load_ids = [&](int tile, int32_t (&dst)[MAXP][kBwdMaxL]) {
#pragma unroll
        for (int ps = 0; ps < MAXP; ++ps) {
            const int64_t row = (int64_t)tile * T + ps * ROWS_PER_PASS + my_row;
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l)
                dst[ps][l] = (ps * ROWS_PER_PASS < T && l < p.L && tile < cfg.n_tiles && row < p.B)
                                 ? (int32_t)__ldg(p.ids + (int64_t)l * p.B + row) : 0;
        }
    };
    load_ids(blockIdx.x, ids_next);
    bool timed_out = false;
    if (cfg.cb_smem && !mbar_wait(&bar[STAGES], 0)) timed_out = true;

    int it = 0;
    for (int tile = blockIdx.x; tile < cfg.n_tiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES;
        int32_t ids_cur[MAXP][kBwdMaxL];
#pragma unroll
        for (int ps = 0; ps < MAXP; ++ps)
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l) ids_cur[ps][l] = ids_next[ps][l];
        load_ids(tile + gridDim.x, ids_next);             // in flight during this tile's arithmetic
        if (!mbar_wait(&bar[s], (it / STAGES) & 1)) timed_out = true;
        const uint8_t* st = sStage + s * cfg.stage_bytes;
#pragma unroll
        for (int ps = 0; ps < MAXP; ++ps) {
            if (ps * ROWS_PER_PASS >= T) break;
            const int trow = ps * ROWS_PER_PASS + my_row;
            const int64_t row = (int64_t)tile * T + trow;
            const bool valid = row < p.B;
            const float gl = (p.g_loss && valid) ? __ldg(p.g_loss + row) : 0.f;
            float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
            // rows past B run the arithmetic on whatever the stage holds (the lane-group shuffles need every
            // lane of the warp) and are masked at the stores
#pragma unroll
            for (int l = kBwdMaxL - 1; l >= 0; --l) {
                if (l < p.L) {
                    const uint8_t* lv = st + (size_t)l * 2 * tile_bytes + (size_t)trow * D * 4 + sub * 16;
                    const float4 r = *reinterpret_cast<const float4*>(lv);
                    const float4 go = p.g_emb ? *reinterpret_cast<const float4*>(lv + tile_bytes)
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
                    const float4 t = (p.g_res && valid) ? ldg_stream4(p.g_res + ((int64_t)l * p.B + row) * D + sub * 4)
                                                        : make_float4(0.f, 0.f, 0.f, 0.f);
                    const int64_t coff = ((int64_t)l * p.K + ids_cur[ps][l]) * D + sub * 4;
                    const float4 e = cfg.cb_smem ? *reinterpret_cast<const float4*>(sCb + coff) : ldg_nc4(p.cb + coff);
                    float4 gc;
                    bwd_level_lc<LPR>(p.mode, p.beta, gl, r, go, t, e, G, gc);
                    if (valid) red_add_v4(p.g_cb + coff, gc.x, gc.y, gc.z, gc.w);
                }
            }
            if (valid) *reinterpret_cast<float4*>(p.g_x + row * D + sub * 4) = G;
        }
        // everybody is done with stage s: refill it with the tile STAGES iterations ahead
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            const int nt = tile + STAGES * gridDim.x;
            if (nt < cfg.n_tiles) issue_tile(nt, s);
        }
    }
    (void)timed_out;
}

template <int D>
static int launch_bwd_tma(const BwdParams& p, cudaStream_t st, bool* served) {
    using namespace bwdtma;
    *served = false;
    if (p.L > kBwdMaxL || getenv("RQ_BWD_NO_TMA")) return 0;
    if ((((uintptr_t)p.res | (uintptr_t)p.g_emb | (uintptr_t)p.cb | (uintptr_t)p.g_x | (uintptr_t)p.g_cb) & 15) != 0) return 0;
    constexpr int RPW = 32 / (D / 4);
    constexpr int ROWS_PER_PASS = (THREADS / 32) * RPW;
    BwdTmaCfg cfg;
    cfg.cb_bytes = 0;
    cfg.cb_smem = 0;
    const int64_t cb_all = (int64_t)p.L * p.K * D * 4;
    int T = 0;
    for (int cand = 4 * ROWS_PER_PASS; cand >= ROWS_PER_PASS; cand /= 2) {
        // prefer 64-row tiles with the codebooks resident; shrink the tile before giving the codebooks up
        const int64_t stage = (int64_t)p.L * 2 * cand * D * 4;
        if (cand > 64 && cand > ROWS_PER_PASS) continue;
        if (STAGES * stage + cb_all + 64 <= kSmemBudget && cb_all <= 96 * 1024) { T = cand; cfg.cb_smem = 1; cfg.cb_bytes = (int)cb_all; break; }
    }
    if (!T) {
        for (int cand = 4 * ROWS_PER_PASS; cand >= ROWS_PER_PASS; cand /= 2) {
            const int64_t stage = (int64_t)p.L * 2 * cand * D * 4;
            if (cand > 128 && cand > ROWS_PER_PASS) continue;
            if (STAGES * stage + 64 <= kSmemBudget) { T = cand; break; }
        }
    }
    if (!T) return 0;
    cfg.T = T;
    cfg.stage_bytes = p.L * 2 * T * D * 4;
    cfg.n_tiles = (int)((p.B + T - 1) / T);
    const size_t smem = (size_t)STAGES * cfg.stage_bytes + cfg.cb_bytes + 64;
    static bool attr_set = false;
    if (!attr_set) {
        RQ_CUDA(cudaFuncSetAttribute(rq_bwd_tma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    const int grid = cfg.n_tiles < sm_count() ? cfg.n_tiles : sm_count();
    rq_bwd_tma_kernel<D><<<grid, THREADS, smem, st>>>(p, cfg);
    RQ_LAUNCHED();
    *served = true;
    return 0;
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
    {
        bool served = false;
        int rc = 0;
        if (D == 32) rc = launch_bwd_tma<32>(p, st, &served);
        else if (D == 64) rc = launch_bwd_tma<64>(p, st, &served);
        else if (D == 16) rc = launch_bwd_tma<16>(p, st, &served);
        if (rc || served) return rc;
    }
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
