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

// Timing what-ifs (skip the reductions / stores / math of a kernel; wrong gradients by construction) and the switches
// that route around a kernel exist only in -DRQ_EXPERIMENTS builds; the release library reads no environment variable
// on this path.
#ifdef RQ_EXPERIMENTS
static bool bwd_knob(const char* name) { return getenv(name) != nullptr; }
static int bwd_dbg() { return getenv("RQ_BWD_DBG") ? atoi(getenv("RQ_BWD_DBG")) : 0; }
#else
static constexpr bool bwd_knob(const char*) { return false; }
static constexpr int bwd_dbg() { return 0; }
#endif

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

// The same level split in two: everything that does not depend on the gradient carried down from the level
// above (norms, r.e, the MUFU chain) is computed for all levels up front — independent work the scheduler can
// interleave — so the serial chain per level is two dot products and a handful of FMAs.
struct BwdLevelPre {
    float4 r, go, t, e;
    float lam, iu, iq, iw;
    int coff;
};
template <int LPR, int MODE>
__device__ __forceinline__ void bwd_level_pre(BwdLevelPre& v) {
    if (MODE == RQ_MODE_ROTATION_TRICK) {
        const float rr = group_sum<LPR>(dot4(v.r, v.r)), ee = group_sum<LPR>(dot4(v.e, v.e));
        const float re = group_sum<LPR>(dot4(v.r, v.e));
        const float nx = fast_sqrt(rr), ne = fast_sqrt(ee);
        v.iu = fast_rcp(nx + 1e-8f);
        v.iq = fast_rcp(ne + 1e-8f);
        v.lam = ne * fast_rcp(nx + 1e-6f);
        float ns = fmaf(v.iu * v.iu, rr, fmaf(2.f * v.iu * v.iq, re, v.iq * v.iq * ee));
        if (ns < 0.0625f) {                             // u ~ -q: the expansion cancels, sum directly (group-uniform)
            const float4 w = make_float4(fmaf(v.r.x, v.iu, v.e.x * v.iq), fmaf(v.r.y, v.iu, v.e.y * v.iq),
                                         fmaf(v.r.z, v.iu, v.e.z * v.iq), fmaf(v.r.w, v.iu, v.e.w * v.iq));
            float x = dot4(w, w);
            const unsigned gmask = (LPR == 32) ? 0xffffffffu : (((1u << LPR) - 1u) << ((threadIdx.x & 31) / LPR * LPR));
#pragma unroll
            for (int o = LPR / 2; o > 0; o >>= 1) x += __shfl_xor_sync(gmask, x, o);
            ns = x;
        }
        v.iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f));
    }
}
template <int LPR, int MODE>
__device__ __forceinline__ void bwd_level_chain(const BwdLevelPre& v, float beta, float gl, float4& G, float4& gc) {
    const float4 go = make_float4(v.go.x - G.x, v.go.y - G.y, v.go.z - G.z, v.go.w - G.w);
    const float g2 = 2.f * gl, b2 = g2 * beta;
    gc = make_float4(g2 * (v.e.x - v.r.x), g2 * (v.e.y - v.r.y), g2 * (v.e.z - v.r.z), g2 * (v.e.w - v.r.w));
    float cg, cr, ce;
    if (MODE == RQ_MODE_ROTATION_TRICK) {
        const float gr = group_sum<LPR>(dot4(go, v.r)), ge = group_sum<LPR>(dot4(go, v.e));
        const float gw = v.lam * v.iw * fmaf(v.iu, gr, v.iq * ge);
        const float gq = v.lam * v.iq * ge;
        cg = v.lam;
        cr = 2.f * v.iu * (gq - gw * v.iw) + b2;
        ce = -2.f * gw * v.iw * v.iq - b2;
    } else if (MODE == RQ_MODE_STE) {
        cg = 1.f; cr = b2; ce = -b2;
    } else {
        gc.x += go.x; gc.y += go.y; gc.z += go.z; gc.w += go.w;
        cg = 0.f; cr = b2; ce = -b2;
    }
    G.x += fmaf(cg, go.x, fmaf(cr, v.r.x, fmaf(ce, v.e.x, v.t.x)));
    G.y += fmaf(cg, go.y, fmaf(cr, v.r.y, fmaf(ce, v.e.y, v.t.y)));
    G.z += fmaf(cg, go.z, fmaf(cr, v.r.z, fmaf(ce, v.e.z, v.t.z)));
    G.w += fmaf(cg, go.w, fmaf(cr, v.r.w, fmaf(ce, v.e.w, v.t.w)));
}

struct BwdTmaCfg {
    int T;            // rows per tile
    int cb_smem;      // codebooks staged in shared memory
    int n_tiles;
    int stage_bytes, cb_bytes;
    int dbg;          // experiments (RQ_BWD_DBG): 1 = no codebook-gradient reductions, 2 = no g_x stores
};

template <int D, int MODE>
__global__ void __launch_bounds__(bwdtma::THREADS, 1) rq_bwd_tma_kernel(BwdParams p, BwdTmaCfg cfg) {
    using namespace bwdtma;
    constexpr int LPR = D / 4, RPW = 32 / LPR;
    constexpr int ROWS_PER_PASS = (THREADS / 32) * RPW;
    extern __shared__ __align__(128) uint8_t smem[];
    // [stage 0 | stage 1 | ids 0 | ids 1 | codebooks | barriers]; a stage = L x (res tile, g_emb tile),
    // tile = T x D fp32; ids s = [L][T] int64
    uint8_t* sStage = smem;
    const int T = cfg.T;
    const int ids_bytes = p.L * T * 8 + T * 4;              // [L][T] int64 ids, then [T] fp32 g_loss
    uint8_t* sIds = smem + STAGES * cfg.stage_bytes;
    const float* sCb = reinterpret_cast<const float*>(sIds + STAGES * ids_bytes);
    uint64_t* bar = reinterpret_cast<uint64_t*>(sIds + STAGES * ids_bytes + cfg.cb_bytes);   // [STAGES] + [1]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = lane % LPR, rsel = lane / LPR;
    const uint32_t tile_bytes = (uint32_t)T * D * 4;
    const bool has_gemb = p.g_emb != nullptr, has_gres = p.g_res != nullptr, cb_smem = cfg.cb_smem != 0;

    auto issue_tile = [&](int tile, int s) {             // one thread
        const int64_t row0 = (int64_t)tile * T;
        const int64_t rows = (p.B - row0 < T) ? (p.B - row0) : T;
        const uint32_t bytes = (uint32_t)rows * D * 4;
        mbar_expect_tx(&bar[s], bytes * (has_gemb ? 2u : 1u) * p.L);
        for (int l = 0; l < p.L; ++l) {
            uint8_t* dst = sStage + s * cfg.stage_bytes + (size_t)l * 2 * tile_bytes;
            bulk_g2s(dst, p.res + ((int64_t)l * p.B + row0) * D, bytes, &bar[s]);
            if (has_gemb) bulk_g2s(dst + tile_bytes, p.g_emb + ((int64_t)l * p.B + row0) * D, bytes, &bar[s]);
        }
    };
    // ids of a tile: 8-byte cp.async per (level, row), one element per thread and round
    auto issue_ids = [&](int tile, int s) {
        if (tile < cfg.n_tiles) {
            for (int i = tid; i < p.L * T; i += THREADS) {
                const int l = i / T, rr = i - l * T;
                const int64_t row = (int64_t)tile * T + rr;
                if (row < p.B) {
                    const uint32_t dst = smem_u32(sIds + s * ids_bytes + i * 8);
                    const int64_t* src = p.ids + (int64_t)l * p.B + row;
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
                }
            }
            if (p.g_loss && tid < T && (int64_t)tile * T + tid < p.B) {
                const uint32_t dst = smem_u32(sIds + s * ids_bytes + p.L * T * 8 + tid * 4);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(p.g_loss + (int64_t)tile * T + tid)
                             : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (tid == 0) {
        for (int i = 0; i <= STAGES; ++i) mbar_init(&bar[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (cb_smem) {
            mbar_expect_tx(&bar[STAGES], (uint32_t)cfg.cb_bytes);
            bulk_g2s(const_cast<float*>(sCb), p.cb, (uint32_t)cfg.cb_bytes, &bar[STAGES]);
        }
        for (int s = 0; s < STAGES; ++s) {
            const int tile = blockIdx.x + s * gridDim.x;
            if (tile < cfg.n_tiles) issue_tile(tile, s);
        }
    }
    for (int s = 0; s < STAGES; ++s) issue_ids(blockIdx.x + s * gridDim.x, s);
    __syncthreads();
    bool timed_out = false;
    if (cb_smem && !mbar_wait(&bar[STAGES], 0)) timed_out = true;
    const int my_row = warp * RPW + rsel;

    int it = 0;
    for (int tile = blockIdx.x; tile < cfg.n_tiles; tile += gridDim.x, ++it) {
        const int s = it % STAGES;
        // this tile's ids: all but the newest cp.async group are complete, then make them CTA-visible
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 1) : "memory");
        __syncthreads();
        if (!mbar_wait(&bar[s], (it / STAGES) & 1)) timed_out = true;
        const uint8_t* st = sStage + s * cfg.stage_bytes;
        const int64_t* ids_s = reinterpret_cast<const int64_t*>(sIds + s * ids_bytes);
#pragma unroll 1
        for (int trow = my_row; trow < ((cfg.dbg & 4) ? 0 : T); trow += ROWS_PER_PASS) {
            const int64_t row = (int64_t)tile * T + trow;
            const bool valid = row < p.B;
            const float gl = (p.g_loss && valid) ? reinterpret_cast<const float*>(ids_s + p.L * T)[trow] : 0.f;
            float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
            // rows past B run the arithmetic on whatever the stage holds (the lane-group shuffles need every
            // lane of the warp) and are masked at the stores
            BwdLevelPre lv[kBwdMaxL];
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l) {
                if (l < p.L) {
                    const uint8_t* src = st + (size_t)l * 2 * tile_bytes + (size_t)trow * D * 4 + sub * 16;
                    lv[l].r = *reinterpret_cast<const float4*>(src);
                    lv[l].go = has_gemb ? *reinterpret_cast<const float4*>(src + tile_bytes) : z4;
                    lv[l].t = (has_gres && valid) ? ldg_stream4(p.g_res + ((int64_t)l * p.B + row) * D + sub * 4) : z4;
                    const int id = valid ? (int)ids_s[l * T + trow] : 0;
                    lv[l].coff = (l * p.K + id) * D + sub * 4;
                    lv[l].e = cb_smem ? *reinterpret_cast<const float4*>(sCb + lv[l].coff) : ldg_nc4(p.cb + lv[l].coff);
                }
            }
#pragma unroll
            for (int l = 0; l < kBwdMaxL; ++l)
                if (l < p.L) bwd_level_pre<LPR, MODE>(lv[l]);
#pragma unroll
            for (int l = kBwdMaxL - 1; l >= 0; --l) {
                if (l < p.L) {
                    float4 gc;
                    bwd_level_chain<LPR, MODE>(lv[l], p.beta, gl, G, gc);
                    if (valid && !(cfg.dbg & 1)) red_add_v4(p.g_cb + lv[l].coff, gc.x, gc.y, gc.z, gc.w);
                }
            }
            if (valid && !(cfg.dbg & 2)) *reinterpret_cast<float4*>(p.g_x + row * D + sub * 4) = G;
        }
        // everybody is done with stage s: refill it with the tile STAGES iterations ahead
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        const int nt = tile + STAGES * gridDim.x;
        if (tid == 0 && nt < cfg.n_tiles) issue_tile(nt, s);
        issue_ids(nt, s);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    (void)timed_out;
}

// ---------------------------------------------------------------------------------------------------------
// Row-per-thread variant for D = 32 (the headline shape).  The lane-group kernels above spend 8 lanes on a
// row, so every warp instruction advances 4 rows; here a thread owns a row (as in the forward kernel) and a
// warp instruction advances 32 — ~4x fewer instructions, no shuffles in the arithmetic.
//
// Every WARP is an independent pipeline over 32-row tiles: the (residual, upstream gradient) rows of one
// (tile, level) are one 8 KB stage, fetched with 16-byte cp.async in the coalesced pattern (8 lanes per
// 128-byte row) into a 128-byte-swizzled layout and read back row-wise by the owning lanes without bank
// conflicts; a 3-stage ring per warp keeps two stages (16 KB) in flight while one is processed, 8 warps per
// SM.  Rows are warp-private, so __syncwarp is the only synchronisation.  Codeword gathers, the codebook-
// gradient reductions (REDG.128) and the g_x stores also use the coalesced pattern: the stage just consumed
// doubles as the staging tile.
namespace bwdrow {
constexpr int D = 32, TW = 32, WARPS = 4, THREADS = 32 * WARPS, STAGES = 3;
constexpr int STAGE_BYTES = 2 * TW * D * 4;            // 8 KB: residual rows, then gradient rows
__device__ __forceinline__ uint32_t sw_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}
}  // namespace bwdrow

template <int MODE>
__global__ void __launch_bounds__(bwdrow::THREADS, 2) rq_bwd_row_kernel(BwdParams p, int n_tiles, int dbg) {
    using namespace bwdrow;
    extern __shared__ __align__(1024) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* ring = smem + warp * (STAGES * STAGE_BYTES);
    const uint32_t ring_u32 = (uint32_t)__cvta_generic_to_shared(ring);
    const bool has_gemb = p.g_emb != nullptr;
    const int chunk = lane & 7;
    const uint32_t co_sw = sw_off(lane >> 3, chunk);        // coalesced pattern: chunk of rows 4 i + lane / 8
    auto co_off = [&](int i) -> uint32_t { return (co_sw + 512u * i) ^ ((i & 1) ? 64u : 0u); };
    const int wglobal = blockIdx.x * WARPS + warp, nwarps = gridDim.x * WARPS;
    const int my_tiles = (n_tiles > wglobal) ? (n_tiles - 1 - wglobal) / nwarps + 1 : 0;
    const int n_items = my_tiles * p.L;                      // (tile, level) pairs in processing order

    // item q -> tile index, level (levels walk downwards inside a tile)
    auto issue_item = [&](int q) {
        if (q < n_items && !(dbg & 8)) {
            const int ti = q / p.L, l = p.L - 1 - (q - ti * p.L);
            const int64_t row0 = ((int64_t)wglobal + (int64_t)ti * nwarps) * TW;
            const uint32_t dst0 = ring_u32 + (q % STAGES) * STAGE_BYTES;
            const float* src_r = p.res + ((int64_t)l * p.B + row0 + (lane >> 3)) * D + chunk * 4;
            const float* src_g = has_gemb ? p.g_emb + ((int64_t)l * p.B + row0 + (lane >> 3)) * D + chunk * 4 : nullptr;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (row0 + (lane >> 3) + 4 * i < p.B) {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + co_off(i)), "l"(src_r + i * 4 * D)
                                 : "memory");
                    if (has_gemb)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + TW * D * 4 + co_off(i)),
                                     "l"(src_g + i * 4 * D)
                                     : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto gather = [&](int l, int id, float4 (&g)[8]) {       // coalesced codeword fetch for the warp's 32 rows
        const float* base = p.cb + (int64_t)l * p.K * D + chunk * 4;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int k = __shfl_sync(0xffffffffu, id, 4 * i + (lane >> 3));
            g[i] = __ldg(reinterpret_cast<const float4*>(base + (int64_t)k * D));
        }
    };
    int ids_n[kBwdMaxL];
    float gl_n = 0.f;
    auto load_ids = [&](int ti) {                            // ids / g_loss of this lane's row in tile ti
        const int64_t row = ((int64_t)wglobal + (int64_t)ti * nwarps) * TW + lane;
        const bool ok = ti < my_tiles && row < p.B;
#pragma unroll
        for (int l = 0; l < kBwdMaxL; ++l) ids_n[l] = (ok && l < p.L) ? (int)__ldg(p.ids + (int64_t)l * p.B + row) : 0;
        gl_n = (ok && p.g_loss) ? __ldg(p.g_loss + row) : 0.f;
    };
    auto pick = [&](const int (&a)[kBwdMaxL], int l) { return (l == 3) ? a[3] : (l == 2) ? a[2] : (l == 1) ? a[1] : a[0]; };

    issue_item(0);
    issue_item(1);
    load_ids(0);
    float4 gnext[8];
    if (my_tiles > 0) gather(p.L - 1, pick(ids_n, p.L - 1), gnext);

    int q = 0;
    for (int ti = 0; ti < my_tiles; ++ti) {
        const int64_t row0 = ((int64_t)wglobal + (int64_t)ti * nwarps) * TW;
        const int64_t row = row0 + lane;
        const bool valid = row < p.B;
        int ids[kBwdMaxL];
#pragma unroll
        for (int l = 0; l < kBwdMaxL; ++l) ids[l] = ids_n[l];
        const float gl = gl_n;
        load_ids(ti + 1);                                    // next tile's ids / g_loss in flight
        const float g2 = 2.f * gl, b2 = g2 * p.beta;
        float G[D];
#pragma unroll
        for (int j = 0; j < D; ++j) G[j] = 0.f;
#pragma unroll 1
        for (int l = (dbg & 4) ? -1 : p.L - 1; l >= 0; --l, ++q) {
            const int id = pick(ids, l);
            issue_item(q + 2);                               // into the stage consumed at q - 1
            asm volatile("cp.async.wait_group 2;" ::: "memory");   // item q landed (two newer groups may be pending)
            __syncwarp();
            uint8_t* st = ring + (q % STAGES) * STAGE_BYTES;
            float e[D], r[D], go[D];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(st + sw_off(lane, c));
                r[4 * c] = v.x; r[4 * c + 1] = v.y; r[4 * c + 2] = v.z; r[4 * c + 3] = v.w;
            }
            if (has_gemb) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float4 v = *reinterpret_cast<const float4*>(st + TW * D * 4 + sw_off(lane, c));
                    go[4 * c] = v.x - G[4 * c]; go[4 * c + 1] = v.y - G[4 * c + 1];
                    go[4 * c + 2] = v.z - G[4 * c + 2]; go[4 * c + 3] = v.w - G[4 * c + 3];
                }
            } else {
#pragma unroll
                for (int j = 0; j < D; ++j) go[j] = -G[j];
            }
            __syncwarp();                                    // the stage is now free: it doubles as the staging tile
            uint8_t* scratch = st;
#pragma unroll
            for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(scratch + co_off(i)) = gnext[i];
            __syncwarp();
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(scratch + sw_off(lane, c));
                e[4 * c] = v.x; e[4 * c + 1] = v.y; e[4 * c + 2] = v.z; e[4 * c + 3] = v.w;
            }
            __syncwarp();
            // next item's codewords go out now (next level of this tile, or the first level of the next tile)
            if (l > 0) gather(l - 1, pick(ids, l - 1), gnext);
            else if (ti + 1 < my_tiles) gather(p.L - 1, pick(ids_n, p.L - 1), gnext);
            float cg, cr, ce;
            if (MODE == RQ_MODE_ROTATION_TRICK) {
                float rr = 0.f, ee = 0.f, re = 0.f, gr = 0.f, ge = 0.f;
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    rr = fmaf(r[j], r[j], rr); ee = fmaf(e[j], e[j], ee); re = fmaf(r[j], e[j], re);
                    gr = fmaf(go[j], r[j], gr); ge = fmaf(go[j], e[j], ge);
                }
                const float nx = fast_sqrt(rr), ne = fast_sqrt(ee);
                const float iu = fast_rcp(nx + 1e-8f), iq = fast_rcp(ne + 1e-8f);
                const float lam = ne * fast_rcp(nx + 1e-6f);
                float ns = fmaf(iu * iu, rr, fmaf(2.f * iu * iq, re, iq * iq * ee));
                if (ns < 0.0625f) {
                    ns = 0.f;
#pragma unroll
                    for (int j = 0; j < D; ++j) {
                        const float w = fmaf(r[j], iu, e[j] * iq);
                        ns = fmaf(w, w, ns);
                    }
                }
                const float iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f));
                const float gw = lam * iw * fmaf(iu, gr, iq * ge);
                const float gq = lam * iq * ge;
                cg = lam;
                cr = 2.f * iu * (gq - gw * iw) + b2;
                ce = -2.f * gw * iw * iq - b2;
            } else if (MODE == RQ_MODE_STE) {
                cg = 1.f; cr = b2; ce = -b2;
            } else {
                cg = 0.f; cr = b2; ce = -b2;
            }
            // this row's codebook gradient 2 (e - r) g_loss (+ go when the output is the codeword itself):
            // staged, then reduced into g_cb by the coalesced pattern (4 codeword rows per REDG.128 instruction)
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float4 v = make_float4(g2 * (e[4 * c] - r[4 * c]), g2 * (e[4 * c + 1] - r[4 * c + 1]),
                                       g2 * (e[4 * c + 2] - r[4 * c + 2]), g2 * (e[4 * c + 3] - r[4 * c + 3]));
                if (MODE == RQ_MODE_EVAL) { v.x += go[4 * c]; v.y += go[4 * c + 1]; v.z += go[4 * c + 2]; v.w += go[4 * c + 3]; }
                *reinterpret_cast<float4*>(scratch + sw_off(lane, c)) = v;
            }
            __syncwarp();
            {
                float* gbase = p.g_cb + (int64_t)l * p.K * D + chunk * 4;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = __shfl_sync(0xffffffffu, id, 4 * i + (lane >> 3));
                    const float4 v = *reinterpret_cast<const float4*>(scratch + co_off(i));
                    if (row0 + (lane >> 3) + 4 * i < p.B && !(dbg & 1)) red_add_v4(gbase + (int64_t)k * D, v.x, v.y, v.z, v.w);
                }
            }
            if (p.g_res && valid) {
                const float* tg = p.g_res + ((int64_t)l * p.B + row) * D;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float4 v = ldg_stream4(tg + 4 * c);
                    G[4 * c] += v.x; G[4 * c + 1] += v.y; G[4 * c + 2] += v.z; G[4 * c + 3] += v.w;
                }
            }
#pragma unroll
            for (int j = 0; j < D; ++j) G[j] += fmaf(cg, go[j], fmaf(cr, r[j], ce * e[j]));
            if (l == 0) {
                // g_x: staged in the same tile and written out coalesced
                __syncwarp();
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    *reinterpret_cast<float4*>(scratch + sw_off(lane, c)) =
                        make_float4(G[4 * c], G[4 * c + 1], G[4 * c + 2], G[4 * c + 3]);
                __syncwarp();
                float* dst = p.g_x + (row0 + (lane >> 3)) * D + chunk * 4;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 v = *reinterpret_cast<const float4*>(scratch + co_off(i));
                    if (row0 + (lane >> 3) + 4 * i < p.B) *reinterpret_cast<float4*>(dst + i * 4 * D) = v;
                }
            }
            __syncwarp();                                    // staging reads done before the stage is refilled
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// ---- D = 32, two lanes per row, double-buffered ---------------------------------------------------------------
// Per warp the work is one serial chain (load -> dot products -> update -> scatter, level after level), so the
// launch takes as long as the slowest warp's chain (tools/time_bwd.py: the phases add up instead of overlapping).
// Here a warp's item is 16 rows x 1 level with TWO lanes per row (16 dims each): the per-item chain is half as
// long, a 6 KB item fits twice in the warp's 12 KB, and the next item (next level, or the next tile's first
// level) is in flight while the current one is computed.
namespace bwdrow3 {
constexpr int D = 32, TW = 16, WARPS = 8, THREADS = 32 * WARPS;
constexpr int TILE_BYTES = TW * D * 4;           // 2 KB
constexpr int ITEM_BYTES = 3 * TILE_BYTES;       // residual rows | gradient rows | codeword rows
constexpr int WARP_BYTES = 2 * ITEM_BYTES;
}  // namespace bwdrow3

template <int MODE>
__global__ void __launch_bounds__(bwdrow3::THREADS, 2) rq_bwd_row3_kernel(BwdParams p, int n_tiles) {
    using namespace bwdrow3;
    using bwdrow::sw_off;
    extern __shared__ __align__(1024) uint8_t smem[];
    pdl_launch_dependents();
    pdl_wait();                     // launched with PDL: scheduled early, reads nothing before the forward has completed
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* wbuf = smem + warp * WARP_BYTES;
    const uint32_t wbuf_u32 = (uint32_t)__cvta_generic_to_shared(wbuf);
    const bool has_gemb = p.g_emb != nullptr;
    const int chunk = lane & 7, sub = lane >> 3;            // coalesced side: 8 lanes per row, 4 rows per instruction
    const int ri = lane >> 1, half = lane & 1;              // compute side: 2 lanes per row, 16 dims each
    const int wglobal = blockIdx.x * WARPS + warp, nwarps = gridDim.x * WARPS;
    uint32_t co[4], ro[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        co[i] = sw_off(4 * i + sub, chunk);
        ro[i] = sw_off(ri, 4 * half + i);
    }

    int ids_c[kBwdMaxL], ids_n[kBwdMaxL];
    float gl_c = 0.f, gl_n = 0.f;
    auto load_ids = [&](int tile, int (&ids)[kBwdMaxL], float& gl) {
        const int64_t row = (int64_t)tile * TW + ri;
        const bool ok = tile < n_tiles && row < p.B;
#pragma unroll
        for (int l = 0; l < kBwdMaxL; ++l) ids[l] = (ok && l < p.L) ? (int)__ldg(p.ids + (int64_t)l * p.B + row) : 0;
        gl = (ok && p.g_loss) ? __ldg(p.g_loss + row) : 0.f;
    };
    auto pick = [&](const int (&a)[kBwdMaxL], int l) { return (l == 3) ? a[3] : (l == 2) ? a[2] : (l == 1) ? a[1] : a[0]; };
    // fill buffer `b` with (tile, level l); `id` = this lane's row id at that level
    auto issue = [&](int b, int tile, int l, int id) {
        const int64_t row0 = (int64_t)tile * TW;
        const uint32_t dst = wbuf_u32 + b * ITEM_BYTES;
        const float* src_r = p.res + ((int64_t)l * p.B + row0 + sub) * D + chunk * 4;
        const float* src_g = has_gemb ? p.g_emb + ((int64_t)l * p.B + row0 + sub) * D + chunk * 4 : nullptr;
        const float* cbl = p.cb + (int64_t)l * p.K * D + chunk * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int k = __shfl_sync(0xffffffffu, id, 2 * (4 * i + sub));
            if (row0 + sub + 4 * i < p.B) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + co[i]), "l"(src_r + i * 4 * D) : "memory");
                if (has_gemb)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + TILE_BYTES + co[i]), "l"(src_g + i * 4 * D)
                                 : "memory");
            }
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 2 * TILE_BYTES + co[i]), "l"(cbl + (int64_t)k * D)
                         : "memory");
        }
    };

    int cur = wglobal;
    load_ids(cur, ids_c, gl_c);
    if (cur < n_tiles) issue(0, cur, p.L - 1, pick(ids_c, p.L - 1));
    asm volatile("cp.async.commit_group;" ::: "memory");
    int buf = 0;
    while (cur < n_tiles) {
        const int nxt = cur + nwarps;
        load_ids(nxt, ids_n, gl_n);
        const int64_t row0 = (int64_t)cur * TW;
        const float g2 = 2.f * gl_c, b2 = g2 * p.beta;
        float G[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) G[j] = 0.f;
#pragma unroll 1
        for (int l = p.L - 1; l >= 0; --l) {
            const int id = pick(ids_c, l);
            if (l > 0) issue(buf ^ 1, cur, l - 1, pick(ids_c, l - 1));
            else if (nxt < n_tiles) issue(buf ^ 1, nxt, p.L - 1, pick(ids_n, p.L - 1));
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncwarp();
            uint8_t* tR = wbuf + buf * ITEM_BYTES;
            uint8_t* tG = tR + TILE_BYTES;
            uint8_t* tE = tG + TILE_BYTES;
            float cg, cr, ce;
            // the row halves stay in registers from the dot products to the update (occupancy is bounded by the
            // 12 KB of shared memory per warp, not by registers)
            float R_[16], E_[16], Gv_[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float4 r4 = *reinterpret_cast<const float4*>(tR + ro[c]);
                const float4 e4 = *reinterpret_cast<const float4*>(tE + ro[c]);
                float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (has_gemb) g4 = *reinterpret_cast<const float4*>(tG + ro[c]);
                R_[4 * c] = r4.x; R_[4 * c + 1] = r4.y; R_[4 * c + 2] = r4.z; R_[4 * c + 3] = r4.w;
                E_[4 * c] = e4.x; E_[4 * c + 1] = e4.y; E_[4 * c + 2] = e4.z; E_[4 * c + 3] = e4.w;
                Gv_[4 * c] = g4.x - G[4 * c]; Gv_[4 * c + 1] = g4.y - G[4 * c + 1];
                Gv_[4 * c + 2] = g4.z - G[4 * c + 2]; Gv_[4 * c + 3] = g4.w - G[4 * c + 3];
            }
            if (MODE == RQ_MODE_ROTATION_TRICK) {
                float rr = 0.f, ee = 0.f, re = 0.f, gr = 0.f, ge = 0.f;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    rr = fmaf(R_[j], R_[j], rr); ee = fmaf(E_[j], E_[j], ee); re = fmaf(R_[j], E_[j], re);
                    gr = fmaf(Gv_[j], R_[j], gr); ge = fmaf(Gv_[j], E_[j], ge);
                }
                rr += __shfl_xor_sync(0xffffffffu, rr, 1); ee += __shfl_xor_sync(0xffffffffu, ee, 1);
                re += __shfl_xor_sync(0xffffffffu, re, 1); gr += __shfl_xor_sync(0xffffffffu, gr, 1);
                ge += __shfl_xor_sync(0xffffffffu, ge, 1);
                const float nx = fast_sqrt(rr), ne = fast_sqrt(ee);
                const float iu = fast_rcp(nx + 1e-8f), iq = fast_rcp(ne + 1e-8f);
                const float lam = ne * fast_rcp(nx + 1e-6f);
                float ns = fmaf(iu * iu, rr, fmaf(2.f * iu * iq, re, iq * iq * ee));
                // |u+q| < 1/4 (codeword antipodal to the residual): recompute without cancellation; both lanes of a
                // row see the same ns, so the branch is uniform per row and the shuffle is taken by whole pairs
                const bool redo = ns < 0.0625f;
                if (__any_sync(0xffffffffu, redo)) {
                    float n2 = 0.f;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float w = fmaf(R_[j], iu, E_[j] * iq);
                        n2 = fmaf(w, w, n2);
                    }
                    n2 += __shfl_xor_sync(0xffffffffu, n2, 1);
                    if (redo) ns = n2;
                }
                const float iw = fast_rcp(fmaxf(fast_sqrt(ns), 1e-6f));
                const float gw = lam * iw * fmaf(iu, gr, iq * ge);
                const float gq = lam * iq * ge;
                cg = lam;
                cr = 2.f * iu * (gq - gw * iw) + b2;
                ce = -2.f * gw * iw * iq - b2;
            } else if (MODE == RQ_MODE_STE) {
                cg = 1.f; cr = b2; ce = -b2;
            } else {
                cg = 0.f; cr = b2; ce = -b2;
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float rvj = R_[4 * c + j], evj = E_[4 * c + j], gvj = Gv_[4 * c + j];
                    v[j] = g2 * (evj - rvj);
                    if (MODE == RQ_MODE_EVAL) v[j] += gvj;
                    G[4 * c + j] += fmaf(cg, gvj, fmaf(cr, rvj, ce * evj));
                }
                *reinterpret_cast<float4*>(tG + ro[c]) = make_float4(v[0], v[1], v[2], v[3]);
            }
            __syncwarp();
            {
                float* gbase = p.g_cb + (int64_t)l * p.K * D + chunk * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int k = __shfl_sync(0xffffffffu, id, 2 * (4 * i + sub));
                    const float4 v = *reinterpret_cast<const float4*>(tG + co[i]);
                    if (row0 + sub + 4 * i < p.B) red_add_v4(gbase + (int64_t)k * D, v.x, v.y, v.z, v.w);
                }
            }
            if (l == 0) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    *reinterpret_cast<float4*>(tR + ro[c]) = make_float4(G[4 * c], G[4 * c + 1], G[4 * c + 2], G[4 * c + 3]);
                __syncwarp();
                float* dst = p.g_x + (row0 + sub) * D + chunk * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 v = *reinterpret_cast<const float4*>(tR + co[i]);
                    if (row0 + sub + 4 * i < p.B) *reinterpret_cast<float4*>(dst + i * 4 * D) = v;
                }
            }
            __syncwarp();                                    // reads of this buffer done before it is refilled
            buf ^= 1;
        }
        cur = nxt;
#pragma unroll
        for (int l = 0; l < kBwdMaxL; ++l) ids_c[l] = ids_n[l];
        gl_c = gl_n;
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

static int launch_bwd_row3(const BwdParams& p, cudaStream_t st, bool* served) {
    using namespace bwdrow3;
    *served = false;
    if (p.L > kBwdMaxL || bwd_knob("RQ_BWD_NO_ROW3")) return 0;
    if (p.g_res) return 0;                                    // residual-output gradients: ring-buffered kernel
    if ((((uintptr_t)p.res | (uintptr_t)p.g_emb | (uintptr_t)p.cb | (uintptr_t)p.g_x | (uintptr_t)p.g_cb) & 15) != 0) return 0;
    const size_t smem = (size_t)WARPS * WARP_BYTES;            // 96 KB: two CTAs (16 warps) per SM
    const int n_tiles = (int)((p.B + TW - 1) / TW);
    const int ctas_needed = (n_tiles + WARPS - 1) / WARPS;
    const int grid = ctas_needed < 2 * sm_count() ? ctas_needed : 2 * sm_count();
    auto launch = [&](auto kern) -> int {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
        RQ_CUDA(launch_pdl(kern, dim3(grid), dim3(THREADS), smem, st, p, n_tiles));
        return 0;
    };
    int rc;
    if (p.mode == RQ_MODE_ROTATION_TRICK) rc = launch(rq_bwd_row3_kernel<RQ_MODE_ROTATION_TRICK>);
    else if (p.mode == RQ_MODE_STE) rc = launch(rq_bwd_row3_kernel<RQ_MODE_STE>);
    else rc = launch(rq_bwd_row3_kernel<RQ_MODE_EVAL>);
    if (rc) return rc;
    RQ_LAUNCHED();
    *served = true;
    return 0;
}

static int launch_bwd_row(const BwdParams& p, cudaStream_t st, bool* served) {
    using namespace bwdrow;
    *served = false;
    if (p.L > kBwdMaxL || bwd_knob("RQ_BWD_NO_ROW")) return 0;
    if ((((uintptr_t)p.res | (uintptr_t)p.g_emb | (uintptr_t)p.cb | (uintptr_t)p.g_x | (uintptr_t)p.g_cb) & 15) != 0) return 0;
    const size_t smem = (size_t)WARPS * STAGES * STAGE_BYTES;  // 96 KB: two CTAs per SM
    const int n_tiles = (int)((p.B + TW - 1) / TW);
    const int ctas_needed = (n_tiles + WARPS - 1) / WARPS;
    const int grid = ctas_needed < 2 * sm_count() ? ctas_needed : 2 * sm_count();
    const int dbg = bwd_dbg();
    auto launch = [&](auto kern) -> int {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
        kern<<<grid, THREADS, smem, st>>>(p, n_tiles, dbg);
        return 0;
    };
    int rc;
    if (p.mode == RQ_MODE_ROTATION_TRICK) rc = launch(rq_bwd_row_kernel<RQ_MODE_ROTATION_TRICK>);
    else if (p.mode == RQ_MODE_STE) rc = launch(rq_bwd_row_kernel<RQ_MODE_STE>);
    else rc = launch(rq_bwd_row_kernel<RQ_MODE_EVAL>);
    if (rc) return rc;
    RQ_LAUNCHED();
    *served = true;
    return 0;
}

template <int D>
static int launch_bwd_tma(const BwdParams& p, cudaStream_t st, bool* served) {
    using namespace bwdtma;
    *served = false;
    if (p.L > kBwdMaxL || bwd_knob("RQ_BWD_NO_TMA")) return 0;
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
        if (STAGES * (stage + (int64_t)p.L * cand * 8 + cand * 4) + cb_all + 64 <= kSmemBudget && cb_all <= 96 * 1024) { T = cand; cfg.cb_smem = 1; cfg.cb_bytes = (int)cb_all; break; }
    }
    if (!T) {
        for (int cand = 4 * ROWS_PER_PASS; cand >= ROWS_PER_PASS; cand /= 2) {
            const int64_t stage = (int64_t)p.L * 2 * cand * D * 4;
            if (cand > 128 && cand > ROWS_PER_PASS) continue;
            if (STAGES * (stage + (int64_t)p.L * cand * 8 + cand * 4) + 64 <= kSmemBudget) { T = cand; break; }
        }
    }
    if (!T) return 0;
    cfg.T = T;
    cfg.dbg = bwd_dbg();
    cfg.stage_bytes = p.L * 2 * T * D * 4;
    cfg.n_tiles = (int)((p.B + T - 1) / T);
    const size_t smem = (size_t)STAGES * (cfg.stage_bytes + p.L * T * 8 + T * 4) + cfg.cb_bytes + 64;
    const int grid = cfg.n_tiles < sm_count() ? cfg.n_tiles : sm_count();
    auto launch = [&](auto kern) -> int {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, 227 * 1024));
        kern<<<grid, THREADS, smem, st>>>(p, cfg);
        return 0;
    };
    int rc;
    if (p.mode == RQ_MODE_ROTATION_TRICK) rc = launch(rq_bwd_tma_kernel<D, RQ_MODE_ROTATION_TRICK>);
    else if (p.mode == RQ_MODE_STE) rc = launch(rq_bwd_tma_kernel<D, RQ_MODE_STE>);
    else rc = launch(rq_bwd_tma_kernel<D, RQ_MODE_EVAL>);
    if (rc) return rc;
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
    // accumulate: g_cb already holds what the codebook gradient is added to (zeros the caller wrote earlier, or a
    // running gradient) — no memset between the forward and this kernel, so the two stay adjacent for PDL
    if (!p.accumulate) RQ_CUDA(cudaMemsetAsync(p.g_cb, 0, (size_t)p.L * p.K * D * sizeof(float), st));
    if (p.B == 0) return 0;
    {
        bool served = false;
        int rc = 0;
        if (D == 32) rc = launch_bwd_row3(p, st, &served);
        if (!rc && !served && D == 32) rc = launch_bwd_row(p, st, &served);
        if (!rc && !served && D == 32) rc = launch_bwd_tma<32>(p, st, &served);
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
