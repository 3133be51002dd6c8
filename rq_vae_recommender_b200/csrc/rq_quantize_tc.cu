// tcgen05 family: fused L-level residual quantiser on the 5th-gen tensor cores (sm_100a).
//
// One persistent CTA per SM, NWG (4, or 2 when the operand tiles of L levels leave no room) 128-thread
// warpgroups.  A warpgroup owns one 128-row tile at a time (UMMA M = 128; one thread = one row = one
// TMEM lane) and walks its L levels with the residual in registers.  TMEM holds two 128x256 fp32
// accumulators; warpgroups w and w+2 take turns on accumulator w&1 (a shared-memory try-lock), so
// while one of them is in its MMA + scan phase the other does its CUDA-core epilogue.
//
//   prologue   every level's codebook is converted once per CTA into the UMMA B operand
//              [-2*hi(c) | -2*lo(c)] (bf16 split, 128-byte swizzled rows, K-major) plus a small
//              [hi, mid, lo] split of ||c||^2 (no-swizzle tile), kept in shared memory.
//   per level  each thread writes [hi(r) | lo(r)] of its residual as the A operand row; one elected
//              thread issues 7 tcgen05.mma (N=256, K=16 each):
//                 hi(r).hi(c) + lo(r).hi(c) + hi(r).lo(c)  (+ 1 . ||c||^2)
//              so the accumulator holds  ||c||^2 - 2 r.c  to ~2^-17 relative.
//   scan       ONE pass over the TMEM row keeps two families of running minima: per 16-column group
//              (A, 16 of them) and per column-mod-16 class (B, 16 of them) — one FMNMX3 per two values
//              each.  Any two distinct columns differ in group or in class, so "a second value within
//              the error bound of the minimum" shows up as a close runner-up in A or in B; when neither
//              has one (the usual case) argmin = (argmin group, argmin class).  Otherwise the warp
//              re-reads the row, flags every code within the bound and re-scores those in fp32 with
//              exactly the arithmetic of the SIMT family, lowest index winning — so ids are
//              bit-identical to the SIMT kernel.
//   epilogue   gather the fp32 codeword, estimator (STE / rotation trick / eval), loss, outputs,
//              residual update in registers — the next level never touches HBM.
#include <stdlib.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"
#include "rq_tc_ptx.cuh"

namespace rq {

namespace tc {
constexpr int D = 32;            // latent dim served by this kernel
constexpr int KC = 256;          // codes per level (UMMA N)
constexpr int TILE_M = 128;      // rows per tile (UMMA M)
constexpr int MAX_L = 4;
constexpr int B_TILE_BYTES = KC * 128;       // [256 codes][-2hi(32) | -2lo(32)] bf16, SW128
constexpr int CC_TILE_BYTES = KC * 32;       // [256 codes][16 bf16], no swizzle, 8x16B core matrices
constexpr int A_TILE_BYTES = TILE_M * 128;   // [128 rows][hi(32) | lo(32)] bf16, SW128
constexpr int ONES_TILE_BYTES = TILE_M * 32; // [128 rows][1,1,1,0...] bf16, no swizzle

struct SmemLayout {
    int b_tiles, cc_tiles, cc_f32, a_tiles, park, ones, misc, total;
};
// nwg == 4: one A tile per accumulator (the two warpgroups sharing it take turns) and a 16 KB fp32
// parking area per warpgroup for the residual rows while the scan needs the registers.
__host__ __device__ inline SmemLayout smem_layout(int L, int nwg) {
    SmemLayout s;
    int off = 0;
    s.b_tiles = off; off += L * B_TILE_BYTES;
    s.a_tiles = off; off += 2 * A_TILE_BYTES;
    s.park = off; off += (nwg == 4 ? 4 : 0) * A_TILE_BYTES;
    s.cc_tiles = off; off += L * CC_TILE_BYTES;
    s.ones = off; off += ONES_TILE_BYTES;
    s.cc_f32 = off; off += L * KC * 4;
    s.misc = off; off += 256;
    s.total = off;
    return s;
}

// byte offset of 16-byte chunk `c` of row `r` inside a 128B-swizzled tile with 128-byte rows
__device__ __forceinline__ uint32_t sw128_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}
// byte offset of 16-byte K-chunk `c` (0/1) of row `r` in a no-swizzle K-major tile:
// core matrix = 8 rows x 16 B (128 B); the two K-chunks are adjacent (LBO = 128 B), 8-row groups
// follow at SBO = 256 B.
__device__ __forceinline__ uint32_t nosw_off(int r, int c) {
    return (uint32_t)((r >> 3) * 256 + c * 128 + (r & 7) * 16);
}
}  // namespace tc


struct TcParams {
    FwdParams f;
    int* err;        // device flag: set when an mbarrier / lock wait timed out
    int n_tiles;
    float* dbg_acc;  // optional [128][256] dump of tile 0 / level 0's accumulator (tests only)
    int dbg_flags;      // experiments only (RQ_TC_DBG_FLAGS): 1 = skip first x load, 2 = skip codebook TMA,
                        // 4 = per-CTA %globaltimer stamps at dbg_tl[2048 + 4 * blockIdx.x],
                        // 8 = never take the re-scoring path, 32 = issue 2 of the 7 MMAs (timing what-ifs, wrong ids)
    long long* dbg_tl;  // optional [32 warps][64 marks] clock64 timeline of CTA 0 (profiling only)
};
#define RQ_TL(mark)                                                                           \
    do {                                                                                      \
        if (P.dbg_tl && blockIdx.x == 0 && lane == 0) P.dbg_tl[warp * 64 + (mark)] = clock64(); \
    } while (0)
// same, but the stamp is taken only once `val` (a float) has arrived
#define RQ_TL_DEP(mark, val)                                                                  \
    do {                                                                                      \
        if (P.dbg_tl && blockIdx.x == 0 && lane == 0)                                         \
            P.dbg_tl[warp * 64 + (mark)] = clock64() + (__float_as_int(val) == 0x7fc12345);  \
    } while (0)

namespace tc {

// top-2 of 16 packed keys (value bits with the entry index in the low 4 mantissa bits)
__device__ __forceinline__ void top2_of16(const float (&k)[16], float& m1, float& m2) {
    float lo[8], hi[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        lo[i] = fminf(k[2 * i], k[2 * i + 1]);
        hi[i] = fmaxf(k[2 * i], k[2 * i + 1]);
    }
#pragma unroll
    for (int s = 1; s < 8; s *= 2) {
#pragma unroll
        for (int i = 0; i + s < 8; i += 2 * s) {
            const float a = fmaxf(lo[i], lo[i + s]);
            hi[i] = ptx::fmin3(a, hi[i], hi[i + s]);
            lo[i] = fminf(lo[i], lo[i + s]);
        }
    }
    m1 = lo[0];
    m2 = hi[0];
}
__device__ __forceinline__ float pack_idx(float v, int i) {
    return __uint_as_float((__float_as_uint(v) & 0xFFFFFFF0u) | (uint32_t)i);
}

}  // namespace tc

template <int NWG>
__global__ void __launch_bounds__(128 * NWG, 1) rq_fwd_tc_kernel(TcParams P) {
    using namespace tc;
    constexpr int THREADS = 128 * NWG;
    constexpr bool SHARE = (NWG == 4);      // warpgroups w and w+2 alternate on accumulator / A tile w&1
    const FwdParams& p = P.f;
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled UMMA tiles need 1024-byte alignment; the launch over-allocates by 1 KB
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const SmemLayout S = smem_layout(p.L, NWG);
    uint8_t* sB = smem + S.b_tiles;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sPark = smem + S.park;
    uint8_t* sCCt = smem + S.cc_tiles;
    uint8_t* sOnes = smem + S.ones;
    float* sCC = reinterpret_cast<float*>(smem + S.cc_f32);
    uint64_t* mbar_full = reinterpret_cast<uint64_t*>(smem + S.misc);        // [4] accumulator ready (per warpgroup)
    uint64_t* mbar_free = reinterpret_cast<uint64_t*>(smem + S.misc + 32);   // [2] accumulator read out (per slot)
    uint64_t* mbar_load = reinterpret_cast<uint64_t*>(smem + S.misc + 128);  // [MAX_L] a level's codebook bulk copy landed
    uint64_t* mbar_x = reinterpret_cast<uint64_t*>(smem + S.misc + 96);     // [4] x tile landed (per warpgroup)
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 64);
    int* cmax_bits = reinterpret_cast<int*>(smem + S.misc + 80);             // [MAX_L] max ||c||^2 per level

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wg = tid >> 7, wtid = tid & 127;
    const int slot = wg & 1;
    RQ_TL(0);
    if (P.dbg_tl && (P.dbg_flags & 4) && tid == 0) {       // per-CTA wall-clock stamps (profiling only)
        P.dbg_tl[2048 + 4 * blockIdx.x] = (long long)ptx::globaltimer();
        P.dbg_tl[2048 + 4 * blockIdx.x + 1] = (long long)ptx::smid();
    }

    // ---- codebooks: one TMA bulk copy per level, fp32 rows straight into that level's B-tile space --------
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) ptx::mbar_init(&mbar_full[i], 1);
        ptx::mbar_init(&mbar_free[0], 4);
        ptx::mbar_init(&mbar_free[1], 4);
#pragma unroll
        for (int i = 0; i < MAX_L; ++i) ptx::mbar_init(&mbar_load[i], 1);
#pragma unroll
        for (int i = 0; i < 4; ++i) ptx::mbar_init(&mbar_x[i], 1);
        ptx::fence_mbar_init();
        const uint32_t level_bytes = (uint32_t)p.K * D * 4;
        // one barrier per level: level l is converted while the later levels are still in flight
        for (int l = 0; l < p.L; ++l) {
            ptx::mbar_arrive_expect_tx(&mbar_load[l], (P.dbg_flags & 2) ? 0u : level_bytes);
            if (!(P.dbg_flags & 2)) ptx::bulk_g2s(sB + l * B_TILE_BYTES, p.cb + (int64_t)l * p.K * D, level_bytes, &mbar_load[l]);
        }
    }
    RQ_TL(45);
    // ---- first tile's rows: HBM loads in flight during the whole prologue ------------------------------
    // tile t -> CTA t % grid, warpgroup (t / grid) % NWG: every SM gets floor or ceil of n_tiles / grid
    const int my_tiles = (P.n_tiles > (int)blockIdx.x) ? (P.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    // tiles of this CTA are numbered 0..my_tiles-1; warpgroup w owns those with index % NWG == w
    const int n_mine = (my_tiles > wg) ? (my_tiles - 1 - wg) / NWG + 1 : 0;
    const int n_partner = SHARE ? ((my_tiles > (wg ^ 2)) ? (my_tiles - 1 - (wg ^ 2)) / NWG + 1 : 0) : 0;
    float r[D];
    const int wrow0 = (warp & 3) * 32;      // first tile row of this warp
    uint8_t* sPw = sPark + wg * A_TILE_BYTES;
    // rows of a tile are contiguous in x ([B, D] row-major): one TMA bulk copy per tile into the warpgroup's
    // parking tile (linear layout), issued by one thread, no load instructions in anybody's way
    auto issue_x_tile = [&](int64_t trow0) {
        const int64_t rows = (p.B - trow0 < TILE_M) ? (p.B - trow0) : TILE_M;
        ptx::mbar_arrive_expect_tx(&mbar_x[wg], (uint32_t)rows * D * 4);
        ptx::bulk_g2s(sPw, p.x + trow0 * D, (uint32_t)rows * D * 4, &mbar_x[wg]);
    };
    if (!SHARE) {
        const int64_t row = ((int64_t)blockIdx.x + (int64_t)gridDim.x * wg) * TILE_M + wtid;
        if (n_mine > 0 && row < p.B) {
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + row * D + j, &r[j]);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
    }
    RQ_TL(46);
    if (tid < MAX_L) cmax_bits[tid] = 0;
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, 512);
    RQ_TL(47);
    // ones tile: k = 0,1,2 -> 1.0 (multiplies hi/mid/lo of ||c||^2), rest 0
    for (int i = tid; i < TILE_M * 2; i += THREADS) {
        int rr = i >> 1, c = i & 1;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0) { v.x = 0x3F803F80u; v.y = 0x00003F80u; }
        *reinterpret_cast<uint4*>(sOnes + nosw_off(rr, c)) = v;
    }
    RQ_TL(48);
    __syncthreads();                                   // barrier inits + cmax zeros visible
    RQ_TL(1);
    if (SHARE && wtid == 0 && n_mine > 0) issue_x_tile(((int64_t)blockIdx.x + (int64_t)gridDim.x * wg) * TILE_M);
    RQ_TL(2);
    // ---- in-place conversion: work item = (level, code, 8-element chunk); the four lanes of a code read
    // its 128-byte fp32 row and rewrite the same 128 bytes as [-2hi | -2lo] bf16 in the swizzled order ------
    {
        const int n_items = p.L * KC * 4;                          // multiple of 2 * THREADS
        constexpr int U = 2;                                       // items in flight per thread
        for (int i0 = tid; i0 < n_items; i0 += U * THREADS) {
            // items i0 and i0 + THREADS belong to the same level (1024 items per level, U * THREADS divides 1024)
            if (!ptx::mbar_wait(&mbar_load[i0 >> 10], 0)) {
                if (P.err) atomicExch(P.err, 1);
            }
            float cv[U][8];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * THREADS;
                const int l = i >> 10, k = (i >> 2) & 255, q = i & 3;
                const uint8_t* src = sB + l * B_TILE_BYTES + k * 128 + q * 32;
                if (k < p.K) {
                    const float4 v0 = *reinterpret_cast<const float4*>(src);
                    const float4 v1 = *reinterpret_cast<const float4*>(src + 16);
                    cv[u][0] = v0.x; cv[u][1] = v0.y; cv[u][2] = v0.z; cv[u][3] = v0.w;
                    cv[u][4] = v1.x; cv[u][5] = v1.y; cv[u][6] = v1.z; cv[u][7] = v1.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) cv[u][j] = 0.f;
                }
            }
            uint32_t h[U][4], lw[U][4];
            float part[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                part[u] = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float a = cv[u][2 * j], b = cv[u][2 * j + 1];
                    part[u] = fmaf(a, a, part[u]);
                    part[u] = fmaf(b, b, part[u]);
                    ptx::split_pair(-2.f * a, -2.f * b, h[u][j], lw[u][j]);
                }
            }
            // ||c||^2 in the canonical order shared with the SIMT family (code_sqnorm in rq_rowmath.cuh):
            // 8-wide sequential partials, then (p0+p1)+(p2+p3)
#pragma unroll
            for (int u = 0; u < U; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], 1);
#pragma unroll
            for (int u = 0; u < U; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], 2);
            __syncwarp();                                           // the quad's reads precede its writes
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * THREADS;
                const int l = i >> 10, k = (i >> 2) & 255, q = i & 3;
                uint8_t* tileB = sB + l * B_TILE_BYTES;
                *reinterpret_cast<uint4*>(tileB + sw128_off(k, q)) = make_uint4(h[u][0], h[u][1], h[u][2], h[u][3]);
                *reinterpret_cast<uint4*>(tileB + sw128_off(k, 4 + q)) = make_uint4(lw[u][0], lw[u][1], lw[u][2], lw[u][3]);
                if (q == 0) {
                    float cc = (k < p.K) ? part[u] : 1e30f;        // padding codes can never be the minimum
                    sCC[l * KC + k] = cc;
                    float hh = ptx::bf16_round(cc), mm = ptx::bf16_round(cc - hh), ll = ptx::bf16_round((cc - hh) - mm);
                    *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 0)) =
                        make_uint4(ptx::pack_bf16x2(hh, mm), ptx::pack_bf16x2(ll, 0.f), 0, 0);
                    *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 1)) = make_uint4(0, 0, 0, 0);
                }
            }
        }
    }
    RQ_TL(3);
    ptx::fence_proxy_async_smem();
    __syncthreads();
    if (warp < p.L) {                                  // max ||c||^2 per level (the scan's error bound)
        float m = 0.f;
        for (int k = lane; k < p.K; k += 32) m = fmaxf(m, sCC[warp * KC + k]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) cmax_bits[warp] = __float_as_int(m);
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    RQ_TL(4);
    const uint32_t tmem_base = *tmem_base_s;
    const uint32_t tmem_acc = tmem_base + (uint32_t)(slot * KC);                     // column offset
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);        // this warp's lanes
    uint8_t* sAw = sA + slot * A_TILE_BYTES;
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TILE_M, KC);
    bool timed_out = false;
    // Global <-> register traffic of a warp's 32 rows goes through its 4 KB of the parking tile so that
    // every global instruction covers 4 whole rows (4 lines) instead of one sector of 32 different lines.
    auto park_store_row = [&](const float (&v)[D]) {              // my row -> parking tile (swizzled, conflict-free)
#pragma unroll
        for (int c = 0; c < 8; ++c)
            *reinterpret_cast<float4*>(sPw + sw128_off(wtid, c)) =
                make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
    };
    auto park_load_row = [&](float (&v)[D]) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const float4 t = *reinterpret_cast<const float4*>(sPw + sw128_off(wtid, c));
            v[4 * c] = t.x; v[4 * c + 1] = t.y; v[4 * c + 2] = t.z; v[4 * c + 3] = t.w;
        }
    };
    // this lane's (row, chunk) in the 4-rows-per-instruction coalesced pattern: row = wrow0 + 4 i + lane/8
    const uint32_t co_row = (uint32_t)(wrow0 + (lane >> 3)), co_sw = sw128_off((int)co_row, lane & 7);
    auto co_off = [&](int i) -> uint32_t { return (co_sw + 512u * i) ^ ((i & 1) ? 64u : 0u); };
    auto park_copy_out = [&](float* dst_tile, int64_t trow0) {     // parking tile -> 4 KB of global, coalesced
        float* dst = dst_tile + (int64_t)co_row * D + (lane & 7) * 4;
        // rows 4 i further down keep (row & 7) parity only for even i: 4 i rows = 512 i bytes, the swizzle term
        // flips bit 2 of the chunk index for odd i
        if (trow0 + TILE_M <= p.B) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 t = *reinterpret_cast<const float4*>(sPw + co_off(i));
                ptx::st_global_v4(dst + i * 4 * D, t);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 t = *reinterpret_cast<const float4*>(sPw + co_off(i));
                if (trow0 + co_row + 4 * i < p.B) ptx::st_global_v4(dst + i * 4 * D, t);
            }
        }
    };
    auto take_x_tile = [&](int it, bool active) {                  // linear parking tile -> my row
        if (!ptx::mbar_wait(&mbar_x[wg], it & 1)) timed_out = true;
        RQ_TL(49);
        // linear -> swizzled in place (conflict-free both ways: 8 lanes cover one 128-byte row), then my row
        float4 t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
            t[i] = *reinterpret_cast<const float4*>(sPw + (wrow0 + 4 * i + (lane >> 3)) * 128 + (lane & 7) * 16);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(sPw + co_off(i)) = t[i];
        __syncwarp();
        if (active) {
            park_load_row(r);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
        __syncwarp();
    };
    auto store_id = [&](int l, int64_t row, bool active, int bi) {
        int64_t* dst = p.ids + (int64_t)l * p.B + row;
        if (p.n_changed) {                                  // k-means: count changed assignments (warp-aggregated)
            const bool ch = active && (!p.has_prev || *dst != (int64_t)bi);
            const unsigned m = __ballot_sync(0xffffffffu, ch);
            if (lane == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
        }
        if (active) *dst = bi;
    };
    const int n_partner_uses = n_partner * p.L;
    const bool is_hi = SHARE && (wg >= 2);
    int use = 0;                                   // this warpgroup's running (tile, level) count

    for (int it = 0; it < n_mine; ++it) {
        const int tile = (int)blockIdx.x + (int)gridDim.x * (wg + NWG * it);
        const int64_t row = (int64_t)tile * TILE_M + wtid;
        const bool active = row < p.B;
        if (SHARE) take_x_tile(it, active);
        float xx = dot_self<D>(r);
        RQ_TL_DEP(50, xx);
        float loss_acc = 0.f;

        for (int l = 0; l < p.L; ++l, ++use) {
            // The use of (accumulator, A tile) `slot` that precedes this one is the partner's: its use
            // `use` when we are the second warpgroup of the pair, `use - 1` when we are the first.
            const int prev = is_hi ? use : use - 1;
            const bool wait_partner = SHARE && prev >= 0 && prev < n_partner_uses;
            // its MMAs must have finished reading the A tile before we overwrite it
            if (wait_partner && !ptx::mbar_wait(&mbar_full[wg ^ 2], prev & 1)) timed_out = true;
            // ---- A operand row: [hi(r) | lo(r)]; the exact row is parked in shared memory ----------------
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t h[4], lw[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) ptx::split_pair(r[c * 8 + 2 * q], r[c * 8 + 2 * q + 1], h[q], lw[q]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, c)) = make_uint4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, 4 + c)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            ptx::fence_proxy_async_smem();
            RQ_TL_DEP(5 + 8 * l, xx);
            ptx::named_bar_sync(1 + wg, 128);
            if (wtid == 0) {
                // the partner must have read its accumulator out: completion #(2*use - 1) resp. #(2*use) of
                // the slot's read-out barrier, parity 1 resp. 0
                if (wait_partner && !ptx::mbar_wait(&mbar_free[slot], is_hi ? 0 : 1)) timed_out = true;
                ptx::tc_fence_after_sync();
                const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sB + l * B_TILE_BYTES);
                // (A k-offset bytes, B k-offset bytes): hi.hi, hi.hi, lo.hi, lo.hi, hi.lo, hi.lo
                const int ak[6] = {0, 32, 64, 96, 0, 32};
                const int bk[6] = {0, 32, 0, 32, 64, 96};
#pragma unroll
                for (int i = 0; i < ((P.dbg_flags & 32) ? 1 : 6); ++i) {
                    uint64_t ad = ptx::make_smem_desc(a0 + ak[i], 16, 1024, ptx::kLayoutSw128);
                    uint64_t bd = ptx::make_smem_desc(b0 + bk[i], 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);
                }
                uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
                uint64_t cd = ptx::make_smem_desc(ptx::smem_u32(sCCt + l * CC_TILE_BYTES), 128, 256, ptx::kLayoutNone);
                ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
                ptx::umma_commit(&mbar_full[wg]);
                RQ_TL(6 + 8 * l);
            }
            // ---- while the MMAs run: this level's input residual goes out (HBM) and to the parking rows ----
            const int64_t off = ((int64_t)l * p.B + row) * D;
            const int64_t trow0 = (int64_t)tile * TILE_M;
            if (SHARE) {
                park_store_row(r);
                if (p.res) {
                    __syncwarp();
                    park_copy_out(p.res + ((int64_t)l * p.B + trow0) * D, trow0);
                }
            } else if (active && p.res) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.res + off + j, &r[j]);
            }
            const float cmax2 = __int_as_float(cmax_bits[l]);
            // |acc - exact| <= 2*(2^-18 + 2*2^-17)|r||c| = 3.8e-5 |r||c| (dropped lo.lo term + the two
            // truncated lo parts) + fp32 accumulation noise; both the winner and a rival may be off,
            // hence 2x, with margin ("error bound")
            const float tol = 1e-4f * sqrtf(xx * cmax2) + 3e-6f * cmax2 + 1e-30f;
            if (!ptx::mbar_wait(&mbar_full[wg], use & 1)) timed_out = true;
            ptx::tc_fence_after_sync();
            RQ_TL(7 + 8 * l);

            if (P.dbg_acc && tile == 0 && l == 0) {
#pragma unroll 1
                for (int c = 0; c < KC / 32; ++c) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32b_x32(tmem_row + c * 32, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) P.dbg_acc[wtid * KC + c * 32 + j] = __uint_as_float(v[j]);
                }
            }
            // ---- scan: group minima (A: columns 16g..16g+15) and class minima (B: column mod 16) -----------
            // The TMEM loads are double-buffered: chunk c+1 is in flight while chunk c is reduced.
            float accA[16], accB[16];
            {
                uint32_t va[32], vb[32];
                auto reduce_chunk = [&](const uint32_t (&v)[32], int c) {
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float* g = f + 16 * h;
                        float a0 = ptx::fmin3(g[0], g[1], g[2]), a1 = ptx::fmin3(g[3], g[4], g[5]);
                        float a2 = ptx::fmin3(g[6], g[7], g[8]), a3 = ptx::fmin3(g[9], g[10], g[11]);
                        float a4 = ptx::fmin3(g[12], g[13], g[14]);
                        a0 = ptx::fmin3(a0, a1, a2);
                        a3 = ptx::fmin3(a3, a4, g[15]);
                        accA[2 * c + h] = fminf(a0, a3);
                    }
                    if (c == 0) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) accB[i] = fminf(f[i], f[16 + i]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) accB[i] = ptx::fmin3(accB[i], f[i], f[16 + i]);
                    }
                };
                ptx::tmem_ld_32x32b_x32(tmem_row, va);
                ptx::tmem_ld_32x32b_x32(tmem_row + 32, vb);
                ptx::tmem_ld_wait_for2(va, vb);
#pragma unroll
                for (int c = 0; c < KC / 32; c += 2) {
                    reduce_chunk(va, c);
                    if (c + 2 < KC / 32) ptx::tmem_ld_32x32b_x32(tmem_row + (c + 2) * 32, va);
                    reduce_chunk(vb, c + 1);
                    if (c + 2 < KC / 32) {
                        ptx::tmem_ld_32x32b_x32(tmem_row + (c + 3) * 32, vb);
                        ptx::tmem_ld_wait_for2(va, vb);
                    }
                }
            }
            RQ_TL_DEP(8 + 8 * l, accB[0]);
            // ---- the accumulator is read out: hand it back (one arrival per warp) --------------------------------
            ptx::tc_fence_before_sync();            // orders this thread's tcgen05.ld before the arrival / next barrier
            if (SHARE) {
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&mbar_free[slot]);
            }
            float m1a, m2a, m1b, m2b;
            {
                float k[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) k[i] = pack_idx(accA[i], i);
                top2_of16(k, m1a, m2a);
#pragma unroll
                for (int i = 0; i < 16; ++i) k[i] = pack_idx(accB[i], i);
                top2_of16(k, m1b, m2b);
            }
            int bi = (int)(((__float_as_uint(m1a) & 15u) << 4) | (__float_as_uint(m1b) & 15u));
            // packing the index moves a value by < 2^-19 |v|: widen the band by 2^-17 |m1|
            const float m1 = fminf(m1a, m1b);
            const float thr = m1 + (tol + 8e-6f * fabsf(m1));
            const bool ambiguous = ((m2a <= thr) || (m2b <= thr)) && !(P.dbg_flags & 8);
            const float* cbl = p.cb + (int64_t)l * p.K * D;
            // the exact residual row comes back from its parking place
            if (SHARE) park_load_row(r);
            if (__any_sync(0xffffffffu, ambiguous)) {
                // ---- rare: every code within the band lies in a group with accA <= thr AND a class with
                // accB <= thr; re-score those (group, class) pairs in fp32, ascending index, strict '<':
                // exactly the SIMT family's arithmetic.  Extra pairs are harmless (their exact distance loses).
                uint32_t gm = 0, cm = 0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    gm |= (accA[i] <= thr) ? (1u << i) : 0u;
                    cm |= (accB[i] <= thr) ? (1u << i) : 0u;
                }
                if (ambiguous) {
                    float best = INFINITY;
#pragma unroll 1
                    for (uint32_t g = gm; g; g &= g - 1) {
                        const int gi = __ffs(g) - 1;
#pragma unroll 1
                        for (uint32_t c = cm; c; c &= c - 1) {
                            const int k = gi * 16 + (__ffs(c) - 1);
                            if (k >= p.K) continue;
                            const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)k * D);
                            float4 cv[D / 4];
#pragma unroll
                            for (int j = 0; j < D / 4; ++j) cv[j] = __ldg(c4 + j);
                            float a = 0.f;
#pragma unroll
                            for (int j = 0; j < D / 4; ++j) {
                                a = fmaf(r[4 * j], cv[j].x, a); a = fmaf(r[4 * j + 1], cv[j].y, a);
                                a = fmaf(r[4 * j + 2], cv[j].z, a); a = fmaf(r[4 * j + 3], cv[j].w, a);
                            }
                            float d = (xx + sCC[l * KC + k]) - 2.f * a;
                            if (d < best) { best = d; bi = k; }
                        }
                    }
                }
            }
            RQ_TL(9 + 8 * l);
            if (bi >= p.K) bi = p.K - 1;        // unreachable unless inputs are NaN; keeps the gather in range

            // ---- gather, estimator, loss, outputs, residual update ---------------------------------------
            // ids-only launches (tokenisation, k-means assignment) need nothing past the id at the last level
            if (l == p.L - 1 && !p.emb && !p.loss) {
                store_id(l, row, active, bi);
                continue;
            }
            float e[D], o[D];
            if (SHARE) {
                // cooperative gather: 8 lanes fetch the 8 chunks of one row's codeword (4 rows per instruction)
                float4 g[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = __shfl_sync(0xffffffffu, bi, 4 * i + (lane >> 3));
                    g[i] = __ldg(reinterpret_cast<const float4*>(cbl + (int64_t)k * D) + (lane & 7));
                }
                __syncwarp();                       // everybody has re-read its parked residual
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<float4*>(sPw + co_off(i)) = g[i];
                __syncwarp();
                park_load_row(e);
            } else {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::ld_global_v8(cbl + (int64_t)bi * D + j, &e[j]);
            }
            RQ_TL_DEP(10 + 8 * l, e[31]);
            loss_acc += level_epilogue<D>(p.mode, p.beta, r, e, xx, o);
            RQ_TL_DEP(11 + 8 * l, o[31]);
            if (SHARE) {
                if (p.emb) {
                    __syncwarp();                   // everybody has read its codeword row
                    park_store_row(o);
                    __syncwarp();
                    park_copy_out(p.emb + ((int64_t)l * p.B + trow0) * D, trow0);
                    __syncwarp();                   // copied out before the next level parks its residual
                }
                store_id(l, row, active, bi);
            } else {
                if (active && p.emb) {
#pragma unroll
                    for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.emb + off + j, &o[j]);
                }
                store_id(l, row, active, bi);
            }
            residual_update<D>(r, o);                                      // rqvae.py:129
            xx = dot_self<D>(r);
            RQ_TL_DEP(12 + 8 * l, xx);
        }
        if (active && p.loss) p.loss[row] = loss_acc;
        // next tile of this warpgroup
        if (it + 1 < n_mine) {
            const int64_t ntrow0 = ((int64_t)blockIdx.x + (int64_t)gridDim.x * (wg + NWG * (it + 1))) * TILE_M;
            if (SHARE) {
                ptx::fence_proxy_async_smem();          // our generic accesses to the parking tile, then the TMA write
                ptx::named_bar_sync(1 + wg, 128);
                if (wtid == 0) issue_x_tile(ntrow0);
            } else if (ntrow0 + wtid < p.B) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + (ntrow0 + wtid) * D + j, &r[j]);
            } else {
#pragma unroll
                for (int j = 0; j < D; ++j) r[j] = 0.f;
            }
        }
    }

    if (timed_out && P.err) atomicExch(P.err, 1);
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
    if (P.dbg_tl && (P.dbg_flags & 4) && tid == 0) P.dbg_tl[2048 + 4 * blockIdx.x + 2] = (long long)ptx::globaltimer();
}

bool tc32_supported(int D, int K, int L, int mode, int distance) {
    if (getenv("RQ_DISABLE_TCGEN05")) return false;
    return D == tc::D && K >= 1 && K <= tc::KC && L >= 1 && L <= tc::MAX_L && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

static int* d_err = nullptr;
static float* g_dbg_acc = nullptr;
static long long* g_dbg_tl = nullptr;
constexpr int kMaxSmem = 227 * 1024;

template <int NWG>
static int launch_tc(const TcParams& P, cudaStream_t st) {
    const tc::SmemLayout S = tc::smem_layout(P.f.L, NWG);
    static bool attr_set = false;
    if (!attr_set) {
        RQ_CUDA(cudaFuncSetAttribute(rq_fwd_tc_kernel<NWG>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
        attr_set = true;
    }
    int grid = P.n_tiles < sm_count() ? P.n_tiles : sm_count();
    rq_fwd_tc_kernel<NWG><<<grid, 128 * NWG, S.total + 1024, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

bool tc_supported(int D, int K, int L, int mode, int distance) {
    return tc32_supported(D, K, L, mode, distance) || tc64_supported(D, K, L, mode, distance);
}
int quantize_forward_tc(const FwdParams& p, int D, cudaStream_t st) {
    return D == 64 ? quantize_forward_tc64(p, st) : quantize_forward_tc32(p, st);
}

int quantize_forward_tc32(const FwdParams& p, cudaStream_t st) {
    if (p.B == 0) return 0;
    if (!d_err) {
        RQ_CUDA(cudaMalloc((void**)&d_err, sizeof(int)));
        RQ_CUDA(cudaMemset(d_err, 0, sizeof(int)));
    }
    TcParams P;
    P.f = p;
    P.err = d_err;
    P.dbg_acc = g_dbg_acc;
    P.dbg_tl = g_dbg_tl;
    P.dbg_flags = getenv("RQ_TC_DBG_FLAGS") ? atoi(getenv("RQ_TC_DBG_FLAGS")) : 0;
    P.n_tiles = (int)((p.B + tc::TILE_M - 1) / tc::TILE_M);
    // four warpgroups when the operand tiles of L levels + four A tiles fit in shared memory (L <= 3)
    const bool four = tc::smem_layout(p.L, 4).total + 1024 <= kMaxSmem && !getenv("RQ_TC_TWO_WARPGROUPS");
    return four ? launch_tc<4>(P, st) : launch_tc<2>(P, st);
}

}  // namespace rq

// Test hooks (declared in include/rq_engine.h): the mbarrier-timeout flag and an accumulator dump.
extern "C" int rq_debug_tc_timeout_flag(void) {
    int v = 0;
    if (rq::d_err) cudaMemcpy(&v, rq::d_err, sizeof(int), cudaMemcpyDeviceToHost);
    return v | rq::tc64_timeout_flag();
}
extern "C" int rq_debug_tc_set_accumulator_dump(float* device_buf_128x256) {
    rq::g_dbg_acc = device_buf_128x256;
    return 0;
}
extern "C" int rq_debug_tc_set_timeline(long long* device_buf_32x64) {
    rq::g_dbg_tl = device_buf_32x64;
    return 0;
}
