// tcgen05 family, D = 32, K <= 256, L <= 3: fused L-level residual quantiser on the 5th-gen tensor cores (sm_100a).
//
// One persistent CTA per SM, four 128-thread warpgroups.  A warpgroup owns one 128-row tile at a time (UMMA M = 128)
// and walks its L levels; TMEM holds two 128x256 fp32 accumulators, warpgroups w and w+2 alternate strictly on
// accumulator / A-operand tile w&1, so while one of them is in its MMA + scan phase the other does its CUDA-core work.
//
// Two thread <-> data layouts alternate inside a warp, which always serves the same 32 rows:
//   * scan     thread = row (the TMEM lane): ONE pass over the 256 accumulator columns keeps 16 group minima and 16
//              class minima (column / 16, column % 16); any two distinct columns differ in group or class, so a second
//              value within the error bound of the minimum shows as a close runner-up among the groups or among the
//              classes.  Without one (the usual case) argmin = (argmin group, argmin class); otherwise the codes of
//              every (group, class) pair within the bound are re-scored in fp32 with exactly the SIMT family's
//              arithmetic, lowest index winning — ids are bit-identical to the exact fp32 kernel.
//   * row math 4 lanes = one row, 8 consecutive floats per lane: every global instruction covers 8 whole rows (1 KB
//              contiguous), dot products finish with two shfl.xor (canonical chunk order of rq_rowmath.cuh), nothing
//              bounces through shared memory.  The residual row lives in a 128-byte-swizzled fp32 tile (conflict-free
//              for both layouts) between levels; ids / ||r||^2 cross between the layouts by warp shuffles.
// Per level a warp does: codeword gather -> estimator (STE / rotation trick / eval) -> next residual -> bf16 hi/lo
// split straight into the A tile -> MMAs issued; the output stores (embedding, residual, loss) and ||r||^2 of the next
// level run in the shadow of those MMAs.
//
// Prologue: only level 0's codebook is converted by all warps before the first MMA ([-2hi(c) | -2lo(c)] bf16 split in
// place over the fp32 rows a TMA bulk copy dropped into the B-tile space, plus a [hi, mid, lo] split of ||c||^2 that a
// K = 16 MMA against a ones tile adds to the accumulator); the later levels are converted by warpgroups 2 and 3 while
// they wait for their first turn on the accumulators, and published through per-level mbarriers.
#include <stdlib.h>

#include "rq_rowmath.cuh"
#include "rq_tc_tiles.cuh"

namespace rq {

namespace tc32 {
using namespace tc;
constexpr int NWG = 4;
constexpr int THREADS = 128 * NWG;
constexpr int MAX_L = 3;
constexpr int R_TILE_BYTES = TILE_M * 128;   // [128 rows][32 fp32], SW128

struct SmemLayout {
    int b_tiles, a_tiles, r_tiles, cc_tiles, ones, cc_f32, misc, km, total;
};
// km: room for the k-means partial sums of the fused assignment + accumulation launch (L = 1): two int32 pieces per
// (cluster, component) + a count per cluster
__host__ __device__ inline SmemLayout smem_layout(int L, bool km = false) {
    SmemLayout s;
    int off = 0;
    s.b_tiles = off; off += L * B_TILE_BYTES;
    s.a_tiles = off; off += 2 * A_TILE_BYTES;
    s.r_tiles = off; off += NWG * R_TILE_BYTES;
    s.cc_tiles = off; off += L * CC_TILE_BYTES;
    s.ones = off; off += ONES_TILE_BYTES;
    s.cc_f32 = off; off += L * KC * 4;
    s.misc = off; off += 256;
    s.km = off; off += km ? (2 * KC * D + KC) * 4 : 0;
    s.total = off;
    return s;
}
}  // namespace tc32

#ifdef RQ_EXPERIMENTS
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
// per-CTA wall-clock stamps (%globaltimer at entry / exit, %smid) at dbg_tl[2048 + 4 * blockIdx.x]
#define RQ_CTA_STAMP(slot_, val_)                                                                \
    do {                                                                                        \
        if (P.dbg_tl && threadIdx.x == 0) P.dbg_tl[2048 + 4 * blockIdx.x + (slot_)] = (long long)(val_); \
    } while (0)
#define RQ_WHATIF(bit) (P.dbg_flags & (bit))
#else
#define RQ_WHATIF(bit) false
#define RQ_TL(mark) do {} while (0)
#define RQ_TL_DEP(mark, val) do {} while (0)
#define RQ_CTA_STAMP(slot_, val_) do {} while (0)
#endif

template <int MODE, bool KM = false>
__global__ void __launch_bounds__(tc32::THREADS, 1) rq_fwd_tc32_kernel(TcParams P) {
    using namespace tc32;
    const FwdParams& p = P.f;
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled UMMA tiles need 1024-byte alignment; the launch over-allocates by 1 KB
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const SmemLayout S = smem_layout(p.L, KM);
    uint8_t* sB = smem + S.b_tiles;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sR = smem + S.r_tiles;
    uint8_t* sCCt = smem + S.cc_tiles;
    uint8_t* sOnes = smem + S.ones;
    float* sCC = reinterpret_cast<float*>(smem + S.cc_f32);
    uint64_t* mbar_full = reinterpret_cast<uint64_t*>(smem + S.misc);          // [4] accumulator ready (per warpgroup)
    uint64_t* mbar_free = reinterpret_cast<uint64_t*>(smem + S.misc + 32);     // [2] accumulator read out (per slot)
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 64);
    int* cmax_bits = reinterpret_cast<int*>(smem + S.misc + 80);               // [MAX_L] max ||c||^2 per level
    uint64_t* mbar_load = reinterpret_cast<uint64_t*>(smem + S.misc + 96);     // [MAX_L] a level's bulk copy landed
    uint64_t* mbar_bready = reinterpret_cast<uint64_t*>(smem + S.misc + 128);  // [MAX_L] a level's operand tiles converted
    uint64_t* mbar_tmem = reinterpret_cast<uint64_t*>(smem + S.misc + 160);    // TMEM allocated, address published

    // the warp index is made provably warp-uniform (shuffle from lane 0): everything derived from it — warpgroup, slot,
    // tile numbers, row offsets, output base pointers — then lives in uniform registers instead of the 128 per thread
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int wg = warp >> 2, wtid = (warp & 3) * 32 + lane;
    const int slot = wg & 1;
    const bool is_hi = wg >= 2;
    const int wrow0 = (warp & 3) * 32;          // first tile row of this warp
    uint8_t* sRw = sR + wg * R_TILE_BYTES;      // this warpgroup's residual tile
    uint8_t* sAw = sA + slot * A_TILE_BYTES;    // the A tile of this warpgroup's accumulator slot
    bool timed_out = false;
    RQ_TL(0);
    RQ_CTA_STAMP(0, ptx::globaltimer());
    RQ_CTA_STAMP(1, ptx::smid());

    // tile t -> CTA t % grid; the CTA's tiles are numbered 0..my_tiles-1, warpgroup w owns those with index % 4 == w
    const int my_tiles = (P.n_tiles > (int)blockIdx.x) ? (P.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int n_mine = (my_tiles > wg) ? (my_tiles - 1 - wg) / NWG + 1 : 0;
    const int n_partner = (my_tiles > (wg ^ 2)) ? (my_tiles - 1 - (wg ^ 2)) / NWG + 1 : 0;
    auto tile_of = [&](int it) { return (int)blockIdx.x + (int)gridDim.x * (wg + NWG * it); };

    // ---- x rows of a tile: 16-byte cp.async straight into the swizzled residual tile (512 contiguous bytes per
    // instruction).  Each warp fetches (and later reads) only its own 32 rows.  The kernel serves whole tiles only: the
    // host hands the last B % 128 rows to the exact family, so nothing in here is predicated on the row index. ---------
    auto prefetch_x = [&](int tile) {
        const float* src = p.x + ((int64_t)tile * TILE_M + wrow0) * D + lane * 4;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            ptx::cp_async16(sRw + sw128_off(wrow0 + 4 * j + (lane >> 3), lane & 7), src + j * 4 * D, 16u);
        ptx::cp_async_commit();
    };
    // ---- prologue.  Warpgroups 0 and 1 own the first turns on the two accumulators: their x tiles are requested
    // first thing (HBM latency + a 64 KB-per-SM burst are the long pole of the first MMA), while warpgroups 2 and 3 —
    // which cannot start before the first scans are done — set the CTA up: barriers, codebook bulk copies (one per level,
    // fp32 rows straight into that level's B-tile space), ones tile, TMEM.  pdl_wait(): with programmatic dependent
    // launch this CTA is scheduled as soon as the SM's previous CTA retires; the preceding kernel's writes (x,
    // codebooks) are visible once it returns.  Nobody waits for the TMEM allocation before it needs the address. ---------
    pdl_launch_dependents();
    if (tid == 2 * 128) {
#pragma unroll
        for (int i = 0; i < 4; ++i) ptx::mbar_init(&mbar_full[i], 1);
        ptx::mbar_init(&mbar_free[0], 4);
        ptx::mbar_init(&mbar_free[1], 4);
#pragma unroll
        for (int i = 0; i < MAX_L; ++i) {
            ptx::mbar_init(&mbar_load[i], 1);
            ptx::mbar_init(&mbar_bready[i], i == 0 ? 16 : 8);       // warps that convert the level
        }
        ptx::mbar_init(mbar_tmem, 1);
        ptx::fence_mbar_init();
    }
    pdl_wait();
    if (!is_hi) {
        if (n_mine > 0) prefetch_x(tile_of(0));
    } else {
        if (tid == 2 * 128) {
            const uint32_t level_bytes = (uint32_t)p.K * D * 4;
            for (int l = 0; l < p.L; ++l) {
                ptx::mbar_arrive_expect_tx(&mbar_load[l], level_bytes);
                ptx::bulk_g2s(sB + l * B_TILE_BYTES, p.cb + (int64_t)l * p.K * D, level_bytes, &mbar_load[l]);
            }
        }
        if (wtid < MAX_L && wg == 2) cmax_bits[wtid] = 0;
        // ones tile: k = 0,1,2 -> 1.0 (multiplies hi/mid/lo of ||c||^2), rest 0
        {
            const int rr = (tid - 256) >> 1, c = tid & 1;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (c == 0) { v.x = 0x3F803F80u; v.y = 0x00003F80u; }
            *reinterpret_cast<uint4*>(sOnes + nosw_off(rr, c)) = v;
        }
        ptx::fence_proxy_async_smem();
    }
    if (KM) {                                          // k-means partials start at zero
        int* km = reinterpret_cast<int*>(smem + S.km);
        for (int i = tid; i < 2 * KC * D + KC; i += THREADS) km[i] = 0;
    }
    __syncthreads();                                   // barrier inits, cmax zeros, ones tile
    if (warp == 9) {
        ptx::tmem_alloc(tmem_base_s, 512);
        ptx::tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(mbar_tmem);
    }
    RQ_TL(1);

    // ---- operand conversion: level 0 by everybody, the later levels by warpgroups 2 and 3 (they cannot start before
    // warpgroups 0 and 1 have scanned their first accumulator).  A level is published on mbar_bready[l]. --------------
    auto convert_level = [&](int l, int first, int stride, int n_iter) {
        if (!ptx::mbar_wait(&mbar_load[l], 0)) timed_out = true;
        float mx = 0.f;
        for (int i = 0; i < n_iter; ++i)
            mx = fmaxf(mx, convert_codebook_items<2>(sB + l * B_TILE_BYTES, sCCt + l * CC_TILE_BYTES, sCC + l * KC, p.K,
                                                     first + 2 * i * stride, stride));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) atomicMax(&cmax_bits[l], __float_as_int(mx));     // non-negative floats order as ints
        ptx::fence_proxy_async_smem();                  // my generic writes -> visible to the MMA's operand reads
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&mbar_bready[l]);
    };
    convert_level(0, tid, THREADS, 1);                  // 1024 items: 2 per thread
    // level 1 by the odd warpgroups (slot 1), level 2 by warpgroups 2 and 3: warpgroup 1 starts about half an
    // accumulator hold time after warpgroup 0, so the two slots stay out of phase (one scans while the other's MMAs
    // run) instead of competing for the tensor pipe and for the ALU pipe at the same moments
#ifndef RQ_V_NOSHIFT
    if (p.L > 1 && (wg & 1)) convert_level(1, (wg >> 1) * 128 + wtid, 256, 2);   // 4 per thread
    if (p.L > 2 && is_hi) convert_level(2, tid - 256, 256, 2);
#else
    if (p.L > 1 && is_hi) convert_level(1, tid - 256, 256, 2);
    if (p.L > 2 && is_hi) convert_level(2, tid - 256, 256, 2);
#endif
    if (is_hi && n_mine > 0) prefetch_x(tile_of(0));
    RQ_TL(2);

    if (!ptx::mbar_wait(mbar_tmem, 0)) timed_out = true;
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
    const uint32_t tmem_acc = tmem_base + (uint32_t)(slot * KC);                     // column offset
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);        // this warp's lanes
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TILE_M, KC);

    // ---- row-math layout: lane = (row 8 i + rsub of the warp's 32 rows, floats 8 q .. 8 q + 7), i = 0..3 -------------
    const int q = lane & 3, rsub = lane >> 2;
    // residual tile: 16-byte chunks 2q, 2q+1 of row wrow0 + 8 i + rsub (row & 7 == rsub); + 1024 i
    const uint32_t offR0 = (uint32_t)(wrow0 * 128 + rsub * 128 + (((2 * q) ^ rsub) << 4));
    uint8_t* const pR0 = sRw + offR0;
    uint8_t* const pR1 = sRw + (offR0 ^ 16u);          // chunk index ^ 1
    // A tile: hi chunk q, lo chunk 4 + q of the same row (chunk index ^ 4 = byte offset ^ 64).  The two rows of a
    // store phase land in the same half of the chunk columns (2-way bank conflict on these 8 stores per level); the
    // alternative, swapping hi / lo by row parity, costs 8 selects per row on the ALU pipe, the loaded one here.
    const uint32_t offA = (uint32_t)(wrow0 * 128 + rsub * 128 + ((q ^ rsub) << 4));
    uint8_t* const pA_hi = sAw + offA;
    uint8_t* const pA_lo = sAw + (offA ^ 64u);         // chunk index ^ 4
    const int lane_gofs = (wrow0 + rsub) * D + q * 8;      // float offset of my first 8 floats inside a [128, D] tile

    // Transposing reduction over the 4 lanes of a row group: v[i] is my chunk partial of row 8 i + rsub; the result is
    // the full sum for row 8 q + rsub — MY row of the group — in the canonical tree order (p0+p1)+(p2+p3) (fp32
    // addition commutes bit for bit).  3 shuffles instead of 8, and the per-row scalar math runs once per row.
    const bool qb0 = q & 1, qb1 = q & 2;
    auto quad_reduce_t = [&](const float (&v)[4]) {
        const float keep0 = qb0 ? v[1] : v[0], send0 = qb0 ? v[0] : v[1];
        const float keep1 = qb0 ? v[3] : v[2], send1 = qb0 ? v[2] : v[3];
        const float t0 = keep0 + __shfl_xor_sync(0xffffffffu, send0, 1);
        const float t1 = keep1 + __shfl_xor_sync(0xffffffffu, send1, 1);
        const float keep = qb1 ? t1 : t0, send = qb1 ? t0 : t1;
        return keep + __shfl_xor_sync(0xffffffffu, send, 2);
    };
    const int quad0 = lane & ~3;                        // lane of my row group that owns row 8 * 0 + rsub
    // scan layout <-> row-math layout: tile row wrow0 + t = 8 (t >> 3) + (t & 7) is owned by lane 4 (t & 7) + (t >> 3)
    const int owner_of_my_scan_row = 4 * (lane & 7) + (lane >> 3);
    auto load_row8 = [&](int i, float (&r)[8]) {        // residual tile -> my 8 floats of row 8 i + rsub
        const float4 a = *reinterpret_cast<const float4*>(pR0 + 1024 * i);
        const float4 b = *reinterpret_cast<const float4*>(pR1 + 1024 * i);
        r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
        r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
    };
    auto store_row8 = [&](int i, const float (&r)[8]) {
        *reinterpret_cast<float4*>(pR0 + 1024 * i) = make_float4(r[0], r[1], r[2], r[3]);
        *reinterpret_cast<float4*>(pR1 + 1024 * i) = make_float4(r[4], r[5], r[6], r[7]);
    };
    auto write_a_row8 = [&](int i, const float (&r)[8]) {   // bf16 hi/lo split -> my 16 + 16 bytes of the A tile
        uint32_t h[4], lw[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) ptx::split_pair_cvt(r[2 * j], r[2 * j + 1], h[j], lw[j]);
        *reinterpret_cast<uint4*>(pA_hi + 1024 * i) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(pA_lo + 1024 * i) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
    };
    auto store_id = [&](int l, int64_t row, int bi) {
        int64_t* dst = p.ids + (int64_t)l * p.B + row;
        if (p.n_changed) {                                  // k-means: count changed assignments (warp-aggregated)
            const bool ch = !p.has_prev || *dst != (int64_t)bi;
            const unsigned m = __ballot_sync(0xffffffffu, ch);
            if (lane == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
        }
        *dst = bi;
    };

    const int n_partner_uses = n_partner * p.L;
    // The use of (accumulator, A tile) `slot` that precedes our use u is the partner's: its use u when we are the
    // second warpgroup of the pair, u - 1 when we are the first.  Its MMAs must have read the A tile before we
    // overwrite it.
    auto wait_partner_mma = [&](int u) {
        const int prev = is_hi ? u : u - 1;
        if (prev >= 0 && prev < n_partner_uses && !ptx::mbar_wait(&mbar_full[wg ^ 2], prev & 1)) timed_out = true;
    };
    const bool ids_only = !p.emb && !p.res && !p.loss;
    int use = 0;                                   // this warpgroup's running (tile, level) count

    for (int it = 0; it < n_mine; ++it) {
        const int tile = tile_of(it);
        const int64_t trow0 = (int64_t)tile * TILE_M;
        const int64_t row = trow0 + wtid;          // scan layout: my row
        // row-math layout: where my 8 floats of row 8 i + rsub live in a [B, D] level
        const int64_t tile_gofs = trow0 * D + lane_gofs;
        float xxq, loss_q = 0.f;                   // ||r||^2 and running loss of MY row (8 q + rsub) of the group

        // ---- stage in: x rows -> A tile of level 0 (+ residuals[0]); the residual tile already holds them -------------
        ptx::cp_async_wait_all();
        __syncwarp();
        RQ_TL(60);
        wait_partner_mma(use);
        {
            float px[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float r[8];
                load_row8(i, r);
                write_a_row8(i, r);
                px[i] = chunk_dot<8>(r, r);
                if (p.res) ptx::st_global_v8(p.res + tile_gofs + i * 8 * D, r);
            }
            xxq = quad_reduce_t(px);
        }
        ptx::fence_proxy_async_smem();
        RQ_TL(3);

        for (int l = 0; l < p.L; ++l, ++use) {
            // ---- A rows of the whole tile are in place: one thread issues this level's MMAs -----------------------
            ptx::named_bar_sync(1 + wg, 128);
            if (wtid == 0) {
                if (!ptx::mbar_wait(&mbar_bready[l], 0)) timed_out = true;
                // the partner must have read its accumulator out: completion #(2*use - 1) resp. #(2*use) of the
                // slot's read-out barrier, parity 1 resp. 0
                const int prev = is_hi ? use : use - 1;
                if (prev >= 0 && prev < n_partner_uses && !ptx::mbar_wait(&mbar_free[slot], is_hi ? 0 : 1)) timed_out = true;
                ptx::tc_fence_after_sync();
                const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sB + l * B_TILE_BYTES);
                // (A k-offset bytes, B k-offset bytes): hi.hi, hi.hi, lo.hi, lo.hi, hi.lo, hi.lo
                const int ak[6] = {0, 32, 64, 96, 0, 32};
                const int bk[6] = {0, 32, 0, 32, 64, 96};
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    if (RQ_WHATIF(1) && i > 0) break;
                    const uint64_t ad = ptx::make_smem_desc(a0 + ak[i], 16, 1024, ptx::kLayoutSw128);
                    const uint64_t bd = ptx::make_smem_desc(b0 + bk[i], 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);
                }
                const uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
                const uint64_t cd = ptx::make_smem_desc(ptx::smem_u32(sCCt + l * CC_TILE_BYTES), 128, 256, ptx::kLayoutNone);
                ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
                ptx::umma_commit(&mbar_full[wg]);
            }
            __syncwarp();
            RQ_TL(4 + 8 * l);
            const float xs = __shfl_sync(0xffffffffu, xxq, owner_of_my_scan_row);     // ||r||^2 of my scan row
            RotX X = {};
            if (MODE == RQ_MODE_ROTATION_TRICK) X = rot_x(xxq);      // residual-only scalars, in the shadow of the MMAs
            if (!ptx::mbar_wait(&mbar_full[wg], use & 1)) timed_out = true;
            ptx::tc_fence_after_sync();
            RQ_TL(6 + 8 * l);
            const float cmax2 = __int_as_float(*reinterpret_cast<volatile int*>(&cmax_bits[l]));
            // |acc - exact| <= 2*(2^-18 + 2*2^-17)|r||c| = 3.8e-5 |r||c| (dropped lo.lo term + the two truncated lo
            // parts) + fp32 accumulation noise; both the winner and a rival may be off, hence 2x, with margin
            const float tol = 1e-4f * fast_sqrt(xs * cmax2) + 3e-6f * cmax2 + 1e-30f;

            if (P.dbg_acc && tile == 0 && l == 0) {
#pragma unroll 1
                for (int c = 0; c < KC / 32; ++c) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32b_x32(tmem_row + c * 32, v);
                    ptx::tmem_ld_wait_for(v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) P.dbg_acc[wtid * KC + c * 32 + j] = __uint_as_float(v[j]);
                }
            }
            // ---- scan ----------------------------------------------------------------------------------------------
            int bi;
            {
                float accA[16], accB[16];
                if (RQ_WHATIF(2)) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) { accA[i] = (float)((lane * 7 + i * 3 + l) & 15) + 1.f; accB[i] = (float)((lane * 5 + i) & 15) + 1.f; }
                } else
                scan_accumulator_row(tmem_row, accA, accB);
                // the accumulator is read out: hand it back (one arrival per warp)
                ptx::tc_fence_before_sync();            // orders this thread's tcgen05.ld before the arrival
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&mbar_free[slot]);
                RQ_TL(7 + 8 * l);
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
                bi = (int)(((__float_as_uint(m1a) & 15u) << 4) | (__float_as_uint(m1b) & 15u));
                // packing the index moves a value by < 2^-19 |v|: widen the band by 2^-17 |m1|
                const float m1 = fminf(m1a, m1b);
                const float thr = m1 + (tol + 8e-6f * fabsf(m1));
                const bool ambiguous = (m2a <= thr) || (m2b <= thr);
                if (__any_sync(0xffffffffu, ambiguous)) {
                    // ---- rare: every code within the band lies in a group with accA <= thr AND a class with
                    // accB <= thr; re-score those (group, class) pairs in fp32, ascending index, strict '<': exactly
                    // the SIMT family's arithmetic.  Extra pairs are harmless (their exact distance loses).
                    if (ambiguous) {
                        uint32_t gm = 0, cm = 0;
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            gm |= (accA[i] <= thr) ? (1u << i) : 0u;
                            cm |= (accB[i] <= thr) ? (1u << i) : 0u;
                        }
                        const float* cbl = p.cb + (int64_t)l * p.K * D;
                        float r[D];
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float4 t = *reinterpret_cast<const float4*>(sRw + sw128_off(wtid, c));
                            r[4 * c] = t.x; r[4 * c + 1] = t.y; r[4 * c + 2] = t.z; r[4 * c + 3] = t.w;
                        }
                        float best = INFINITY;
#pragma unroll 1
                        for (uint32_t g = gm; g; g &= g - 1) {
                            const int gi = __ffs(g) - 1;
#pragma unroll 1
                            for (uint32_t c = cm; c; c &= c - 1) {
                                const int k = gi * 16 + (__ffs(c) - 1);
                                if (k >= p.K) continue;
                                const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)k * D);
                                float a = 0.f;
#pragma unroll
                                for (int j = 0; j < D / 4; ++j) {
                                    const float4 cv = __ldg(c4 + j);
                                    a = fmaf(r[4 * j], cv.x, a); a = fmaf(r[4 * j + 1], cv.y, a);
                                    a = fmaf(r[4 * j + 2], cv.z, a); a = fmaf(r[4 * j + 3], cv.w, a);
                                }
                                const float d = (xs + sCC[l * KC + k]) - 2.f * a;
                                if (d < best) { best = d; bi = k; }
                            }
                        }
                    }
                    __syncwarp();
                }
            }
            if (bi >= p.K) bi = p.K - 1;        // unreachable unless inputs are NaN; keeps the gather in range
            store_id(l, row, bi);
            RQ_TL(8 + 8 * l);
            const bool last = (l == p.L - 1);
            // ids-only launches (tokenisation, k-means assignment) need nothing past the id at the last level
            if (last && ids_only) {
                if (KM) {
                    // ---- k-means: the tile's rows are still in the residual tile and their ids are known — add them
                    // to the CTA's partial sums here instead of reading x a second time.  Fixed point on the integer
                    // unit as in rq_kmeans_accumulate_fx_kernel (rq_kmeans.cu): the warp takes its 32 rows one at a
                    // time, lane l component l (one conflict-free LDS, two conflict-free reductions per row), the counts
                    // are one reduction for the 32 rows.  range = 2^e >= 16 max||c||.
                    const uint32_t km_hi = ptx::smem_u32(smem + S.km) + lane * 4;
                    const uint32_t km_lo = km_hi + KC * D * 4;
                    const uint32_t km_cnt = ptx::smem_u32(smem + S.km) + 2 * KC * D * 4;
                    int e2 = (int)((__float_as_uint(fast_sqrt(cmax2)) >> 23) & 0xffu) + 5;
                    e2 = e2 < 30 ? 30 : (e2 > 220 ? 220 : e2);
                    const float S_fx = 4194304.f / __uint_as_float((unsigned)e2 << 23);
                    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(km_cnt + (uint32_t)bi * 4u), "r"(1u) : "memory");
                    bool bad = false;
                    const uint8_t* rows = sRw + wrow0 * 128 + (lane & 3) * 4;
#pragma unroll 8
                    for (int t = 0; t < 32; ++t) {
                        const int k = __shfl_sync(0xffffffffu, bi, t);
                        const float v = *reinterpret_cast<const float*>(rows + (t >> 3) * 1024 + (t & 7) * 128 +
                                                                        (((lane >> 2) ^ (t & 7)) << 4));
                        const float sc = v * S_fx;
                        const bool ok = fabsf(sc) < 4194304.f;
                        const int f = ok ? (__float_as_int(sc + 12582912.f) - 0x4B400000) : 0;
                        const uint32_t off = (uint32_t)k * (D * 4);
                        asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(km_hi + off), "r"(f >> 11) : "memory");
                        asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(km_lo + off), "r"(f & 0x7ff) : "memory");
                        bad |= !ok;
                    }
                    if (__any_sync(0xffffffffu, bad)) {        // rare: components outside the fixed-point range
#pragma unroll 1
                        for (int t = 0; t < 32; ++t) {
                            const int k = __shfl_sync(0xffffffffu, bi, t);
                            const float v = *reinterpret_cast<const float*>(rows + (t >> 3) * 1024 + (t & 7) * 128 +
                                                                            (((lane >> 2) ^ (t & 7)) << 4));
                            if (!(fabsf(v * S_fx) < 4194304.f)) atomicAdd(p.km_sums + k * D + lane, v);
                        }
                    }
                    __syncwarp();                               // every lane has read the rows: the tile may be refilled
                }
                if (it + 1 < n_mine) prefetch_x(tile_of(it + 1));       // the residual tile is free (__syncwarp above)
                continue;
            }

            // ---- row math, two rows per lane at a time: gather, estimator, loss, outputs, next residual, next A rows ----
            const float* cbl = p.cb + (int64_t)l * p.K * D + q * 8;
            // my 8 floats of row 8 i + rsub in this level's embedding / the next level's residual output: + 1024 i bytes
            float* const emb_l = p.emb ? p.emb + (int64_t)l * p.B * D + tile_gofs : nullptr;
            float* const res_n = p.res ? p.res + (int64_t)(l + 1) * p.B * D + tile_gofs : nullptr;
            float e[4][8], r[4][8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {               // all four gathers in flight (L2 latency)
                const int k = __shfl_sync(0xffffffffu, bi, 8 * i + rsub);
                if (RQ_WHATIF(4)) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) e[i][j] = (float)(k + j) * 0.01f;
                } else
                ptx::ld_global_v8(cbl + (int64_t)k * D, e[i]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) load_row8(i, r[i]);
            RQ_TL_DEP(5 + 8 * l, e[3][7]);
            if (p.loss) {
                float pd[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) pd[i] = chunk_sqdiff<8>(r[i], e[i]);
                const float dsq = quad_reduce_t(pd);
                loss_q += dsq + p.beta * dsq;                                                  // loss.py:40-42
            }
            if (MODE == RQ_MODE_ROTATION_TRICK) {
                // o = ca r + ce e (rq_rowmath.cuh: level_epilogue), in place over e.  The scalars of row 8 i + rsub are
                // computed by lane i of the group and handed to the other three by shuffles.
                float pe[4], pr[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    pe[i] = chunk_dot<8>(e[i], e[i]);
                    pr[i] = chunk_dot<8>(r[i], e[i]);
                }
                const float e2 = quad_reduce_t(pe), re = quad_reduce_t(pr);
                float ne, iq;
                float ns = rot_prepare(X, xxq, e2, re, ne, iq);
                if (__any_sync(0xffffffffu, ns < kRotCancel)) {        // rare: |u + q| small, recompute it directly
                    float pn[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        pn[i] = chunk_sq_axpby<8>(r[i], e[i], __shfl_sync(0xffffffffu, X.iu, quad0 + i),
                                                  __shfl_sync(0xffffffffu, iq, quad0 + i));
                    const float nd = quad_reduce_t(pn);
                    if (ns < kRotCancel) ns = nd;
                }
                float ca, ce;
                rot_finish(X, xxq, re, ne, iq, ns, ca, ce);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float cai = __shfl_sync(0xffffffffu, ca, quad0 + i), cei = __shfl_sync(0xffffffffu, ce, quad0 + i);
                    const f32x2 ca2 = pack2(cai, cai), ce2 = pack2(cei, cei);
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        unpack2(fma2(ca2, pack2(r[i][2 * j], r[i][2 * j + 1]), mul2(ce2, pack2(e[i][2 * j], e[i][2 * j + 1]))),
                                e[i][2 * j], e[i][2 * j + 1]);
                }
            } else if (MODE == RQ_MODE_STE) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {                              // quantize.py:132: x + (emb - x)
                        const f32x2 rp = pack2(r[i][2 * j], r[i][2 * j + 1]), ep = pack2(e[i][2 * j], e[i][2 * j + 1]);
                        unpack2(add2(rp, sub2(ep, rp)), e[i][2 * j], e[i][2 * j + 1]);
                    }
            }                                                                  // EVAL: the codeword itself, quantize.py:149
            // the A tile is shared with the partner warpgroup: its MMAs of the use before ours must have read it before
            // the next level's rows go in (long done by now in the steady state)
            if (!last) wait_partner_mma(use + 1);
            RQ_TL(11 + 8 * l);
            {
                float px[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (emb_l && !RQ_WHATIF(8)) ptx::st_global_v8(emb_l + i * 8 * D, e[i]);
                    if (!last) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)                            // rqvae.py:129
                            unpack2(sub2(pack2(r[i][2 * j], r[i][2 * j + 1]), pack2(e[i][2 * j], e[i][2 * j + 1])),
                                    r[i][2 * j], r[i][2 * j + 1]);
                        store_row8(i, r[i]);
                        write_a_row8(i, r[i]);
                        px[i] = chunk_dot<8>(r[i], r[i]);
                        if (res_n && !RQ_WHATIF(8)) ptx::st_global_v8(res_n + i * 8 * D, r[i]);
                    }
                }
                if (!last) xxq = quad_reduce_t(px);
            }
            RQ_TL(9 + 8 * l);
            if (!last) {
                ptx::fence_proxy_async_smem();          // my A rows -> visible to the MMAs' operand reads
            } else if (it + 1 < n_mine) {
                __syncwarp();                           // everybody has read its rows: fetch the next tile's rows
                prefetch_x(tile_of(it + 1));
            }
            RQ_TL(10 + 8 * l);
        }
        // ---- tail of the tile: the loss (row-math layout -> one row per lane, coalesced) ----------------------------------
        if (p.loss) {
            const float ls = __shfl_sync(0xffffffffu, loss_q, owner_of_my_scan_row);
            p.loss[row] = ls;
        }
    }

    if (timed_out && P.err) atomicExch(P.err, 1);
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 9) ptx::tmem_dealloc(tmem_base, 512);
    if (KM) {                                           // the CTA's partials: one float reduction per (cluster, component)
        const int* km = reinterpret_cast<const int*>(smem + S.km);
        const float cm2 = __int_as_float(*reinterpret_cast<volatile int*>(&cmax_bits[0]));
        int e2 = (int)((__float_as_uint(fast_sqrt(cm2)) >> 23) & 0xffu) + 5;
        e2 = e2 < 30 ? 30 : (e2 > 220 ? 220 : e2);
        const double inv = (double)__uint_as_float((unsigned)e2 << 23) / 4194304.0;
        for (int i = tid; i < p.K * D; i += THREADS) {
            const long long tot = (long long)km[i] * 2048 + (long long)km[KC * D + i];
            if (tot != 0) atomicAdd(p.km_sums + i, (float)((double)tot * inv));
        }
        for (int k = tid; k < p.K; k += THREADS) {
            const unsigned c = (unsigned)km[2 * KC * D + k];
            if (c) atomicAdd(p.km_counts + k, (unsigned long long)c);
        }
    }
    RQ_CTA_STAMP(2, ptx::globaltimer());
}

static int* d_err32 = nullptr;
static float* g_dbg_acc32 = nullptr;
#ifdef RQ_EXPERIMENTS
static long long* g_dbg_tl32 = nullptr;
#endif

bool tc32_fused_supported(int D, int K, int L, int mode, int distance) {
    return D == tc::D && K >= 1 && K <= tc::KC && L >= 1 && L <= tc32::MAX_L && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

void tc32_set_accumulator_dump(float* buf) { g_dbg_acc32 = buf; }
int tc32_timeout_flag() {
    int v = 0;
    if (d_err32) cudaMemcpy(&v, d_err32, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}

int quantize_forward_tc32_fused(const FwdParams& p, cudaStream_t st) {
    if (p.B == 0) return 0;
    const int64_t n_tiles = p.B / tc::TILE_M;
    if (n_tiles > 0) {
        if (!d_err32) {
            RQ_CUDA(cudaMalloc((void**)&d_err32, sizeof(int)));
            RQ_CUDA(cudaMemset(d_err32, 0, sizeof(int)));
        }
        TcParams P = {};
        P.f = p;
        P.err = d_err32;
        P.dbg_acc = g_dbg_acc32;
#ifdef RQ_EXPERIMENTS
        P.dbg_tl = g_dbg_tl32;
        P.dbg_flags = getenv("RQ_TC_WHATIF") ? atoi(getenv("RQ_TC_WHATIF")) : 0;    // experiments build only
#endif
        P.n_tiles = (int)n_tiles;
        const bool km = p.km_sums != nullptr;
        const tc32::SmemLayout S = tc32::smem_layout(p.L, km);
        const int grid = P.n_tiles < sm_count() ? P.n_tiles : sm_count();
        const size_t smem = (size_t)S.total + 1024;
        if (km) {
            // fused k-means assignment + accumulation: single level, ids only (the caller checked), 2^20 rows per CTA at most
            if (p.L != 1 || p.emb || p.res || p.loss || !p.km_counts || smem > (size_t)tc::kMaxSmem ||
                (int64_t)P.n_tiles * tc::TILE_M > (int64_t)grid << 20) {
                set_error("fused k-means accumulation: unsupported launch shape");
                return RQ_ERR_UNSUPPORTED_SHAPE;
            }
            RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc32_kernel<RQ_MODE_EVAL, true>, tc::kMaxSmem));
            RQ_CUDA(launch_pdl(rq_fwd_tc32_kernel<RQ_MODE_EVAL, true>, dim3(grid), dim3(tc32::THREADS), smem, st, P));
        } else
        switch (p.mode) {
            case RQ_MODE_ROTATION_TRICK:
                RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc32_kernel<RQ_MODE_ROTATION_TRICK>, tc::kMaxSmem));
                RQ_CUDA(launch_pdl(rq_fwd_tc32_kernel<RQ_MODE_ROTATION_TRICK>, dim3(grid), dim3(tc32::THREADS), smem, st, P));
                break;
            case RQ_MODE_STE:
                RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc32_kernel<RQ_MODE_STE>, tc::kMaxSmem));
                RQ_CUDA(launch_pdl(rq_fwd_tc32_kernel<RQ_MODE_STE>, dim3(grid), dim3(tc32::THREADS), smem, st, P));
                break;
            default:
                RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc32_kernel<RQ_MODE_EVAL>, tc::kMaxSmem));
                RQ_CUDA(launch_pdl(rq_fwd_tc32_kernel<RQ_MODE_EVAL>, dim3(grid), dim3(tc32::THREADS), smem, st, P));
                break;
        }
        RQ_LAUNCHED();
    }
    // the last B % 128 rows: exact family (bit-identical ids, same row arithmetic), one small launch
    if (n_tiles * tc::TILE_M < p.B) {
        FwdParams t = p;
        t.row0 = n_tiles * tc::TILE_M;
        return quantize_forward_simt(t, tc::D, st);
    }
    return 0;
}

}  // namespace rq

// Profiling hook: only builds with -DRQ_EXPERIMENTS (python -m rq_vae_recommender_b200.build --experiments) carry the
// clock64 marks; the release library answers RQ_ERR_UNSUPPORTED_SHAPE.
extern "C" int rq_debug_tc_set_timeline(long long* device_buf_32x64) {
#ifdef RQ_EXPERIMENTS
    rq::g_dbg_tl32 = device_buf_32x64;
    return 0;
#else
    (void)device_buf_32x64;
    return RQ_ERR_UNSUPPORTED_SHAPE;
#endif
}
