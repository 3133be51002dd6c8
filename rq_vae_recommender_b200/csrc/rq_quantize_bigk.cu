// Codebooks larger than one UMMA tile (K > 256) at D = 64 — BASELINE config 5: 4 levels x 1024 codes.
//
// The operand tile of 1024 x 64 codes (128 KB in fp16) leaves no room for more than one level per SM, so
// the levels run one after the other (level-major) and every level is a composition of launches:
//
//   rq_fwd_tc64_kernel with L = 1, chunked: CTA b scores every row of the current residual against the
//       256-code chunk b % n_chunks of the level's codebook        -> per chunk, winner id and its exact
//                                                                          fp32 distance per row (best_dist)
//   rq_bigk_merge_kernel: the chunk whose winner has the smallest exact distance (lowest chunk on a tie,
//       i.e. the lowest code index), codeword gather, estimator, loss, outputs, next residual.
//
// Each chunk winner is the exact fp32 argmin of its chunk (rq_quantize_tc64.cu) and its distance is computed
// with the SIMT family's arithmetic, so the merged id is the SIMT family's id bit for bit; the row epilogue is
// the shared level_epilogue<D>.  The residual travels through HBM between levels (it is an output of the
// training forward anyway); scratch comes from the library's stream-ordered pool.
#include <stdlib.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"
#include "rq_tc64_tiles.cuh"

namespace rq {

constexpr int kBigkChunk = 256;
constexpr int kBigkThreads = 128;

struct BigkMergeParams {
    const float* r_in;        // [B, D] residual entering the level
    const float* cb;          // [K, D] this level's codebook
    const int64_t* cand_ids;  // [n_chunks][B] winner inside each chunk
    const float* cand_d;      // [n_chunks][B] its exact distance
    int64_t B;
    int K, n_chunks, mode, first;
    float beta;
    float* emb;               // [B, D] or nullptr
    float* res;               // [B, D] or nullptr
    int64_t* ids;             // [B]
    float* loss;              // [B] or nullptr (accumulated over the levels)
    float* r_out;             // [B, D] or nullptr (last level)
};

template <int D>
__global__ void __launch_bounds__(kBigkThreads) rq_bigk_merge_kernel(BigkMergeParams m) {
    const int64_t row = (int64_t)blockIdx.x * kBigkThreads + threadIdx.x;
    if (row >= m.B) return;
    int bc = 0;
    float bd = __ldg(m.cand_d + row);
    for (int c = 1; c < m.n_chunks; ++c) {
        const float d = __ldg(m.cand_d + (int64_t)c * m.B + row);
        if (d < bd) { bd = d; bc = c; }                       // strict: the lowest chunk wins a tie
    }
    int k = bc * kBigkChunk + (int)__ldg(m.cand_ids + (int64_t)bc * m.B + row);
    if (k >= m.K) k = m.K - 1;
    float r[D], e[D];
    load_row<D>(m.r_in + row * D, r);
    load_row<D>(m.cb + (int64_t)k * D, e);
    const float xx = dot_self<D>(r);
    if (m.res) store_row<D>(m.res + row * D, r);
    const float lvl = level_epilogue<D>(m.mode, m.beta, r, e, xx, e);     // estimator value in place
    if (m.emb) store_row<D>(m.emb + row * D, e);
    m.ids[row] = k;
    if (m.loss) m.loss[row] = m.first ? lvl : m.loss[row] + lvl;
    if (m.r_out) {
        residual_update<D>(r, e);                                          // rqvae.py:129
        store_row<D>(m.r_out + row * D, r);
    }
}

bool tc64_bigk_supported(int D, int K, int L, int mode, int distance) {
    // also serves K <= 256 with more levels than the fused d=64 kernel keeps resident (L > 3)
    return D == 64 && K >= 1 && L >= 1 && (K > kBigkChunk || L > 3) && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

static int quantize_forward_tc64_bigk_chunked(const FwdParams& p, cudaStream_t st) {
    constexpr int D = 64;
    if (p.B == 0) return 0;
    const int n_chunks = (p.K + kBigkChunk - 1) / kBigkChunk;
    int64_t* cand_ids = (int64_t*)engine_pool_alloc((size_t)n_chunks * p.B * sizeof(int64_t), st);
    float* cand_d = (float*)engine_pool_alloc((size_t)n_chunks * p.B * sizeof(float), st);
    float* rbuf[2] = {nullptr, nullptr};
    if (p.L > 1) {
        rbuf[0] = (float*)engine_pool_alloc((size_t)p.B * D * sizeof(float), st);
        if (p.L > 2) rbuf[1] = (float*)engine_pool_alloc((size_t)p.B * D * sizeof(float), st);
    }
    int rc = 0;
    if (!cand_ids || !cand_d || (p.L > 1 && !rbuf[0]) || (p.L > 2 && !rbuf[1])) {
        set_error("rq_quantize_forward: out of device memory for the K > 256 scratch buffers");
        rc = (int)cudaErrorMemoryAllocation;
    }
    const float* r_in = p.x;
    for (int l = 0; l < p.L && rc == 0; ++l) {
        const float* cbl = p.cb + (int64_t)l * p.K * D;
        {
            // one launch scores every row against all chunks: CTA b serves chunk b % n_chunks
            FwdParams q;
            q.x = r_in; q.cb = cbl; q.B = p.B; q.K = p.K;
            q.L = 1; q.mode = RQ_MODE_EVAL; q.dist = RQ_DIST_L2; q.beta = p.beta; q.temp = p.temp;
            q.gu = nullptr; q.seed = 0; q.offset = 0;
            q.emb = nullptr; q.res = nullptr; q.loss = nullptr; q.KC = 0;
            q.ids = cand_ids;
            q.best_dist = cand_d;
            q.n_chunks = n_chunks;
            rc = quantize_forward_tc64(q, st);
        }
        if (rc) break;
        BigkMergeParams m;
        m.r_in = r_in; m.cb = cbl; m.cand_ids = cand_ids; m.cand_d = cand_d; m.B = p.B; m.K = p.K;
        m.n_chunks = n_chunks; m.mode = p.mode; m.first = (l == 0); m.beta = p.beta;
        m.emb = p.emb ? p.emb + (int64_t)l * p.B * D : nullptr;
        m.res = p.res ? p.res + (int64_t)l * p.B * D : nullptr;
        m.ids = p.ids + (int64_t)l * p.B;
        m.loss = p.loss;
        m.r_out = (l + 1 < p.L) ? rbuf[l & 1] : nullptr;
        const unsigned grid = (unsigned)((p.B + kBigkThreads - 1) / kBigkThreads);
        rq_bigk_merge_kernel<D><<<grid, kBigkThreads, 0, st>>>(m);
        RQ_LAUNCHED();
        r_in = m.r_out;
    }
    if (cand_ids) cudaFreeAsync(cand_ids, st);
    if (cand_d) cudaFreeAsync(cand_d, st);
    if (rbuf[0]) cudaFreeAsync(rbuf[0], st);
    if (rbuf[1]) cudaFreeAsync(rbuf[1], st);
    return rc;
}


// ======================================================================================================================
// Fused form (L <= 8): ONE preparation launch + ONE persistent launch for all levels and all code chunks.
//
//   rq_bigk_prepare_kernel   one CTA per level: ||c||^2 (canonical order), the level's power-of-two scale t, and per
//                            256-code chunk the images of the UMMA operand tiles exactly as rq_fwd_tc64_kernel builds
//                            them in shared memory ([-2 t c] fp16, 128-byte swizzled; the [hi, mid, lo] split of
//                            t^2 ||c||^2), laid out back to back in global scratch (40 KB per chunk, L2-resident).
//   rq_fwd_tc64_bigk_kernel  persistent CTA per SM, two warpgroups, each with its own TMEM accumulator and its own row
//                            tile (thread = row, the residual in registers across the levels as in rq_fwd_tc64_kernel).
//                            The two warpgroups walk the same (level, chunk) sequence, so the chunk images stream
//                            through ONE ring of shared-memory stages filled by cp.async.bulk: a stage is refilled by
//                            whichever warpgroup is second to finish its MMAs on it (a counter, nobody polls).  Per
//                            chunk: 9 tcgen05.mma, the single-pass scan, exact re-scoring inside the error band, the
//                            exact fp32 distance of the chunk's winner; the running minimum over the chunks (strict '<':
//                            the lowest chunk wins a tie) is the SIMT family's id bit for bit.  The MMAs of chunk c + 1
//                            are issued as soon as the accumulator of chunk c has been read, so they run under chunk
//                            c's re-scoring.  The A rows are written once per level, the row never leaves the SM
//                            between levels, and the merge kernel and its [n_chunks, B] scratch are gone.
namespace bigk {
using namespace tc64;
// One chunk's operand images: [-2 t c] as fp16 hi and lo parts (two 32 KB tiles) + the ||c||^2 tile.  With the lo part
// the product carries 22 bits of the codeword (error 3 * 2^-22 |t r||2 t c| instead of 2^-11): the scan's band shrinks
// 25-fold and the re-scoring path, a fifth of the launch with hi-only images, all but disappears; 13 MMAs per chunk
// instead of 9.
constexpr int STAGE_BYTES = 2 * B_TILE_BYTES + CC_TILE_BYTES;  // 72 KB
constexpr int N_STAGES = 2;
constexpr int BIGK_MAX_L = 8;
constexpr int SCRATCH_F32 = 64;                                // floats per warp: one parked row

struct BigkSmem {
    int stages, a_tiles, ones, cc_f32, scratch, misc, total;
};
__host__ __device__ inline BigkSmem bigk_smem_layout(int K) {
    BigkSmem s;
    int off = 0;
    s.stages = off; off += N_STAGES * STAGE_BYTES;
    s.a_tiles = off; off += NWG * A_TILE_BYTES;
    s.ones = off; off += ONES_TILE_BYTES;
    s.cc_f32 = off;                                              // (||c||^2 is read from L2 in the rare paths)
    s.scratch = off; off += NWG * 4 * SCRATCH_F32 * 4;
    s.misc = off; off += 256;
    s.total = off;
    return s;
}
}  // namespace bigk

struct BigkPrepParams {
    const float* cb;     // [L, K, D]
    int K, L, n_chunks;
    uint8_t* img;        // [L][n_chunks][STAGE_BYTES]
    float* cc;           // [L][n_chunks * 256] ||c||^2 (1e30 for padding codes)
    float* lvl;          // [2 * BIGK_MAX_L]: max ||c||^2 per level, then the scale t per level
};

__global__ void __launch_bounds__(1024) rq_bigk_prepare_kernel(BigkPrepParams q) {
    using namespace bigk;
    __shared__ float s_max[32];
    __shared__ float s_t;
    const int l = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int Kp = q.n_chunks * KC;
    const float* cbl = q.cb + (int64_t)l * q.K * D;
    float* ccl = q.cc + (int64_t)l * Kp;
    // pass 1: ||c||^2 in the canonical order of code_sqnorm<64> (8-wide sequential partials, xor-shuffles 1, 2, 4)
    float mx = 0.f;
    for (int i = tid; i < Kp * 8; i += blockDim.x) {
        const int k = i >> 3, c8 = i & 7;
        float part = 0.f;
        if (k < q.K) {
            const float4* src = reinterpret_cast<const float4*>(cbl + (int64_t)k * D + c8 * 8);
            const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
            part = fmaf(v0.x, v0.x, part); part = fmaf(v0.y, v0.y, part);
            part = fmaf(v0.z, v0.z, part); part = fmaf(v0.w, v0.w, part);
            part = fmaf(v1.x, v1.x, part); part = fmaf(v1.y, v1.y, part);
            part = fmaf(v1.z, v1.z, part); part = fmaf(v1.w, v1.w, part);
        }
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        part += __shfl_xor_sync(0xffffffffu, part, 2);
        part += __shfl_xor_sync(0xffffffffu, part, 4);
        if (c8 == 0) ccl[k] = (k < q.K) ? part : 1e30f;
        if (k < q.K) mx = fmaxf(mx, part);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_max[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        float m = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = fmaxf(m, s_max[w]);
        // t = 2^-(floor(log2 max||c||) + 1): t * max||c|| in [0.5, 1); exponent clamped so t stays normal
        int e = (int)((__float_as_uint(sqrtf(m)) >> 23) & 255u) - 127;
        e = (m > 0.f) ? max(-100, min(100, e)) : -1;
        const float t = __int_as_float((127 - (e + 1)) << 23);
        q.lvl[l] = m;
        q.lvl[BIGK_MAX_L + l] = t;
        s_t = t;
    }
    __syncthreads();                                               // also orders the ccl writes before the reads below
    const float t = s_t, m2t = -2.f * t;
    // pass 2: operand images per chunk
    for (int i = tid; i < Kp * 8; i += blockDim.x) {
        const int k = i >> 3, c8 = i & 7, ch = k >> 8, kk = k & 255;
        uint8_t* st = q.img + ((int64_t)l * q.n_chunks + ch) * STAGE_BYTES;
        uint4 w = make_uint4(0, 0, 0, 0), wl = make_uint4(0, 0, 0, 0);
        if (k < q.K) {
            const float4* src = reinterpret_cast<const float4*>(cbl + (int64_t)k * D + c8 * 8);
            const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
            split_pair_h(m2t * v0.x, m2t * v0.y, w.x, wl.x); split_pair_h(m2t * v0.z, m2t * v0.w, w.y, wl.y);
            split_pair_h(m2t * v1.x, m2t * v1.y, w.z, wl.z); split_pair_h(m2t * v1.z, m2t * v1.w, w.w, wl.w);
        }
        *reinterpret_cast<uint4*>(st + sw128_off(kk, c8)) = w;
        *reinterpret_cast<uint4*>(st + B_TILE_BYTES + sw128_off(kk, c8)) = wl;
        if (c8 == 0) {
            uint4 c0 = make_uint4(0, 0, 0, 0);
            if (k < q.K) {
                const float v = t * t * ccl[k];                   // <= 1
                const float hh = h_round(v), mm = h_round(v - hh), ll = h_round((v - hh) - mm);
                c0.x = pack_h2(hh, mm); c0.y = pack_h2(ll, 0.f);
            } else {
                c0.x = pack_h2(60000.f, 0.f);                     // padding codes: larger than any real score, finite
            }
            *reinterpret_cast<uint4*>(st + 2 * B_TILE_BYTES + nosw_off(kk, 0)) = c0;
            *reinterpret_cast<uint4*>(st + 2 * B_TILE_BYTES + nosw_off(kk, 1)) = make_uint4(0, 0, 0, 0);
        }
    }
}

struct BigkParams {
    FwdParams f;
    int* err;
    int n_pairs;          // pairs of consecutive 128-row tiles (one per warpgroup)
    int n_chunks;
    const uint8_t* img;   // [L][n_chunks][STAGE_BYTES]
    const float* cc;      // [L][n_chunks * 256]
    const float* lvl;     // [2 * BIGK_MAX_L]
};

__global__ void __launch_bounds__(tc64::THREADS, 1) rq_fwd_tc64_bigk_kernel(BigkParams P) {
    using namespace bigk;
    const FwdParams& p = P.f;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const BigkSmem S = bigk_smem_layout(p.K);
    uint8_t* sStage = smem + S.stages;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sOnes = smem + S.ones;
    uint64_t* mbar_full = reinterpret_cast<uint64_t*>(smem + S.misc);            // [2] accumulator ready
    uint64_t* mbar_ready = reinterpret_cast<uint64_t*>(smem + S.misc + 16);      // [N_STAGES] chunk images landed
    uint32_t* stage_rel = reinterpret_cast<uint32_t*>(smem + S.misc + 64);       // [N_STAGES] releases (2 per use)
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 96);
    uint32_t* wg_events = reinterpret_cast<uint32_t*>(smem + S.misc + 104);      // [2] see warp_arrive below
    float* s_cmax = reinterpret_cast<float*>(smem + S.misc + 128);               // [BIGK_MAX_L]
    float* s_scale = reinterpret_cast<float*>(smem + S.misc + 160);              // [BIGK_MAX_L]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wg = tid >> 7, wtid = tid & 127;
    const int Kp = P.n_chunks * KC;
    float* scratch = reinterpret_cast<float*>(smem + S.scratch) + (wg * 4 + (warp & 3)) * SCRATCH_F32;
    const int steps_per_tile = p.L * P.n_chunks;
    const int my_pairs = (P.n_pairs > (int)blockIdx.x) ? (P.n_pairs - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int total_steps = my_pairs * steps_per_tile;
    bool timed_out = false;

    // chunk images of step s -> stage s % N_STAGES (one thread; completion as transaction bytes on mbar_ready)
    auto fill_stage = [&](int s) {
        const int lc = s % steps_per_tile;                        // = level * n_chunks + chunk
        const int st = s % N_STAGES;
        ptx::mbar_arrive_expect_tx(&mbar_ready[st], (uint32_t)STAGE_BYTES);
        ptx::bulk_g2s(sStage + st * STAGE_BYTES, P.img + (int64_t)lc * STAGE_BYTES, (uint32_t)STAGE_BYTES, &mbar_ready[st]);
    };

    if (tid == 0) {
        ptx::mbar_init(&mbar_full[0], 1);
        ptx::mbar_init(&mbar_full[1], 1);
        for (int i = 0; i < N_STAGES; ++i) {
            ptx::mbar_init(&mbar_ready[i], 1);
            stage_rel[i] = 0;
        }
        wg_events[0] = wg_events[1] = 0;
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, 512);
    if (tid < BIGK_MAX_L) {
        s_cmax[tid] = (tid < p.L) ? P.lvl[tid] : 0.f;
        s_scale[tid] = (tid < p.L) ? P.lvl[BIGK_MAX_L + tid] : 1.f;
    }
    // ---- first tile's rows: loads in flight during the prologue ----------------------------------------------------
    auto tile_of = [&](int it) { return 2 * ((int)blockIdx.x + (int)gridDim.x * it) + wg; };
    float r[D];
    auto load_rows = [&](int it) {
        const int64_t row = (int64_t)tile_of(it) * TILE_M + wtid;
        if (row < p.B) {
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + row * D + j, &r[j]);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
    };
    // The row of the NEXT tile is pulled into L2 while the current tile is worked on (the row itself lives in registers,
    // so the real load can only be issued once the current tile is finished; without the prefetch its HBM latency is
    // exposed at the top of every tile: a fifth of the warp time of a 4M-item tokenisation launch).
    auto prefetch_rows_l2 = [&](int it) {
        const int64_t row = (int64_t)tile_of(it) * TILE_M + wtid;
        if (row < p.B) {
            const float* ptr = p.x + row * D;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr + 32));
        }
    };
    if (my_pairs > 0) load_rows(0);
    for (int i = tid; i < TILE_M * 2; i += THREADS) {          // ones tile: k = 0,1,2 -> 1.0
        int rr = i >> 1, c = i & 1;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0) { v.x = 0x3C003C00u; v.y = 0x00003C00u; }
        *reinterpret_cast<uint4*>(sOnes + nosw_off(rr, c)) = v;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    if (tid == 0)
        for (int s = 0; s < N_STAGES && s < total_steps; ++s) fill_stage(s);

    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
    const uint32_t tmem_acc = tmem_base + (uint32_t)(wg * KC);
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);
    uint8_t* sAw = sA + wg * A_TILE_BYTES;
    const uint32_t idesc = make_idesc_f16_f32(TILE_M, KC);

    // the MMAs of step s (one thread of the warpgroup): A = this warpgroup's rows, B = the stage's chunk images
    auto issue_step = [&](int s) {
        const int st = s % N_STAGES;
        if (!ptx::mbar_wait(&mbar_ready[st], (uint32_t)((s / N_STAGES) & 1))) timed_out = true;
        ptx::tc_fence_after_sync();
        const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sStage + st * STAGE_BYTES);
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            // (A block, B tile): hi.hi, lo.hi, hi.lo — four K-steps of 16 each
            const int ab = (i >> 2) == 1, bb = (i >> 2) == 2;
            uint64_t ad = ptx::make_smem_desc(a0 + ab * A_BLOCK_BYTES + (i & 3) * 32, 16, 1024, ptx::kLayoutSw128);
            uint64_t bd = ptx::make_smem_desc(b0 + bb * B_TILE_BYTES + (i & 3) * 32, 16, 1024, ptx::kLayoutSw128);
            ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);         // kind::f16; the descriptor says fp16
        }
        uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
        uint64_t cd = ptx::make_smem_desc(b0 + 2 * B_TILE_BYTES, 128, 256, ptx::kLayoutNone);
        ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
        ptx::umma_commit(&mbar_full[wg]);
    };
    // this warpgroup's MMAs of step s are complete: release the stage; the second of the two releases refills it
    auto release_stage = [&](int s) {
        const uint32_t old = ptx::atom_add_acq_rel_shared(&stage_rel[s % N_STAGES], 1u);
        if ((old & 1u) && s + N_STAGES < total_steps) fill_stage(s + N_STAGES);
    };

    // A warp has done its part of what the MMAs of step s wait for (its A rows are written, or it has read the previous
    // chunk's accumulator out): one arrival per warp on the warpgroup's counter, and the warp that completes the four
    // issues.  Every arrival for the next event comes after a wait on this step's MMAs, so events never interleave and
    // no warp waits at a barrier for a neighbour that is busy re-scoring.
    auto warp_arrive = [&](int s) {
        ptx::tc_fence_before_sync();                // this thread's tcgen05.ld precede the arrival
        __syncwarp();
        if (lane == 0) {
            const uint32_t old = ptx::atom_add_acq_rel_shared(&wg_events[wg], 1u);
            if ((old & 3u) == 3u) issue_step(s);
        }
        __syncwarp();
    };
    int step = 0;
    for (int it = 0; it < my_pairs; ++it) {
        const int tile = tile_of(it);
        const int64_t row = (int64_t)tile * TILE_M + wtid;
        const bool active = row < p.B;
        if (it + 1 < my_pairs) prefetch_rows_l2(it + 1);
        float xx = dot_self<D>(r);
        float loss_acc = 0.f;

        for (int l = 0; l < p.L; ++l) {
            const float ts = s_scale[l];
            const float trn = ts * sqrtf(xx);
            // |t r| >= 2^14 would overflow fp16: such a row skips the tensor-core score and is re-scored against all
            // codes in fp32 below (rq_fwd_tc64_kernel)
            const bool wild = !(trn < 16384.f);
            // ---- A operand row: [hi(t r) | lo(t r)], two 128-byte swizzled K-blocks; written once per level ---------
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t h[4], lw[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    split_pair_h(ts * r[c * 8 + 2 * q], ts * r[c * 8 + 2 * q + 1], h[q], lw[q]);
                    if (wild) { h[q] = 0u; lw[q] = 0u; }
                }
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, c)) = make_uint4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<uint4*>(sAw + A_BLOCK_BYTES + sw128_off(wtid, c)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            // this level's ||c||^2 for the re-scoring arithmetic: read from L2 where needed (the rare paths)
            const float* sCC = P.cc + (int64_t)l * Kp;
            ptx::fence_proxy_async_smem();
            warp_arrive(step);                      // A rows in place: the last warp to get here issues the MMAs
            // ---- while the MMAs run: this level's input residual goes out ------------------------------------------
            const int64_t off = ((int64_t)l * p.B + row) * D;
            if (active && p.res) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.res + off + j, &r[j]);
            }
            // Error of the accumulator in its own (scaled) units, T = t r, C = -2 t c (|C| <= 2 tcm), both held as
            // hi + lo fp16 pairs: |T - (Th + Tl)| <= 2^-22 |T|, |C - (Ch + Cl)| <= 2^-22 |C|, the dropped Tl.Cl term
            // <= 2^-22 |T||C|: 3 * 2^-22 * 2 = 1.43e-6 |T| tcm, plus fp32 accumulation over 13 MMAs and the subnormal
            // floors of the lo parts (the absolute terms, as in rq_fwd_tc64_kernel).  Winner and rival may both be off
            // (2x); the factor below leaves a further 14x for the accumulation order of the tensor core.
            const float tcm = ts * sqrtf(s_cmax[l]);
            const float tol = 4e-5f * trn * tcm + 2e-6f * trn + 8e-6f;
            const float* cb_level = p.cb + (int64_t)l * p.K * D;
            float lead_lo = INFINITY, lead_hi = INFINITY, lead_d = 0.f;     // the leading chunk winner so far
            int lead_k = 0;
            bool lead_exact = false;                                         // lead_d holds its exact distance

            for (int ch = 0; ch < P.n_chunks; ++ch, ++step) {
                const int k0 = ch * KC;
                const int Kc = (p.K - k0 < KC) ? (p.K - k0) : KC;
                const float* cbl = cb_level + (int64_t)k0 * D;
                const float* sCCc = sCC + k0;
                if (!ptx::mbar_wait(&mbar_full[wg], (uint32_t)(step & 1))) timed_out = true;
                ptx::tc_fence_after_sync();
                if (wtid == 0) release_stage(step);

                // ---- scan: group minima (columns 16g..16g+15) and class minima (column mod 16) -----------------------
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
                // the accumulator is read out: the next chunk's MMAs (same A rows) go in under this chunk's re-scoring
                if (ch + 1 < P.n_chunks) warp_arrive(step + 1);
                else ptx::tc_fence_before_sync();
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
                const float m1 = fminf(m1a, m1b);
                const float thr = m1 + (tol + 8e-6f * fabsf(m1));
                const bool ambiguous = wild || (m2a <= thr) || (m2b <= thr);
                if (__any_sync(0xffffffffu, ambiguous)) {
                    // every code within the band lies in a group with accA <= thr AND a class with accB <= thr;
                    // re-score those pairs in fp32, ascending index, strict '<': the SIMT family's arithmetic.  The warp
                    // re-scores one ambiguous row at a time, its candidate pairs spread over the 32 lanes
                    // (rq_fwd_tc64_kernel); the row is parked in the warp's scratch row.
                    uint32_t gm = 0, cm = 0;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        gm |= (accA[i] <= thr) ? (1u << i) : 0u;
                        cm |= (accB[i] <= thr) ? (1u << i) : 0u;
                    }
                    if (wild) { gm = 0xFFFFu; cm = 0xFFFFu; }
                    for (unsigned todo = __ballot_sync(0xffffffffu, ambiguous); todo; todo &= todo - 1) {
                        const int src = __ffs(todo) - 1;
                        if (lane == src) {
#pragma unroll
                            for (int j = 0; j < D; j += 4)
                                *reinterpret_cast<float4*>(scratch + j) = make_float4(r[j], r[j + 1], r[j + 2], r[j + 3]);
                        }
                        __syncwarp();
                        const uint32_t g = __shfl_sync(0xffffffffu, gm, src), c = __shfl_sync(0xffffffffu, cm, src);
                        const float xs = __shfl_sync(0xffffffffu, xx, src);
                        const int nc = __popc(c), np = __popc(g) * nc;
                        float best = INFINITY;
                        int bk = __shfl_sync(0xffffffffu, bi, src);      // stays if nothing compares below +inf (NaN input)
                        for (int t = lane; t < np; t += 32) {
                            const int k = (int)__fns(g, 0, t / nc + 1) * 16 + (int)__fns(c, 0, t % nc + 1);
                            if (k >= Kc) continue;
                            const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)k * D);
                            float a = 0.f;
#pragma unroll
                            for (int j = 0; j < D / 4; ++j) {
                                const float4 cv = __ldg(c4 + j);
                                const float4 rv = *reinterpret_cast<const float4*>(scratch + 4 * j);
                                a = fmaf(rv.x, cv.x, a); a = fmaf(rv.y, cv.y, a);
                                a = fmaf(rv.z, cv.z, a); a = fmaf(rv.w, cv.w, a);
                            }
                            const float d = (xs + __ldg(sCCc + k)) - 2.f * a;
                            if (d < best) { best = d; bk = k; }         // a lane's candidates come in ascending index
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            const float od = __shfl_xor_sync(0xffffffffu, best, o);
                            const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
                            if (od < best || (od == best && ok < bk)) { best = od; bk = ok; }
                        }
                        if (lane == src) bi = bk;
                        __syncwarp();                                   // scratch is reused by the next row
                    }
                }
                if (bi >= Kc) bi = Kc - 1;      // unreachable unless inputs are NaN; keeps the gather in range
                // ---- running minimum over the chunks.  The chunks of a level share the scale t, so their tensor-core
                // scores are comparable: the score of this chunk's winner lies in [c_lo, c_hi] (the index packing, and the
                // band when the winner came out of the re-scoring), and two scores further apart than tol order the exact
                // distances (tol = twice the error bound of one score).  Only a near tie between the leader and the
                // newcomer is settled by their exact fp32 distances (the SIMT family's sequential fmaf chain, gathered
                // codewords), strict '<': the earlier chunk — the lower code index — keeps a tie.  A gather per chunk and
                // row, as the merge kernel needed, would make the L1 the busiest unit of the launch.
                {
                    const float c_lo = m1, c_hi = ambiguous ? thr : m1 + 8e-6f * fabsf(m1);
                    bool take = (ch == 0), close = false;
                    if (ch > 0) {
                        take = c_hi + tol < lead_lo;
                        const bool worse = c_lo > lead_hi + tol;
                        close = wild || (!take && !worse);
                        if (wild) take = false;
                    }
                    if (__any_sync(0xffffffffu, close)) {
                        if (close) {
                            auto exact_dist = [&](int k) {
                                const float4* c4 = reinterpret_cast<const float4*>(cb_level + (int64_t)k * D);
                                float a = 0.f;
#pragma unroll
                                for (int j = 0; j < D / 4; ++j) {
                                    const float4 cv = __ldg(c4 + j);
                                    a = fmaf(r[4 * j], cv.x, a); a = fmaf(r[4 * j + 1], cv.y, a);
                                    a = fmaf(r[4 * j + 2], cv.z, a); a = fmaf(r[4 * j + 3], cv.w, a);
                                }
                                return (xx + __ldg(sCC + k)) - 2.f * a;
                            };
                            if (!lead_exact) { lead_d = exact_dist(lead_k); lead_exact = true; }
                            const float d = exact_dist(k0 + bi);
                            if (d < lead_d) { lead_d = d; lead_k = k0 + bi; lead_lo = c_lo; lead_hi = c_hi; }
                        }
                    }
                    if (take) { lead_k = k0 + bi; lead_lo = c_lo; lead_hi = c_hi; lead_exact = false; }
                }
            }
            const int best_k = lead_k;
            // ---- id out; then gather, estimator, loss, outputs, residual update (as rq_fwd_tc64_kernel) ------------------
            const int bi = (best_k < p.K) ? best_k : p.K - 1;
            {
                int64_t* dst = p.ids + (int64_t)l * p.B + row;
                if (p.n_changed) {
                    const bool chg = active && (!p.has_prev || *dst != (int64_t)bi);
                    const unsigned m = __ballot_sync(0xffffffffu, chg);
                    if (lane == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
                }
                if (active) *dst = bi;
            }
            // ids-only launches (tokenisation, k-means assignment) need nothing past the id at the last level
            if (l == p.L - 1 && !p.emb && !p.loss) continue;
            float e[D];
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(cb_level + (int64_t)bi * D + j, &e[j]);
            loss_acc += level_epilogue<D>(p.mode, p.beta, r, e, xx, e);          // estimator value in place
            if (active && p.emb) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.emb + off + j, &e[j]);
            }
            residual_update<D>(r, e);                                      // rqvae.py:129
            xx = dot_self<D>(r);
        }
        if (active && p.loss) p.loss[row] = loss_acc;
        if (it + 1 < my_pairs) load_rows(it + 1);
    }

    if (timed_out && P.err) atomicExch(P.err, 1);
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

static int* d_err_bigk = nullptr;

int tc64_bigk_timeout_flag() {
    int v = 0;
    if (d_err_bigk) cudaMemcpy(&v, d_err_bigk, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}

int quantize_forward_tc64_bigk(const FwdParams& p, cudaStream_t st) {
    using namespace bigk;
    if (p.B == 0) return 0;
    const int n_chunks = (p.K + kBigkChunk - 1) / kBigkChunk;
    const BigkSmem S = bigk_smem_layout(p.K);
    if (p.L > BIGK_MAX_L || S.total + 1024 > 227 * 1024) return quantize_forward_tc64_bigk_chunked(p, st);
    if (!d_err_bigk) {
        RQ_CUDA(cudaMalloc((void**)&d_err_bigk, sizeof(int)));
        RQ_CUDA(cudaMemset(d_err_bigk, 0, sizeof(int)));
    }
    const size_t img_bytes = (size_t)p.L * n_chunks * STAGE_BYTES;
    const size_t cc_bytes = (size_t)p.L * n_chunks * KC * sizeof(float);
    uint8_t* scratch = (uint8_t*)engine_pool_alloc(img_bytes + cc_bytes + 2 * BIGK_MAX_L * sizeof(float), st);
    if (!scratch) {
        set_error("rq_quantize_forward: out of device memory for the K > 256 operand images");
        return (int)cudaErrorMemoryAllocation;
    }
    BigkPrepParams q;
    q.cb = p.cb; q.K = p.K; q.L = p.L; q.n_chunks = n_chunks;
    q.img = scratch;
    q.cc = reinterpret_cast<float*>(scratch + img_bytes);
    q.lvl = reinterpret_cast<float*>(scratch + img_bytes + cc_bytes);
    rq_bigk_prepare_kernel<<<p.L, 1024, 0, st>>>(q);
    RQ_LAUNCHED();
    BigkParams P;
    P.f = p;
    P.err = d_err_bigk;
    const int64_t n_tiles = (p.B + TILE_M - 1) / TILE_M;
    P.n_pairs = (int)((n_tiles + 1) / 2);
    P.n_chunks = n_chunks;
    P.img = q.img; P.cc = q.cc; P.lvl = q.lvl;
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc64_bigk_kernel, 227 * 1024));
    const int grid = P.n_pairs < sm_count() ? P.n_pairs : sm_count();
    rq_fwd_tc64_bigk_kernel<<<grid, THREADS, S.total + 1024, st>>>(P);
    RQ_LAUNCHED();
    cudaFreeAsync(scratch, st);
    return 0;
}

}  // namespace rq
