// tcgen05 family: fused L-level residual quantiser on the 5th-gen tensor cores (sm_100a).
//
// Per CTA (one per SM, persistent over 128-row tiles; two 128-thread warpgroups, each owning one
// tile at a time and one 128x256 fp32 accumulator in TMEM — 2 x 256 = all 512 columns):
//
//   prologue   every level's codebook is converted once per CTA into the UMMA B operand
//              [-2*hi(c) | -2*lo(c)] (bf16 split, 128-byte swizzled rows, K-major) plus a small
//              [hi, mid, lo] split of ||c||^2 (no-swizzle tile) and kept in shared memory.
//   per level  each thread owns one row: it writes [hi(r) | lo(r)] of its residual as the A
//              operand row, one elected thread issues 7 tcgen05.mma (N=256, K=16 each):
//                 hi(r).hi(c) + lo(r).hi(c) + hi(r).lo(c)  (+ 1 . ||c||^2)
//              so the accumulator holds  ||c||^2 - 2 r.c  to ~2^-17 relative.  The epilogue reads
//              its row from TMEM twice: pass 1 takes the row minimum (FMNMX3), pass 2 flags every
//              code within a proven error bound of it.  A single flagged code (the usual case) is
//              the argmin; several are re-scored in fp32 with exactly the arithmetic of the SIMT
//              family, lowest index winning ties — so ids are bit-identical to the SIMT kernel.
//              Then: gather the fp32 codeword, mode epilogue (STE / rotation trick / eval), loss,
//              outputs, residual update in registers — the next level never touches HBM.
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
constexpr int THREADS = 256;     // two warpgroups
constexpr int B_TILE_BYTES = KC * 128;       // [256 codes][-2hi(32) | -2lo(32)] bf16, SW128
constexpr int CC_TILE_BYTES = KC * 32;       // [256 codes][16 bf16], no swizzle, 8x16B core matrices
constexpr int A_TILE_BYTES = TILE_M * 128;   // [128 rows][hi(32) | lo(32)] bf16, SW128
constexpr int ONES_TILE_BYTES = TILE_M * 32; // [128 rows][1,1,1,0...] bf16, no swizzle

struct SmemLayout {
    int b_tiles, cc_tiles, cc_f32, a_tiles, ones, misc, total;
};
__host__ __device__ inline SmemLayout smem_layout(int L) {
    SmemLayout s;
    int off = 0;
    s.b_tiles = off; off += L * B_TILE_BYTES;
    s.a_tiles = off; off += 2 * A_TILE_BYTES;
    s.cc_tiles = off; off += L * CC_TILE_BYTES;
    s.ones = off; off += ONES_TILE_BYTES;
    s.cc_f32 = off; off += L * KC * 4;
    s.misc = off; off += 128;
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
    int* err;        // device flag: set when an mbarrier wait timed out
    int n_tiles;
    float* dbg_acc;  // optional [128][256] dump of tile 0 / level 0's accumulator (tests only)
};

__global__ void __launch_bounds__(tc::THREADS, 1) rq_fwd_tc_kernel(TcParams P) {
    using namespace tc;
    const FwdParams& p = P.f;
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled UMMA tiles need 1024-byte alignment; the launch over-allocates by 1 KB
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const SmemLayout S = smem_layout(p.L);
    uint8_t* sB = smem + S.b_tiles;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sCCt = smem + S.cc_tiles;
    uint8_t* sOnes = smem + S.ones;
    float* sCC = reinterpret_cast<float*>(smem + S.cc_f32);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + S.misc);          // [2]
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 16);
    int* cmax_bits = reinterpret_cast<int*>(smem + S.misc + 32);           // [MAX_L] max ||c||^2 per level

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wg = tid >> 7, wtid = tid & 127;

    // ---- prologue: barriers, TMEM, operand tiles ---------------------------------------------------
    if (tid == 0) {
        ptx::mbar_init(&mbar[0], 1);
        ptx::mbar_init(&mbar[1], 1);
        ptx::fence_mbar_init();
    }
    if (tid < MAX_L) cmax_bits[tid] = 0;
    if (warp == 0) ptx::tmem_alloc(tmem_base_s, 512);
    // ones tile: k = 0,1,2 -> 1.0 (multiplies hi/mid/lo of ||c||^2), rest 0
    for (int i = tid; i < TILE_M * 2; i += THREADS) {
        int r = i >> 1, c = i & 1;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0) { v.x = 0x3F803F80u; v.y = 0x00003F80u; }
        *reinterpret_cast<uint4*>(sOnes + nosw_off(r, c)) = v;
    }
    __syncthreads();
    // B operand: work item = (level, code, 8-element chunk); every thread first issues ALL of its
    // 32-byte codebook loads (the prologue is pure latency otherwise), then converts.
    {
        constexpr int ITEMS = MAX_L * 4;          // L * 256 codes * 4 chunks / 256 threads
        float cv[ITEMS][8];
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int i = it * THREADS + tid, l = i >> 10, k = (i >> 2) & 255, q = i & 3;
            if (l < p.L && k < p.K) {
                ptx::ld_global_v8(p.cb + ((int64_t)l * p.K + k) * D + q * 8, cv[it]);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) cv[it][j] = 0.f;
            }
        }
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int i = it * THREADS + tid, l = i >> 10, k = (i >> 2) & 255, q = i & 3;
            if (l >= p.L) break;                   // uniform across the CTA (l = it >> 2)
            uint32_t h[4], lw[4];
            float part = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float a = cv[it][2 * j], b = cv[it][2 * j + 1];
                part = fmaf(a, a, part);
                part = fmaf(b, b, part);
                ptx::split_pair(-2.f * a, -2.f * b, h[j], lw[j]);
            }
            uint8_t* tileB = sB + l * B_TILE_BYTES;
            *reinterpret_cast<uint4*>(tileB + sw128_off(k, q)) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(tileB + sw128_off(k, 4 + q)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            // ||c||^2 in the canonical order shared with the SIMT family (code_sqnorm in rq_rowmath.cuh):
            // 8-wide sequential partials, then (p0+p1)+(p2+p3)
            part += __shfl_xor_sync(0xffffffffu, part, 1);
            part += __shfl_xor_sync(0xffffffffu, part, 2);
            if (q == 0) {
                float cc = (k < p.K) ? part : 1e30f;     // padding codes can never be the minimum
                sCC[l * KC + k] = cc;
                float hh = ptx::bf16_round(cc), mm = ptx::bf16_round(cc - hh), ll = ptx::bf16_round((cc - hh) - mm);
                *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 0)) =
                    make_uint4(ptx::pack_bf16x2(hh, mm), ptx::pack_bf16x2(ll, 0.f), 0, 0);
                *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 1)) = make_uint4(0, 0, 0, 0);
                if (k < p.K) atomicMax(&cmax_bits[l], __float_as_int(cc));
            }
        }
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_s;
    const uint32_t tmem_acc = tmem_base + (uint32_t)(wg * KC);                       // column offset
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);        // this warp's lanes
    uint8_t* sAw = sA + wg * A_TILE_BYTES;
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TILE_M, KC);
    uint32_t phase = 0;
    bool timed_out = false;

    for (int tile = blockIdx.x * 2 + wg; tile < P.n_tiles; tile += gridDim.x * 2) {
        const int64_t row = (int64_t)tile * TILE_M + wtid;
        const bool active = row < p.B;
        float r[D];
        if (active) {
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + row * D + j, &r[j]);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
        float loss_acc = 0.f;

        for (int l = 0; l < p.L; ++l) {
            // ---- A operand row: [hi(r) | lo(r)] ------------------------------------------------------
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t h[4], lw[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) ptx::split_pair(r[c * 8 + 2 * q], r[c * 8 + 2 * q + 1], h[q], lw[q]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, c)) = make_uint4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, 4 + c)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before_sync();        // orders this thread's earlier tcgen05.ld before the barrier
            ptx::named_bar_sync(1 + wg, 128);
            if (wtid == 0) {
                ptx::tc_fence_after_sync();
                const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sB + l * B_TILE_BYTES);
                // (A k-offset bytes, B k-offset bytes): hi.hi, hi.hi, lo.hi, lo.hi, hi.lo, hi.lo
                const int ak[6] = {0, 32, 64, 96, 0, 32};
                const int bk[6] = {0, 32, 0, 32, 64, 96};
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    uint64_t ad = ptx::make_smem_desc(a0 + ak[i], 16, 1024, ptx::kLayoutSw128);
                    uint64_t bd = ptx::make_smem_desc(b0 + bk[i], 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);
                }
                uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
                uint64_t cd = ptx::make_smem_desc(ptx::smem_u32(sCCt + l * CC_TILE_BYTES), 128, 256, ptx::kLayoutNone);
                ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
                ptx::umma_commit(&mbar[wg]);
            }
            // scalars needed by the epilogue while the MMAs run
            const float xx = dot_self<D>(r);
            const float cmax2 = __int_as_float(cmax_bits[l]);
            // |acc - exact| <= 2*(2^-18 + 2*2^-17)|r||c| = 3.8e-5 |r||c| (dropped lo.lo term + the two
            // truncated lo parts) + fp32 accumulation noise; both the winner and a rival may be off,
            // hence 2x, with margin ("error bound")
            const float tol = 1e-4f * sqrtf(xx * cmax2) + 3e-6f * cmax2 + 1e-30f;
            if (!ptx::mbar_wait(&mbar[wg], phase)) timed_out = true;
            phase ^= 1;
            ptx::tc_fence_after_sync();

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
            // ---- pass 1: row minimum of the accumulator ------------------------------------------------
            // Both passes double-buffer the TMEM reads: chunk c+1 is in flight while chunk c is reduced.
            uint32_t va[32], vb[32];
            float m0 = INFINITY, m1 = INFINITY;
            ptx::tmem_ld_32x32b_x32(tmem_row, va);
            ptx::tmem_ld_wait_for(va);
#pragma unroll
            for (int c = 0; c < KC / 32; c += 2) {
                ptx::tmem_ld_32x32b_x32(tmem_row + (c + 1) * 32, vb);
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    m0 = ptx::fmin3(m0, __uint_as_float(va[j]), __uint_as_float(va[j + 1]));
                    m1 = ptx::fmin3(m1, __uint_as_float(va[j + 2]), __uint_as_float(va[j + 3]));
                }
                ptx::tmem_ld_wait_for(vb);
                // the last prefetch already fetches chunk 0 again for pass 2
                ptx::tmem_ld_32x32b_x32(tmem_row + ((c + 2) & (KC / 32 - 1)) * 32, va);
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    m0 = ptx::fmin3(m0, __uint_as_float(vb[j]), __uint_as_float(vb[j + 1]));
                    m1 = ptx::fmin3(m1, __uint_as_float(vb[j + 2]), __uint_as_float(vb[j + 3]));
                }
                ptx::tmem_ld_wait_for(va);
            }
            const float thr = fminf(m0, m1) + tol;
            // ---- pass 2: flag every code with acc <= thr ---------------------------------------------
            uint32_t cand[KC / 32];
#pragma unroll
            for (int c = 0; c < KC / 32; c += 2) {
                ptx::tmem_ld_32x32b_x32(tmem_row + (c + 1) * 32, vb);
                uint32_t msk = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j)      // sign(thr - v) = 1 <=> v > thr ; column j ends at bit 31-j
                    msk = __funnelshift_l(__float_as_uint(thr - __uint_as_float(va[j])), msk, 1);
                cand[c] = ~msk;
                ptx::tmem_ld_wait_for(vb);
                if (c + 2 < KC / 32) ptx::tmem_ld_32x32b_x32(tmem_row + (c + 2) * 32, va);
                msk = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    msk = __funnelshift_l(__float_as_uint(thr - __uint_as_float(vb[j])), msk, 1);
                cand[c + 1] = ~msk;
                if (c + 2 < KC / 32) ptx::tmem_ld_wait_for(va);
            }
            int n_cand = 0;
#pragma unroll
            for (int c = 0; c < KC / 32; ++c) n_cand += __popc(cand[c]);
            int bi = 0;
            const float* cbl = p.cb + (int64_t)l * p.K * D;
            if (n_cand == 1) {
#pragma unroll
                for (int c = 0; c < KC / 32; ++c)
                    if (cand[c]) bi = c * 32 + __clz(cand[c]);
            } else {
                // fp32 re-score in ascending index order, strict '<': exactly the SIMT family's arithmetic
                float best = INFINITY;
#pragma unroll 1
                for (int c = 0; c < KC / 32; ++c) {
                    uint32_t mk = cand[c];
                    while (mk) {
                        int pos = __clz(mk);
                        mk &= ~(0x80000000u >> pos);
                        int k = c * 32 + pos;
                        if (k >= p.K) continue;
                        const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)k * D);
                        float a = 0.f;
#pragma unroll
                        for (int j = 0; j < D / 4; ++j) {
                            float4 cv = __ldg(c4 + j);
                            a = fmaf(r[4 * j], cv.x, a); a = fmaf(r[4 * j + 1], cv.y, a);
                            a = fmaf(r[4 * j + 2], cv.z, a); a = fmaf(r[4 * j + 3], cv.w, a);
                        }
                        float d = (xx + sCC[l * KC + k]) - 2.f * a;
                        if (d < best) { best = d; bi = k; }
                    }
                }
            }
            if (bi >= p.K) bi = p.K - 1;        // unreachable unless inputs are NaN; keeps the gather in range

            // ---- gather, estimator, loss, outputs, residual update ---------------------------------------
            float e[D], o[D];
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(cbl + (int64_t)bi * D + j, &e[j]);
            loss_acc += level_epilogue<D>(p.mode, p.beta, r, e, o);
            if (active) {
                const int64_t off = ((int64_t)l * p.B + row) * D;
                if (p.res) {
#pragma unroll
                    for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.res + off + j, &r[j]);
                }
                if (p.emb) {
#pragma unroll
                    for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.emb + off + j, &o[j]);
                }
                p.ids[(int64_t)l * p.B + row] = bi;
            }
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = r[j] - o[j];
        }
        if (active && p.loss) p.loss[row] = loss_acc;
    }

    if (timed_out && P.err) atomicExch(P.err, 1);
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tmem_base, 512);
}

bool tc_supported(int D, int K, int L, int mode, int distance) {
    if (getenv("RQ_DISABLE_TCGEN05")) return false;
    return D == tc::D && K >= 1 && K <= tc::KC && L >= 1 && L <= tc::MAX_L && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

static int* d_err = nullptr;
static float* g_dbg_acc = nullptr;

int quantize_forward_tc(const FwdParams& p, int D, cudaStream_t st) {
    (void)D;
    if (p.B == 0) return 0;
    if (!d_err) {
        RQ_CUDA(cudaMalloc((void**)&d_err, sizeof(int)));
        RQ_CUDA(cudaMemset(d_err, 0, sizeof(int)));
    }
    TcParams P;
    P.f = p;
    P.err = d_err;
    P.dbg_acc = g_dbg_acc;
    P.n_tiles = (int)((p.B + tc::TILE_M - 1) / tc::TILE_M);
    const tc::SmemLayout S = tc::smem_layout(p.L);
    static bool attr_set = false;
    if (!attr_set) {
        RQ_CUDA(cudaFuncSetAttribute(rq_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    int grid = (P.n_tiles + 1) / 2;
    if (grid > sm_count()) grid = sm_count();
    rq_fwd_tc_kernel<<<grid, tc::THREADS, S.total + 1024, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

}  // namespace rq

// Test hooks (declared in include/rq_engine.h): the mbarrier-timeout flag and an accumulator dump.
extern "C" int rq_debug_tc_timeout_flag(void) {
    int v = 0;
    if (rq::d_err) cudaMemcpy(&v, rq::d_err, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}
extern "C" int rq_debug_tc_set_accumulator_dump(float* device_buf_128x256) {
    rq::g_dbg_acc = device_buf_128x256;
    return 0;
}

namespace rq {

}  // namespace rq
