// Tile geometry and small device helpers shared by the D = 32 tcgen05 kernels (rq_quantize_tc32.cu: the
// four-warpgroup kernel of the headline shape; rq_quantize_tc.cu: the two-warpgroup variant for L = 4).
#pragma once
#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_tc_ptx.cuh"

namespace rq {

struct TcParams {
    FwdParams f;
    int* err;        // device flag: set when an mbarrier wait timed out
    int n_tiles;
    float* dbg_acc;  // optional [128][256] dump of tile 0 / level 0's accumulator (tests only)
#ifdef RQ_EXPERIMENTS
    int dbg_flags;      // what-if launches (wrong ids by construction): see tools/ab_flags.sh
    long long* dbg_tl;  // optional [32 warps][64 marks] clock64 timeline of CTA 0
#endif
};

namespace tc {
constexpr int D = 32;            // latent dim served by these kernels
constexpr int KC = 256;          // codes per level (UMMA N)
constexpr int TILE_M = 128;      // rows per tile (UMMA M)
constexpr int B_TILE_BYTES = KC * 128;       // [256 codes][-2hi(32) | -2lo(32)] bf16, SW128
constexpr int CC_TILE_BYTES = KC * 32;       // [256 codes][16 bf16], no swizzle, 8x16B core matrices
constexpr int A_TILE_BYTES = TILE_M * 128;   // [128 rows][hi(32) | lo(32)] bf16, SW128
constexpr int ONES_TILE_BYTES = TILE_M * 32; // [128 rows][1,1,1,0...] bf16, no swizzle
constexpr int kMaxSmem = 227 * 1024;

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
// (v & ~15) | i in ONE LOP3: the mask travels in a register (an immediate mask plus an immediate index would take two)
__device__ __forceinline__ float pack_idx(float v, int i) {
    uint32_t r, mask = 0xFFFFFFF0u;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(__float_as_uint(v)), "r"(mask), "r"((uint32_t)i));
    return __uint_as_float(r);
}

// One pass over a 256-column fp32 accumulator row in TMEM (this thread's lane, columns at `tmem_row`): minima of the
// 16 column groups (accA[g]: columns 16g..16g+15) and of the 16 column classes (accB[c]: column mod 16 == c).  The
// TMEM loads are double-buffered: chunk c+1 is in flight while chunk c is reduced.
__device__ __forceinline__ void scan_accumulator_row(uint32_t tmem_row, float (&accA)[16], float (&accB)[16]) {
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

// Converts items (code k, 8-element chunk q) of one level's codebook IN PLACE from the fp32 rows the bulk copy
// dropped into the B-tile space to the UMMA B operand [-2hi(c) | -2lo(c)] (bf16 split, 128-byte swizzled rows);
// the four lanes of a code (consecutive lanes of one warp) read its 128 bytes before any of them writes.  Also
// fills the level's fp32 ||c||^2 (canonical order of code_sqnorm), its [hi, mid, lo] bf16 tile, and returns
// this thread's max ||c||^2 over the real codes it handled (0 when none).  `item0 + u * stride`, u < U, must
// address whole quads inside one warp.
template <int U>
__device__ __forceinline__ float convert_codebook_items(uint8_t* tileB, uint8_t* tileCC, float* cc_f32, int K, int item0,
                                                        int stride) {
    float cv[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int i = item0 + u * stride;
        const int k = i >> 2, q = i & 3;
        const uint8_t* src = tileB + k * 128 + q * 32;
        if (k < K) {
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
#pragma unroll
    for (int u = 0; u < U; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], 1);
#pragma unroll
    for (int u = 0; u < U; ++u) part[u] += __shfl_xor_sync(0xffffffffu, part[u], 2);
    __syncwarp();                                           // the quad's reads precede its writes
    float mx = 0.f;
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const int i = item0 + u * stride;
        const int k = i >> 2, q = i & 3;
        *reinterpret_cast<uint4*>(tileB + sw128_off(k, q)) = make_uint4(h[u][0], h[u][1], h[u][2], h[u][3]);
        *reinterpret_cast<uint4*>(tileB + sw128_off(k, 4 + q)) = make_uint4(lw[u][0], lw[u][1], lw[u][2], lw[u][3]);
        if (q == 0) {
            const float cc = (k < K) ? part[u] : 1e30f;     // padding codes can never be the minimum
            if (k < K) mx = fmaxf(mx, cc);
            cc_f32[k] = cc;
            const float hh = ptx::bf16_round(cc), mm = ptx::bf16_round(cc - hh), ll = ptx::bf16_round((cc - hh) - mm);
            *reinterpret_cast<uint4*>(tileCC + nosw_off(k, 0)) =
                make_uint4(ptx::pack_bf16x2(hh, mm), ptx::pack_bf16x2(ll, 0.f), 0, 0);
            *reinterpret_cast<uint4*>(tileCC + nosw_off(k, 1)) = make_uint4(0, 0, 0, 0);
        }
    }
    return mx;
}

}  // namespace tc
}  // namespace rq
