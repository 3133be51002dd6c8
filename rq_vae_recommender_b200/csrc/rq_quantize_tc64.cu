// tcgen05 family, D = 64 (ml32m-shape latents: configs/rqvae_ml32m.gin of the reference, d=64, 3x256).
//
// Same plan as rq_quantize_tc.cu — persistent CTA per SM, one thread = one row = one TMEM lane, the
// residual in registers across the L levels, single-pass scan with group/class minima, exact fp32
// re-scoring of every code inside the error band so that ids are bit-identical to the SIMT family —
// but the operand format is chosen for the shared-memory budget of 64-wide rows:
//
//   * operands are FP16 with a per-level power-of-two scale t (t*max||c|| in [0.5,1)): the B tile of a
//     level is [-2 t c] rounded to fp16 (one 128-byte swizzled row per code, 32 KB per level, 96 KB
//     for L = 3), the A tile is [hi(t r) | lo(t r)] (two 64-element K-blocks, 32 KB per warpgroup).
//     Eight tcgen05.mma (K=16) per level: hi(tr).(-2tc) over 64 dims, lo(tr).(-2tc) over 64 dims,
//     plus one 1 x [t^2 ||c||^2 split in three fp16] — the accumulator holds t^2 (||c||^2 - 2 r.c)
//     up to the rounding of c: |err| <= 2 |tr| 2^-12 |tc| (+ subnormal terms), see `tol` below.
//     (bf16 operands would need the lo(c) tile as well — 64 KB per level, L = 3 does not fit — or an
//     8x wider band; fp16's 11-bit significand keeps the band at ~1e-3 |r||c|.)
//   * two warpgroups per CTA, each with its own TMEM accumulator (no sharing protocol needed); 255
//     registers per thread hold the 64-wide row through the scan.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"
#include "rq_tc_ptx.cuh"
#include "rq_tc64_tiles.cuh"

namespace rq {


struct Tc64Params {
    FwdParams f;
    int* err;
    int n_tiles;
};

__global__ void __launch_bounds__(tc64::THREADS, 1) rq_fwd_tc64_kernel(Tc64Params P) {
    using namespace tc64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const SmemLayout S = smem_layout(P.f.L);
    uint8_t* sB = smem + S.b_tiles;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sCCt = smem + S.cc_tiles;
    uint8_t* sOnes = smem + S.ones;
    float* sCC = reinterpret_cast<float*>(smem + S.cc_f32);
    uint64_t* mbar_full = reinterpret_cast<uint64_t*>(smem + S.misc);        // [2] accumulator ready
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 16);
    float* s_cmax = reinterpret_cast<float*>(smem + S.misc + 32);            // [MAX_L] max ||c||^2
    float* s_scale = reinterpret_cast<float*>(smem + S.misc + 48);           // [MAX_L] t (power of two)

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wg = tid >> 7, wtid = tid & 127;
    // chunked launch (rq_quantize_bigk.cu): this CTA scores every row against ONE 256-code chunk of the codebook
    const int n_chunks = P.f.n_chunks, chunk = (int)blockIdx.x % n_chunks;
    const int cta = (int)blockIdx.x / n_chunks, n_cta = (int)gridDim.x / n_chunks;   // position among the chunk's CTAs
    FwdParams p = P.f;
    if (n_chunks > 1) {
        p.cb = P.f.cb + (int64_t)chunk * KC * D;
        p.K = (P.f.K - chunk * KC < KC) ? (P.f.K - chunk * KC) : KC;
        p.ids = P.f.ids + (int64_t)chunk * P.f.B;
        p.best_dist = P.f.best_dist ? P.f.best_dist + (int64_t)chunk * P.f.B : nullptr;
    }

    if (tid == 0) {
        ptx::mbar_init(&mbar_full[0], 1);
        ptx::mbar_init(&mbar_full[1], 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, 512);

    // ---- first tile's rows: loads in flight during the prologue ------------------------------------------
    const int my_tiles = (P.n_tiles > cta) ? (P.n_tiles - 1 - cta) / n_cta + 1 : 0;
    const int n_mine = (my_tiles > wg) ? (my_tiles - 1 - wg) / NWG + 1 : 0;
    float r[D];
    {
        const int64_t row = ((int64_t)cta + (int64_t)n_cta * wg) * TILE_M + wtid;
        if (n_mine > 0 && row < p.B) {
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + row * D + j, &r[j]);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
    }
    for (int i = tid; i < TILE_M * 2; i += THREADS) {          // ones tile: k = 0,1,2 -> 1.0
        int rr = i >> 1, c = i & 1;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0) { v.x = 0x3C003C00u; v.y = 0x00003C00u; }
        *reinterpret_cast<uint4*>(sOnes + nosw_off(rr, c)) = v;
    }

    // ---- codebooks, pass 1: ||c||^2 in the canonical order of code_sqnorm<64> (8-wide sequential partials,
    // balanced tree = xor-shuffles 1, 2, 4 over the 8 lanes of a code) ------------------------------------------
    const int n_items = p.L * KC * 8;                              // (level, code, 8-element chunk)
    for (int i = tid; i < n_items; i += THREADS) {
        const int l = i >> 11, k = (i >> 3) & 255, q = i & 7;
        float part = 0.f;
        if (k < p.K) {
            const float4* src = reinterpret_cast<const float4*>(p.cb + ((int64_t)l * p.K + k) * D + q * 8);
            const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
            part = fmaf(v0.x, v0.x, part); part = fmaf(v0.y, v0.y, part);
            part = fmaf(v0.z, v0.z, part); part = fmaf(v0.w, v0.w, part);
            part = fmaf(v1.x, v1.x, part); part = fmaf(v1.y, v1.y, part);
            part = fmaf(v1.z, v1.z, part); part = fmaf(v1.w, v1.w, part);
        }
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        part += __shfl_xor_sync(0xffffffffu, part, 2);
        part += __shfl_xor_sync(0xffffffffu, part, 4);
        if (q == 0) sCC[l * KC + k] = (k < p.K) ? part : 1e30f;
    }
    __syncthreads();
    if (warp < p.L) {
        float m = 0.f;
        for (int k = lane; k < p.K; k += 32) m = fmaxf(m, sCC[warp * KC + k]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) {
            s_cmax[warp] = m;
            // t = 2^-(floor(log2 max||c||) + 1): t * max||c|| in [0.5, 1); exponent clamped so t stays normal
            int e = (int)((__float_as_uint(sqrtf(m)) >> 23) & 255u) - 127;
            e = (m > 0.f) ? max(-100, min(100, e)) : -1;
            s_scale[warp] = __int_as_float((127 - (e + 1)) << 23);
        }
    }
    __syncthreads();
    // ---- pass 2: B tile rows [-2 t c] in fp16, and the [hi, mid, lo] fp16 split of t^2 ||c||^2 ------------------
    for (int i = tid; i < n_items; i += THREADS) {
        const int l = i >> 11, k = (i >> 3) & 255, q = i & 7;
        const float m2t = -2.f * s_scale[l];
        uint4 w = make_uint4(0, 0, 0, 0);
        if (k < p.K) {
            const float4* src = reinterpret_cast<const float4*>(p.cb + ((int64_t)l * p.K + k) * D + q * 8);
            const float4 v0 = __ldg(src), v1 = __ldg(src + 1);
            w.x = pack_h2(m2t * v0.x, m2t * v0.y); w.y = pack_h2(m2t * v0.z, m2t * v0.w);
            w.z = pack_h2(m2t * v1.x, m2t * v1.y); w.w = pack_h2(m2t * v1.z, m2t * v1.w);
        }
        *reinterpret_cast<uint4*>(sB + l * B_TILE_BYTES + sw128_off(k, q)) = w;
        if (q == 0) {
            uint4 c0 = make_uint4(0, 0, 0, 0);
            if (k < p.K) {
                const float v = s_scale[l] * s_scale[l] * sCC[l * KC + k];         // <= 1
                const float hh = h_round(v), mm = h_round(v - hh), ll = h_round((v - hh) - mm);
                c0.x = pack_h2(hh, mm); c0.y = pack_h2(ll, 0.f);
            } else {
                c0.x = pack_h2(60000.f, 0.f);        // padding codes: larger than any real score, finite
            }
            *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 0)) = c0;
            *reinterpret_cast<uint4*>(sCCt + l * CC_TILE_BYTES + nosw_off(k, 1)) = make_uint4(0, 0, 0, 0);
        }
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();

    const uint32_t tmem_base = *tmem_base_s;
    const uint32_t tmem_acc = tmem_base + (uint32_t)(wg * KC);
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);
    uint8_t* sAw = sA + wg * A_TILE_BYTES;
    const uint32_t idesc = make_idesc_f16_f32(TILE_M, KC);
    bool timed_out = false;
    int use = 0;

    for (int it = 0; it < n_mine; ++it) {
        const int tile = cta + n_cta * (wg + NWG * it);
        const int64_t row = (int64_t)tile * TILE_M + wtid;
        const bool active = row < p.B;
        float xx = dot_self<D>(r);
        float loss_acc = 0.f;

        for (int l = 0; l < p.L; ++l, ++use) {
            const float ts = s_scale[l];
            const float trn = ts * sqrtf(xx);
            // |t r| >= 2^14 (the row dwarfs every codeword by > 4 orders of magnitude) would overflow fp16:
            // such a row skips the tensor-core score and is re-scored against all codes in fp32 below
            const bool wild = !(trn < 16384.f);
            // ---- A operand row: [hi(t r) | lo(t r)], two 128-byte swizzled K-blocks ------------------------
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
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before_sync();            // our tcgen05.ld of the previous level precede the next MMA
            ptx::named_bar_sync(1 + wg, 128);
            if (wtid == 0) {
                ptx::tc_fence_after_sync();
                const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sB + l * B_TILE_BYTES);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    // i < 4: hi block, i >= 4: lo block; both against the same four K-steps of the B row
                    uint64_t ad = ptx::make_smem_desc(a0 + (i >> 2) * A_BLOCK_BYTES + (i & 3) * 32, 16, 1024, ptx::kLayoutSw128);
                    uint64_t bd = ptx::make_smem_desc(b0 + (i & 3) * 32, 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);     // kind::f16; the descriptor says fp16
                }
                uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
                uint64_t cd = ptx::make_smem_desc(ptx::smem_u32(sCCt + l * CC_TILE_BYTES), 128, 256, ptx::kLayoutNone);
                ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
                ptx::umma_commit(&mbar_full[wg]);
            }
            // ---- while the MMAs run: this level's input residual goes out ----------------------------------
            const int64_t off = ((int64_t)l * p.B + row) * D;
            if (active && p.res) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.res + off + j, &r[j]);
            }
            // Error of the accumulator in its own (scaled) units, T = t r, C = t c, |C| <= tcm < 1:
            //   rounding of -2tc to fp16:           2 |T| (2^-12 |C| + 8 * 2^-25)
            //   hi+lo representation of T:          2 |C| (2^-24 |T| + 8 * 2^-25)
            //   t^2||c||^2 split, fp32 accumulation: < 4e-6
            // winner and rival may both be off: 2x.  2^-10 = 9.77e-4.
            const float tcm = ts * sqrtf(s_cmax[l]);
            const float tol = 1.07e-3f * trn * tcm + 2e-6f * trn + 8e-6f;
            if (!ptx::mbar_wait(&mbar_full[wg], use & 1)) timed_out = true;
            ptx::tc_fence_after_sync();

            // ---- scan: group minima (columns 16g..16g+15) and class minima (column mod 16) -------------------
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
            ptx::tc_fence_before_sync();
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
            const float* cbl = p.cb + (int64_t)l * p.K * D;
            if (__any_sync(0xffffffffu, ambiguous)) {
                // every code within the band lies in a group with accA <= thr AND a class with accB <= thr;
                // re-score those pairs in fp32, ascending index, strict '<': the SIMT family's arithmetic
                uint32_t gm = 0, cm = 0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    gm |= (accA[i] <= thr) ? (1u << i) : 0u;
                    cm |= (accB[i] <= thr) ? (1u << i) : 0u;
                }
                if (wild) { gm = 0xFFFFu; cm = 0xFFFFu; }
                // The warp re-scores one ambiguous row at a time, its candidate pairs spread over the 32 lanes
                // (a lane-serial loop would hold the other 31 rows — and through the warpgroup barrier the other
                // three warps — for one 64-long FMA chain per candidate).  The row is parked in the warp's own,
                // currently idle, A-tile rows; every lane runs the same sequential fmaf chain as the SIMT family
                // on its candidate, then a lexicographic (distance, index) minimum picks the lowest index among
                // equals — the result of the serial ascending scan.
                float* scratch = reinterpret_cast<float*>(sAw + (warp & 3) * 32 * 128);
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
                    int bk = __shfl_sync(0xffffffffu, bi, src);          // stays if nothing compares below +inf (NaN input)
                    for (int t = lane; t < np; t += 32) {
                        const int k = (int)__fns(g, 0, t / nc + 1) * 16 + (int)__fns(c, 0, t % nc + 1);
                        if (k >= p.K) continue;
                        const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)k * D);
                        float a = 0.f;
#pragma unroll
                        for (int j = 0; j < D / 4; ++j) {
                            const float4 cv = __ldg(c4 + j);
                            const float4 rv = *reinterpret_cast<const float4*>(scratch + 4 * j);
                            a = fmaf(rv.x, cv.x, a); a = fmaf(rv.y, cv.y, a);
                            a = fmaf(rv.z, cv.z, a); a = fmaf(rv.w, cv.w, a);
                        }
                        const float d = (xs + sCC[l * KC + k]) - 2.f * a;
                        if (d < best) { best = d; bk = k; }             // a lane's candidates come in ascending index
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const float od = __shfl_xor_sync(0xffffffffu, best, o);
                        const int ok = __shfl_xor_sync(0xffffffffu, bk, o);
                        if (od < best || (od == best && ok < bk)) { best = od; bk = ok; }
                    }
                    if (lane == src) bi = bk;
                    __syncwarp();                                       // scratch is reused by the next row
                }
            }
            if (bi >= p.K) bi = p.K - 1;        // unreachable unless inputs are NaN; keeps the gather in range

            // ---- id out; then gather, estimator, loss, outputs, residual update ---------------------------------
            {
                int64_t* dst = p.ids + (int64_t)l * p.B + row;
                if (p.n_changed) {
                    const bool ch = active && (!p.has_prev || *dst != (int64_t)bi);
                    const unsigned m = __ballot_sync(0xffffffffu, ch);
                    if (lane == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
                }
                if (active) *dst = bi;
            }
            if (p.best_dist && l == p.L - 1) {
                // exact distance of the winner, the SIMT family's arithmetic (sequential fmaf chain)
                const float4* c4 = reinterpret_cast<const float4*>(cbl + (int64_t)bi * D);
                float a = 0.f;
#pragma unroll
                for (int j = 0; j < D / 4; ++j) {
                    const float4 cv = __ldg(c4 + j);
                    a = fmaf(r[4 * j], cv.x, a); a = fmaf(r[4 * j + 1], cv.y, a);
                    a = fmaf(r[4 * j + 2], cv.z, a); a = fmaf(r[4 * j + 3], cv.w, a);
                }
                if (active) p.best_dist[row] = (xx + sCC[l * KC + bi]) - 2.f * a;
            }
            // ids-only launches (tokenisation, k-means assignment) need nothing past the id at the last level
            if (l == p.L - 1 && !p.emb && !p.loss) continue;
            float e[D];
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(cbl + (int64_t)bi * D + j, &e[j]);
            loss_acc += level_epilogue<D>(p.mode, p.beta, r, e, xx, e);          // estimator value in place
            if (active && p.emb) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.emb + off + j, &e[j]);
            }
            residual_update<D>(r, e);                                      // rqvae.py:129
            xx = dot_self<D>(r);
        }
        if (active && p.loss) p.loss[row] = loss_acc;
        if (it + 1 < n_mine) {
            const int64_t nrow = ((int64_t)cta + (int64_t)n_cta * (wg + NWG * (it + 1))) * TILE_M + wtid;
            if (nrow < p.B) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + nrow * D + j, &r[j]);
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
}

bool tc64_supported(int D, int K, int L, int mode, int distance) {
    return D == tc64::D && K >= 1 && K <= tc64::KC && L >= 1 && L <= tc64::MAX_L && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}

static int* d_err64 = nullptr;

int tc64_timeout_flag() {
    int v = 0;
    if (d_err64) cudaMemcpy(&v, d_err64, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}

int quantize_forward_tc64(const FwdParams& p, cudaStream_t st) {
    if (p.B == 0) return 0;
    if (!d_err64) {
        RQ_CUDA(cudaMalloc((void**)&d_err64, sizeof(int)));
        RQ_CUDA(cudaMemset(d_err64, 0, sizeof(int)));
    }
    Tc64Params P;
    P.f = p;
    P.err = d_err64;
    P.n_tiles = (int)((p.B + tc64::TILE_M - 1) / tc64::TILE_M);
    const tc64::SmemLayout S = tc64::smem_layout(p.L);
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc64_kernel, 227 * 1024));
    int grid = P.n_tiles < sm_count() ? P.n_tiles : sm_count();
    if (p.n_chunks > 1) {
        // equal share of the SMs per chunk (at least one CTA each)
        int per = sm_count() / p.n_chunks;
        if (per < 1) per = 1;
        if (per > P.n_tiles) per = P.n_tiles;
        grid = per * p.n_chunks;
    }
    rq_fwd_tc64_kernel<<<grid, tc64::THREADS, S.total + 1024, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

}  // namespace rq
