// tcgen05 family, D = 32: dispatch + the two-warpgroup kernel that serves L = 4 (the operand tiles of four levels
// leave no room for the residual tiles of the four-warpgroup kernel in rq_quantize_tc32.cu, which serves L <= 3).
//
// rq_fwd_tc32_l4_kernel: one persistent CTA per SM, two 128-thread warpgroups, each with its own TMEM accumulator
// (128x256 fp32) and A tile.  A thread owns one row for all L levels with the residual in registers (255-register
// budget at 256 threads); operand conversion, MMAs, scan and re-scoring are the ones of the four-warpgroup kernel
// (rq_tc_tiles.cuh), so ids are bit-identical to it and to the SIMT family.
#include "rq_rowmath.cuh"
#include "rq_tc_tiles.cuh"

namespace rq {

namespace tcl4 {
using namespace tc;
constexpr int NWG = 2;
constexpr int THREADS = 128 * NWG;
constexpr int MAX_L = 4;
struct SmemLayout {
    int b_tiles, a_tiles, cc_tiles, ones, cc_f32, misc, total;
};
__host__ __device__ inline SmemLayout smem_layout(int L) {
    SmemLayout s;
    int off = 0;
    s.b_tiles = off; off += L * B_TILE_BYTES;
    s.a_tiles = off; off += NWG * A_TILE_BYTES;
    s.cc_tiles = off; off += L * CC_TILE_BYTES;
    s.ones = off; off += ONES_TILE_BYTES;
    s.cc_f32 = off; off += L * KC * 4;
    s.misc = off; off += 256;
    s.total = off;
    return s;
}
}  // namespace tcl4

__global__ void __launch_bounds__(tcl4::THREADS, 1) rq_fwd_tc32_l4_kernel(TcParams P) {
    using namespace tcl4;
    const FwdParams& p = P.f;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    const SmemLayout S = smem_layout(p.L);
    uint8_t* sB = smem + S.b_tiles;
    uint8_t* sA = smem + S.a_tiles;
    uint8_t* sCCt = smem + S.cc_tiles;
    uint8_t* sOnes = smem + S.ones;
    float* sCC = reinterpret_cast<float*>(smem + S.cc_f32);
    uint64_t* mbar_full = reinterpret_cast<uint64_t*>(smem + S.misc);          // [2] accumulator ready (per warpgroup)
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(smem + S.misc + 64);
    int* cmax_bits = reinterpret_cast<int*>(smem + S.misc + 80);               // [MAX_L] max ||c||^2 per level
    uint64_t* mbar_load = reinterpret_cast<uint64_t*>(smem + S.misc + 128);    // [MAX_L] a level's bulk copy landed

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wg = tid >> 7, wtid = tid & 127;
    bool timed_out = false;

    if (tid == 0) {
        ptx::mbar_init(&mbar_full[0], 1);
        ptx::mbar_init(&mbar_full[1], 1);
#pragma unroll
        for (int i = 0; i < MAX_L; ++i) ptx::mbar_init(&mbar_load[i], 1);
        ptx::fence_mbar_init();
        const uint32_t level_bytes = (uint32_t)p.K * D * 4;
        for (int l = 0; l < p.L; ++l) {
            ptx::mbar_arrive_expect_tx(&mbar_load[l], level_bytes);
            ptx::bulk_g2s(sB + l * B_TILE_BYTES, p.cb + (int64_t)l * p.K * D, level_bytes, &mbar_load[l]);
        }
    }
    // tile t -> CTA t % grid, warpgroup (t / grid) % NWG
    const int my_tiles = (P.n_tiles > (int)blockIdx.x) ? (P.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const int n_mine = (my_tiles > wg) ? (my_tiles - 1 - wg) / NWG + 1 : 0;
    float r[D];
    auto load_x_row = [&](int it) {
        const int64_t row = ((int64_t)blockIdx.x + (int64_t)gridDim.x * (wg + NWG * it)) * TILE_M + wtid;
        if (it < n_mine && row < p.B) {
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(p.x + row * D + j, &r[j]);
        } else {
#pragma unroll
            for (int j = 0; j < D; ++j) r[j] = 0.f;
        }
    };
    load_x_row(0);                                     // in flight during the whole prologue
    if (tid < MAX_L) cmax_bits[tid] = 0;
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, 512);
    for (int i = tid; i < TILE_M * 2; i += THREADS) {  // ones tile: k = 0,1,2 -> 1.0, rest 0
        const int rr = i >> 1, c = i & 1;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (c == 0) { v.x = 0x3F803F80u; v.y = 0x00003F80u; }
        *reinterpret_cast<uint4*>(sOnes + nosw_off(rr, c)) = v;
    }
    __syncthreads();                                   // barrier inits + cmax zeros visible
    for (int l = 0; l < p.L; ++l) {                    // 1024 items per level: 4 per thread
        if (!ptx::mbar_wait(&mbar_load[l], 0)) timed_out = true;
        float mx = 0.f;
#pragma unroll 1
        for (int i = 0; i < 2; ++i)
            mx = fmaxf(mx, convert_codebook_items<2>(sB + l * B_TILE_BYTES, sCCt + l * CC_TILE_BYTES, sCC + l * KC, p.K,
                                                     tid + 2 * i * THREADS, THREADS));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) atomicMax(&cmax_bits[l], __float_as_int(mx));
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_s;
    const uint32_t tmem_acc = tmem_base + (uint32_t)(wg * KC);
    const uint32_t tmem_row = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16);
    uint8_t* sAw = sA + wg * A_TILE_BYTES;
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TILE_M, KC);
    int use = 0;

    for (int it = 0; it < n_mine; ++it) {
        const int tile = (int)blockIdx.x + (int)gridDim.x * (wg + NWG * it);
        const int64_t row = (int64_t)tile * TILE_M + wtid;
        const bool active = row < p.B;
        float xx = dot_self<D>(r);
        float loss_acc = 0.f;
        for (int l = 0; l < p.L; ++l, ++use) {
            // ---- A operand row: [hi(r) | lo(r)] (our own previous MMAs were awaited before the last scan) -----------
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t h[4], lw[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) ptx::split_pair(r[c * 8 + 2 * q], r[c * 8 + 2 * q + 1], h[q], lw[q]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, c)) = make_uint4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<uint4*>(sAw + sw128_off(wtid, 4 + c)) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            ptx::fence_proxy_async_smem();
            ptx::named_bar_sync(1 + wg, 128);
            if (wtid == 0) {
                ptx::tc_fence_after_sync();
                const uint32_t a0 = ptx::smem_u32(sAw), b0 = ptx::smem_u32(sB + l * B_TILE_BYTES);
                const int ak[6] = {0, 32, 64, 96, 0, 32};
                const int bk[6] = {0, 32, 0, 32, 64, 96};
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const uint64_t ad = ptx::make_smem_desc(a0 + ak[i], 16, 1024, ptx::kLayoutSw128);
                    const uint64_t bd = ptx::make_smem_desc(b0 + bk[i], 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, i > 0);
                }
                const uint64_t od = ptx::make_smem_desc(ptx::smem_u32(sOnes), 128, 256, ptx::kLayoutNone);
                const uint64_t cd = ptx::make_smem_desc(ptx::smem_u32(sCCt + l * CC_TILE_BYTES), 128, 256, ptx::kLayoutNone);
                ptx::umma_bf16(tmem_acc, od, cd, idesc, 1);
                ptx::umma_commit(&mbar_full[wg]);
            }
            const int64_t off = ((int64_t)l * p.B + row) * D;
            if (active && p.res) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.res + off + j, &r[j]);
            }
            const float cmax2 = __int_as_float(cmax_bits[l]);
            const float tol = 1e-4f * fast_sqrt(xx * cmax2) + 3e-6f * cmax2 + 1e-30f;      // see rq_quantize_tc32.cu
            if (!ptx::mbar_wait(&mbar_full[wg], use & 1)) timed_out = true;
            ptx::tc_fence_after_sync();
            float accA[16], accB[16];
            scan_accumulator_row(tmem_row, accA, accB);
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
            const bool ambiguous = (m2a <= thr) || (m2b <= thr);
            const float* cbl = p.cb + (int64_t)l * p.K * D;
            if (ambiguous) {                        // re-score the (group, class) pairs within the band, exact fp32
                uint32_t gm = 0, cm = 0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    gm |= (accA[i] <= thr) ? (1u << i) : 0u;
                    cm |= (accB[i] <= thr) ? (1u << i) : 0u;
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
                        const float d = (xx + sCC[l * KC + k]) - 2.f * a;
                        if (d < best) { best = d; bi = k; }
                    }
                }
            }
            if (bi >= p.K) bi = p.K - 1;        // unreachable unless inputs are NaN; keeps the gather in range
            {
                int64_t* dst = p.ids + (int64_t)l * p.B + row;
                if (p.n_changed) {
                    const bool ch = active && (!p.has_prev || *dst != (int64_t)bi);
                    const unsigned m = __ballot_sync(0xffffffffu, ch);
                    if (lane == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
                }
                if (active) *dst = bi;
            }
            if (l == p.L - 1 && !p.emb && !p.loss) continue;
            float e[D];
#pragma unroll
            for (int j = 0; j < D; j += 8) ptx::ld_global_v8(cbl + (int64_t)bi * D + j, &e[j]);
            loss_acc += level_epilogue<D>(p.mode, p.beta, r, e, xx, e);          // estimator value in place
            if (active && p.emb) {
#pragma unroll
                for (int j = 0; j < D; j += 8) ptx::st_global_v8(p.emb + off + j, &e[j]);
            }
            residual_update<D>(r, e);                                            // rqvae.py:129
            xx = dot_self<D>(r);
        }
        if (active && p.loss) p.loss[row] = loss_acc;
        load_x_row(it + 1);
    }

    if (timed_out && P.err) atomicExch(P.err, 1);
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

// ---- host side --------------------------------------------------------------------------------------------------
bool tc32_fused_supported(int D, int K, int L, int mode, int distance);
int quantize_forward_tc32_fused(const FwdParams& p, cudaStream_t st);
void tc32_set_accumulator_dump(float* buf);
int tc32_timeout_flag();

static int* d_err = nullptr;

bool tc32_supported(int D, int K, int L, int mode, int distance) {
    return D == tc::D && K >= 1 && K <= tc::KC && L >= 1 && L <= tcl4::MAX_L && distance == RQ_DIST_L2 &&
           (mode == RQ_MODE_EVAL || mode == RQ_MODE_STE || mode == RQ_MODE_ROTATION_TRICK);
}
bool tc_supported(int D, int K, int L, int mode, int distance) {
    return tc32_supported(D, K, L, mode, distance) || tc64_supported(D, K, L, mode, distance);
}
int quantize_forward_tc(const FwdParams& p, int D, cudaStream_t st) {
    // D = 64: the streamed-chunk kernel (rq_quantize_bigk.cu) also serves K <= 256 — one chunk per level, hi + lo fp16
    // operand images from L2 instead of hi-only tiles resident in shared memory: 2.82 against 2.49 G items/s on the
    // 4M-item ids-only tokenisation (tools/time_tok64.py), because the narrower band takes the re-scoring and its
    // warpgroup-barrier stalls out.  rq_fwd_tc64_kernel remains as the chunk scorer of the L > 8 fallback
    // (RQ_V_TC64_RESIDENT builds keep it as the K <= 256 kernel for A/B runs).
#ifndef RQ_V_TC64_RESIDENT
    if (D == 64) return quantize_forward_tc64_bigk(p, st);
#endif
    return D == 64 ? quantize_forward_tc64(p, st) : quantize_forward_tc32(p, st);
}

int quantize_forward_tc32(const FwdParams& p, cudaStream_t st) {
    if (p.B == 0) return 0;
    if (tc32_fused_supported(tc::D, p.K, p.L, p.mode, p.dist)) return quantize_forward_tc32_fused(p, st);
    if (!d_err) {
        RQ_CUDA(cudaMalloc((void**)&d_err, sizeof(int)));
        RQ_CUDA(cudaMemset(d_err, 0, sizeof(int)));
    }
    TcParams P = {};
    P.f = p;
    P.err = d_err;
    P.n_tiles = (int)((p.B + tc::TILE_M - 1) / tc::TILE_M);
    const tcl4::SmemLayout S = tcl4::smem_layout(p.L);
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_fwd_tc32_l4_kernel, tc::kMaxSmem));
    const int grid = P.n_tiles < sm_count() ? P.n_tiles : sm_count();
    rq_fwd_tc32_l4_kernel<<<grid, tcl4::THREADS, S.total + 1024, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

}  // namespace rq

// Test hooks (declared in include/rq_engine.h): the mbarrier-timeout flag and an accumulator dump.
extern "C" int rq_debug_tc_timeout_flag(void) {
    int v = 0;
    if (rq::d_err) cudaMemcpy(&v, rq::d_err, sizeof(int), cudaMemcpyDeviceToHost);
    return v | rq::tc32_timeout_flag() | rq::tc64_timeout_flag() | rq::tc64_bigk_timeout_flag() | rq::linear_tf32_timeout_flag();
}
extern "C" int rq_debug_tc_set_accumulator_dump(float* device_buf_128x256) {
    rq::tc32_set_accumulator_dump(device_buf_128x256);
    return 0;
}
