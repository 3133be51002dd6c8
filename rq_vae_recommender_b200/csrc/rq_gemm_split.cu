// fp32-grade GEMM on the 5th-gen tensor cores: C[M, N] (+)= alpha * A[M, Kd] * B[N, Kd]^T, all operands fp32 row-major
// with the contraction index contiguous ("K-major").  Used by the Gumbel-softmax path (rq_gumbel_tc.cu), whose softmax
// needs logits at fp32 accuracy: every fp32 value v is cut into three bf16 planes v = v1 + v2 + v3 (8 + 8 + 8 significand
// bits, both remainders exact in fp32) while its tile is staged into shared memory, and the product is the six bf16 MMAs
// a1.b1 + a1.b2 + a2.b1 + a2.b2 + a1.b3 + a3.b1 accumulated in fp32 in TMEM: the dropped terms are below 2^-24 |a||b|,
// i.e. the result is as good as an fp32 FMA chain (tests/test_gpu_gumbel.py checks it against float64).
//
// One CTA per 128 x NT output tile (and per K split): 256 threads stage a 64-wide K-block of A and B (global fp32 ->
// registers -> three 128-byte-swizzled bf16 tiles each), one thread issues the 24 tcgen05.mma of the block, the next
// block is staged when those have completed (single-buffered: the callers are bound by the [M, N] traffic to HBM or run
// a handful of K-blocks per CTA).  Epilogue: TMEM -> registers -> 32-byte global stores, or float atomics for split-K.
#include <cuda.h>
#include <string.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_tc_ptx.cuh"

namespace rq {

namespace sg {
constexpr int TM = 128;           // rows per tile (UMMA M)
constexpr int KB = 64;            // contraction elements per staged block (one 128-byte swizzled bf16 row)
constexpr int THREADS = 256;
__device__ __forceinline__ uint32_t sw128_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}
// three bf16 planes of four consecutive fp32 values: 8 bytes per plane
__device__ __forceinline__ void split3_store(const float4& v, uint8_t* p1, uint8_t* p2, uint8_t* p3, bool third = true) {
    const uint32_t h01 = ptx::pack_bf16x2(v.x, v.y), h23 = ptx::pack_bf16x2(v.z, v.w);
    const float r0 = v.x - __uint_as_float(h01 << 16), r1 = v.y - __uint_as_float(h01 & 0xFFFF0000u);
    const float r2 = v.z - __uint_as_float(h23 << 16), r3 = v.w - __uint_as_float(h23 & 0xFFFF0000u);
    const uint32_t m01 = ptx::pack_bf16x2(r0, r1), m23 = ptx::pack_bf16x2(r2, r3);
    if (!third) {
        *reinterpret_cast<uint2*>(p1) = make_uint2(h01, h23);
        *reinterpret_cast<uint2*>(p2) = make_uint2(m01, m23);
        return;
    }
    const float s0 = r0 - __uint_as_float(m01 << 16), s1 = r1 - __uint_as_float(m01 & 0xFFFF0000u);
    const float s2 = r2 - __uint_as_float(m23 << 16), s3 = r3 - __uint_as_float(m23 & 0xFFFF0000u);
    const uint32_t l01 = ptx::pack_bf16x2(s0, s1), l23 = ptx::pack_bf16x2(s2, s3);
    *reinterpret_cast<uint2*>(p1) = make_uint2(h01, h23);
    *reinterpret_cast<uint2*>(p2) = make_uint2(m01, m23);
    *reinterpret_cast<uint2*>(p3) = make_uint2(l01, l23);
}
// four consecutive fp32 of row `row` starting at column k of a [rows, Kd] matrix (zeros outside)
__device__ __forceinline__ float4 load4_guarded(const float* base, int64_t row, int64_t rows, int64_t ld, int64_t k, int64_t Kd) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < rows) {
        const float* src = base + row * ld + k;
        if (k + 3 < Kd) v = __ldg(reinterpret_cast<const float4*>(src));
        else {
            if (k < Kd) v.x = __ldg(src);
            if (k + 1 < Kd) v.y = __ldg(src + 1);
            if (k + 2 < Kd) v.z = __ldg(src + 2);
        }
    }
    return v;
}
// Stages ROWS x 64 fp32 (rows row0.., columns k0..) as bf16 planes: ALL of a thread's loads are issued before the first
// conversion, so a K-block costs one memory latency, not one per float4 (the first version: 8.7 us per block).
template <int ROWS>
__device__ __forceinline__ void tile_load(const float* base, int64_t row0, int64_t rows, int64_t ld, int64_t k0, int64_t Kd,
                                          float4 (&v)[ROWS * 16 / THREADS], int tid) {
#pragma unroll
    for (int j = 0; j < ROWS * 16 / THREADS; ++j) {
        const int i = tid + j * THREADS;
        v[j] = load4_guarded(base, row0 + (i >> 4), rows, ld, k0 + (i & 15) * 4, Kd);
    }
}
template <int ROWS>
__device__ __forceinline__ void tile_store(const float4 (&v)[ROWS * 16 / THREADS], uint8_t* planes, int plane_bytes, bool third,
                                           int tid) {
#pragma unroll
    for (int j = 0; j < ROWS * 16 / THREADS; ++j) {
        const int i = tid + j * THREADS;
        const int r = i >> 4, c4 = i & 15;
        const uint32_t o = sw128_off(r, c4 >> 1) + (c4 & 1) * 8;
        split3_store(v[j], planes + o, planes + plane_bytes + o, planes + 2 * plane_bytes + o, third);
    }
}
template <int ROWS>
__device__ __forceinline__ void stage_tile(const float* base, int64_t row0, int64_t rows, int64_t ld, int64_t k0, int64_t Kd,
                                           uint8_t* planes, int plane_bytes, bool third, int tid) {
    float4 v[ROWS * 16 / THREADS];
    tile_load<ROWS>(base, row0, rows, ld, k0, Kd, v, tid);
    tile_store<ROWS>(v, planes, plane_bytes, third, tid);
}
// ---- operands given transposed: T[Kd, R] row-major, R (= M or N) contiguous ("MN-major") --------------------------------
// The tile keeps that orientation in shared memory: 64-element (128-byte) swizzle atoms along R, 8 k-rows per 1024-byte
// group, the 16-byte piece index XORed with k & 7 (UMMA canonical MN-major SW128: LBO = 8192 bytes from one atom column to
// the next, SBO = 1024 bytes from one group of 8 k to the next).  No transposed copy of the matrix has to exist in HBM.
__device__ __forceinline__ uint32_t mn128_off(int k, int c) {
    return (uint32_t)((c >> 6) * 8192 + (k >> 3) * 1024 + (k & 7) * 128 + ((((c & 63) >> 3) ^ (k & 7)) << 4) + ((c & 7) << 1));
}
template <int COLS>
__device__ __forceinline__ void tile_load_t(const float* base, int64_t col0, int64_t cols, int64_t ld, int64_t k0, int64_t Kd,
                                            float4 (&v)[COLS * 16 / THREADS], int tid) {
#pragma unroll
    for (int j = 0; j < COLS * 16 / THREADS; ++j) {
        const int i = tid + j * THREADS;
        v[j] = load4_guarded(base, k0 + i / (COLS / 4), Kd, ld, col0 + (i % (COLS / 4)) * 4, cols);
    }
}
template <int COLS>
__device__ __forceinline__ void tile_store_t(const float4 (&v)[COLS * 16 / THREADS], uint8_t* planes, int plane_bytes, bool third,
                                             int tid) {
#pragma unroll
    for (int j = 0; j < COLS * 16 / THREADS; ++j) {
        const int i = tid + j * THREADS;
        const uint32_t o = mn128_off(i / (COLS / 4), (i % (COLS / 4)) * 4);
        split3_store(v[j], planes + o, planes + plane_bytes + o, planes + 2 * plane_bytes + o, third);
    }
}
// sum over the warp's 32 lanes of v[0..32): lane j returns the total of element j (31 shuffles; v is destroyed)
__device__ __forceinline__ float warp_colsum32(float (&v)[32]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int j = 0; j < half; ++j) {
            const float keep = up ? v[j + half] : v[j];
            const float send = up ? v[j] : v[j + half];
            v[j] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
    }
    return v[0];
}
}  // namespace sg

struct SplitGemmParams {
    const float* A;      // [M, Kd], row stride lda
    const float* B;      // [N, Kd], row stride ldb
    float* C;            // [M, N], row stride ldc
    int64_t M, N, Kd, lda, ldb, ldc;
    float alpha;
    int atomic;          // 1: C += alpha * A B^T with float atomics (split-K, or accumulation into a live buffer)
    int kb_per_split;    // K-blocks per blockIdx.z
    int terms;           // 6: three planes, fp32-grade (error ~2^-22 |a||b|); 3: two planes, a1.b1 + a1.b2 + a2.b1 (2^-16)
    // optional epilogue of the B-resident kernel (the softmax Jacobian of the Gumbel backward, fused into dW = de C^T):
    // C[m, n] = ep_scale * ep_w[m, n] * (acc[m, n] - ep_t[m]) and ep_csum[n] += sum_m C[m, n]; ep_w has C's shape / stride
    const float* ep_w;
    const float* ep_t;
    float ep_scale;
    float* ep_csum;
};

// TRANS: A is given as At[Kd, M] and B as Bt[Kd, N] (lda / ldb are their row strides), see mn128_off.
template <int NT, bool TRANS = false>
__global__ void __launch_bounds__(sg::THREADS, 1) rq_split_gemm_kernel(SplitGemmParams g) {
    using namespace sg;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int A_PLANE = TM * 128, B_PLANE = NT * 128;
    uint8_t* sA = smem;                           // [3][128 rows][64 bf16]
    uint8_t* sB = smem + 3 * A_PLANE;             // [3][NT rows][64 bf16]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 3 * A_PLANE + 3 * B_PLANE);
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(mbar + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t m0 = (int64_t)blockIdx.x * TM, n0 = (int64_t)blockIdx.y * NT;
    const int64_t n_kb = (g.Kd + KB - 1) / KB;
    const int64_t kb0 = (int64_t)blockIdx.z * g.kb_per_split;
    const int64_t kb1 = (kb0 + g.kb_per_split < n_kb) ? kb0 + g.kb_per_split : n_kb;
    if (kb0 >= kb1) return;

    if (tid == 0) {
        ptx::mbar_init(mbar, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, NT < 32 ? 32 : NT);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_acc = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
    // kind::f16 instruction descriptor, bf16 x bf16 -> fp32, K-major operands
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TM, NT) | (TRANS ? (1u << 15) | (1u << 16) : 0u);   // a_major, b_major = MN
    bool timed_out = false;

    // The loads of block kb + 1 are issued right after the MMAs of block kb, so their latency runs under those MMAs and
    // the wait for them (NT = 64: 12 float4 per thread in flight; the 256-wide tile has no registers to spare for it).
    constexpr bool PIPE = (NT == 64);
    float4 ra[TM * 16 / THREADS], rb[NT * 16 / THREADS];
    if (PIPE && TRANS) {
        tile_load_t<TM>(g.A, m0, g.M, g.lda, kb0 * KB, g.Kd, ra, tid);
        tile_load_t<NT>(g.B, n0, g.N, g.ldb, kb0 * KB, g.Kd, rb, tid);
    } else if (PIPE) {
        tile_load<TM>(g.A, m0, g.M, g.lda, kb0 * KB, g.Kd, ra, tid);
        tile_load<NT>(g.B, n0, g.N, g.ldb, kb0 * KB, g.Kd, rb, tid);
    }
    for (int64_t kb = kb0; kb < kb1; ++kb) {
        if (kb > kb0) {                                            // the tiles are still being read by the last block's MMAs
            if (!ptx::mbar_wait(mbar, (uint32_t)((kb - kb0 - 1) & 1))) timed_out = true;
            ptx::tc_fence_after_sync();
        }
        const int64_t k0 = kb * KB;
        // ---- stage: 16 float4 per row, a warp reads two rows' 256 contiguous bytes per instruction ----------------------
        if (PIPE && TRANS) {
            tile_store_t<TM>(ra, sA, A_PLANE, g.terms == 6, tid);
            tile_store_t<NT>(rb, sB, B_PLANE, g.terms == 6, tid);
        } else if (PIPE) {
            tile_store<TM>(ra, sA, A_PLANE, g.terms == 6, tid);
            tile_store<NT>(rb, sB, B_PLANE, g.terms == 6, tid);
        } else {
            stage_tile<TM>(g.A, m0, g.M, g.lda, k0, g.Kd, sA, A_PLANE, g.terms == 6, tid);
            stage_tile<NT>(g.B, n0, g.N, g.ldb, k0, g.Kd, sB, B_PLANE, g.terms == 6, tid);
        }
        ptx::fence_proxy_async_smem();
        ptx::tc_fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            ptx::tc_fence_after_sync();
            const uint32_t a0 = ptx::smem_u32(sA), b0 = ptx::smem_u32(sB);
            // (A plane, B plane) of the six kept terms, smallest first: the tensor core's fp32 accumulation truncates, so
            // the small corrections are summed among themselves before the leading product comes in
            const int pa[6] = {2, 0, 1, 1, 0, 0};
            const int pb[6] = {0, 2, 1, 0, 1, 0};
#pragma unroll
            for (int t = 0; t < 6; ++t) {
                if (g.terms != 6 && t < 3) continue;               // two planes: the last three products only
#pragma unroll
                for (int s = 0; s < 4; ++s) {                      // four K = 16 steps of 32 bytes
                    // a K = 16 step is 32 bytes along a K-major row, two 1024-byte groups of 8 k in the MN-major tile
                    const uint64_t ad = TRANS ? ptx::make_smem_desc(a0 + pa[t] * A_PLANE + s * 2048, 8192, 1024, ptx::kLayoutSw128)
                                              : ptx::make_smem_desc(a0 + pa[t] * A_PLANE + s * 32, 16, 1024, ptx::kLayoutSw128);
                    const uint64_t bd = TRANS ? ptx::make_smem_desc(b0 + pb[t] * B_PLANE + s * 2048, 8192, 1024, ptx::kLayoutSw128)
                                              : ptx::make_smem_desc(b0 + pb[t] * B_PLANE + s * 32, 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(tmem_acc, ad, bd, idesc, (kb > kb0 || t > (g.terms == 6 ? 0 : 3) || s > 0) ? 1u : 0u);
                }
            }
            ptx::umma_commit(mbar);
        }
        if (PIPE && TRANS && kb + 1 < kb1) {
            tile_load_t<TM>(g.A, m0, g.M, g.lda, k0 + KB, g.Kd, ra, tid);
            tile_load_t<NT>(g.B, n0, g.N, g.ldb, k0 + KB, g.Kd, rb, tid);
        } else if (PIPE && kb + 1 < kb1) {
            tile_load<TM>(g.A, m0, g.M, g.lda, k0 + KB, g.Kd, ra, tid);
            tile_load<NT>(g.B, n0, g.N, g.ldb, k0 + KB, g.Kd, rb, tid);
        }
    }
    if (!ptx::mbar_wait(mbar, (uint32_t)((kb1 - kb0 - 1) & 1))) timed_out = true;
    ptx::tc_fence_after_sync();

    // ---- epilogue: warp w reads TMEM lanes 32 (w & 3) .., columns of half (w >> 2) ----------------------------------------
    {
        const int q = warp & 3, h = warp >> 2;
        const int64_t row = m0 + q * 32 + lane;
        constexpr int HALF = NT / 2;
        const uint32_t trow = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * HALF);
#pragma unroll 1
        for (int c = 0; c < HALF; c += 32) {
            uint32_t v[32];
            ptx::tmem_ld_32x32b_x32(trow + c, v);
            ptx::tmem_ld_wait_for(v);
            if (row < g.M) {
                float* dst = g.C + row * g.ldc + n0 + h * HALF + c;
                const int64_t ncol = g.N - (n0 + h * HALF + c);        // columns left in this row
                if (!g.atomic && ncol >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 31) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        float o[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) o[u] = g.alpha * __uint_as_float(v[j + u]);
                        ptx::st_global_v8(dst + j, o);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (j < ncol) {
                            const float o = g.alpha * __uint_as_float(v[j]);
                            if (g.atomic) atomicAdd(dst + j, o);
                            else dst[j] = o;
                        }
                    }
                }
            }
        }
    }
    (void)timed_out;
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_acc, NT < 32 ? 32 : NT);
}

// The same product for Kd <= 64 (one K-block: the [B, D] x [K, D]^T logits of the Gumbel path) as a persistent kernel:
// a CTA converts its 256-row B tile ONCE and walks over M tiles with two A buffers and two TMEM accumulators, so the
// MMAs of tile i + 1 run under the epilogue (128 KB of stores) of tile i.  grid = (N tiles, CTAs per N tile).
// map_c (use_tma != 0): C as a tensor map with 32 x 32 boxes - the epilogue then goes through a swizzled 4 KB staging tile
// per warp and TMA stores (whole 128-byte lines) instead of thread-per-row 32-byte stores.
__global__ void __launch_bounds__(sg::THREADS, 1) rq_split_gemm_bres_kernel(const __grid_constant__ CUtensorMap map_c,
                                                                            int use_tma, SplitGemmParams g) {
    using namespace sg;
    constexpr int NT = 256;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int A_PLANE = TM * 128, B_PLANE = NT * 128;
    uint8_t* sB = smem;                                   // [3][256 rows][64 bf16]
    uint8_t* sA = smem + 3 * B_PLANE;                     // [2 buffers][3][128 rows][64 bf16]
    uint8_t* my_out = smem + 3 * B_PLANE + 6 * A_PLANE + (threadIdx.x >> 5) * 4096;       // [8 warps][32 rows][128 B]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 3 * B_PLANE + 6 * A_PLANE + 8 * 4096);     // [2]
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(mbar + 2);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n0 = (int64_t)blockIdx.x * NT;
    // 32 values of this thread's row -> staging tile (16-byte pieces XOR-swizzled by the row) -> one TMA store
    auto put_chunk = [&](const float (&o)[32], int64_t col0, int64_t row0) {
        if (lane == 0) ptx::bulk_wait_group_read0();
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(my_out + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            ptx::tma_store_2d(&map_c, (int)col0, (int)row0, my_out);
            ptx::bulk_commit_group();
        }
    };
    const int64_t m_tiles = (g.M + TM - 1) / TM;
    const int64_t n_mine = (m_tiles > (int64_t)blockIdx.y) ? (m_tiles - 1 - blockIdx.y) / gridDim.y + 1 : 0;
    if (tid == 0) {
        ptx::mbar_init(&mbar[0], 1);
        ptx::mbar_init(&mbar[1], 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc(tmem_base_s, 512);
    // B tile: converted once
    stage_tile<NT>(g.B, n0, g.N, g.ldb, 0, g.Kd, sB, B_PLANE, g.terms == 6, tid);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
    const uint32_t idesc = ptx::make_idesc_bf16_f32(TM, NT);
    bool timed_out = false;

    auto stage_a = [&](int64_t it) {
        const int64_t m0 = ((int64_t)blockIdx.y + it * gridDim.y) * TM;
        uint8_t* dstA = sA + (it & 1) * 3 * A_PLANE;
        stage_tile<TM>(g.A, m0, g.M, g.lda, 0, g.Kd, dstA, A_PLANE, g.terms == 6, tid);
        ptx::fence_proxy_async_smem();
        ptx::tc_fence_before_sync();                 // this thread's tcgen05.ld of the accumulator about to be reused
        __syncthreads();
        if (tid == 0) {
            ptx::tc_fence_after_sync();
            const uint32_t a0 = ptx::smem_u32(dstA), b0 = ptx::smem_u32(sB);
            const uint32_t acc = tmem_base + (uint32_t)((it & 1) * NT);
            const int pa[6] = {2, 0, 1, 1, 0, 0};
            const int pb[6] = {0, 2, 1, 0, 1, 0};
            bool first = true;
#pragma unroll
            for (int t = 0; t < 6; ++t) {
                if (g.terms != 6 && t < 3) continue;
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    const uint64_t ad = ptx::make_smem_desc(a0 + pa[t] * A_PLANE + s * 32, 16, 1024, ptx::kLayoutSw128);
                    const uint64_t bd = ptx::make_smem_desc(b0 + pb[t] * B_PLANE + s * 32, 16, 1024, ptx::kLayoutSw128);
                    ptx::umma_bf16(acc, ad, bd, idesc, first ? 0u : 1u);
                    first = false;
                }
            }
            ptx::umma_commit(&mbar[it & 1]);
        }
    };

    float ep_cs[NT / 2 / 32] = {0.f, 0.f, 0.f, 0.f};   // column sums of the fused epilogue (lane = column within a chunk)
    if (n_mine > 0) stage_a(0);
    for (int64_t it = 0; it < n_mine; ++it) {
        // the next tile's rows go in (other A buffer, other accumulator: both were released before the last barrier)
        if (it + 1 < n_mine) stage_a(it + 1);
        if (!ptx::mbar_wait(&mbar[it & 1], (uint32_t)((it >> 1) & 1))) timed_out = true;
        ptx::tc_fence_after_sync();
        const int64_t m0 = ((int64_t)blockIdx.y + it * gridDim.y) * TM;
        const int q = warp & 3, h = warp >> 2;
        const int64_t row = m0 + q * 32 + lane;
        constexpr int HALF = NT / 2;
        const uint32_t trow = tmem_base + (uint32_t)((it & 1) * NT) + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * HALF);
        if (g.ep_w) {
            const float tr = row < g.M ? __ldg(g.ep_t + row) : 0.f;
#pragma unroll
            for (int ci = 0; ci < HALF / 32; ++ci) {
                const int c = ci * 32;
                const int64_t col0 = n0 + h * HALF + c;
                const int64_t ncol = g.N - col0;
                const bool fast = row < g.M && ncol >= 32;
                float w[32];
                if (fast) {                                          // this row's 128 bytes of W, in flight under the tcgen05.ld
                    const float4* src = reinterpret_cast<const float4*>(g.ep_w + row * g.ldc + col0);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 t4 = __ldg(src + j);
                        w[4 * j] = t4.x; w[4 * j + 1] = t4.y; w[4 * j + 2] = t4.z; w[4 * j + 3] = t4.w;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) w[j] = (row < g.M && j < ncol) ? __ldg(g.ep_w + row * g.ldc + col0 + j) : 0.f;
                }
                uint32_t v[32];
                ptx::tmem_ld_32x32b_x32(trow + c, v);
                ptx::tmem_ld_wait_for(v);
                float o[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) o[j] = (row < g.M && j < ncol) ? g.ep_scale * (w[j] * (__uint_as_float(v[j]) - tr)) : 0.f;
                float* dst = g.C + row * g.ldc + col0;
                if (use_tma) {
                    if (ncol > 0) put_chunk(o, col0, m0 + q * 32);
                } else if (fast && ((reinterpret_cast<uintptr_t>(dst) & 31) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) ptx::st_global_v8(dst + j, &o[j]);
                } else if (row < g.M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncol) dst[j] = o[j];
                }
                ep_cs[ci] += warp_colsum32(o);
            }
            continue;
        }
#pragma unroll 1
        for (int c = 0; c < HALF; c += 32) {
            uint32_t v[32];
            ptx::tmem_ld_32x32b_x32(trow + c, v);
            ptx::tmem_ld_wait_for(v);
            if (use_tma) {
                if (n0 + h * HALF + c < g.N) {
                    float o[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) o[j] = g.alpha * __uint_as_float(v[j]);
                    put_chunk(o, n0 + h * HALF + c, m0 + q * 32);
                }
            } else if (row < g.M) {
                float* dst = g.C + row * g.ldc + n0 + h * HALF + c;
                const int64_t ncol = g.N - (n0 + h * HALF + c);
                if (ncol >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 31) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        float o[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) o[u] = g.alpha * __uint_as_float(v[j + u]);
                        ptx::st_global_v8(dst + j, o);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncol) dst[j] = g.alpha * __uint_as_float(v[j]);
                }
            }
        }
    }
    if (use_tma && lane == 0) ptx::bulk_wait_group0();
    if (g.ep_w && g.ep_csum) {                                       // this CTA's column sums: one atomic per warp and column
        const int h = warp >> 2;
#pragma unroll
        for (int ci = 0; ci < NT / 2 / 32; ++ci) {
            const int64_t col = n0 + h * (NT / 2) + ci * 32 + lane;
            if (col < g.N) atomicAdd(g.ep_csum + col, ep_cs[ci]);
        }
    }
    (void)timed_out;
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

// C[M, N] (+)= alpha * A[M, Kd] * B[N, Kd]^T.  k_splits > 1 (or atomic) accumulates with float atomics into C, which the
// caller has initialised.  Rows of A and B must be 16-byte aligned (lda, ldb multiples of 4 floats).
int split_gemm(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd, int64_t lda, int64_t ldb,
               int64_t ldc, float alpha, int atomic, int k_splits, cudaStream_t st, int terms) {
    if (M <= 0 || N <= 0 || Kd <= 0) return 0;
    if ((lda & 3) || (ldb & 3) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15)) {
        set_error("split_gemm: operand rows must be 16-byte aligned");
        return RQ_ERR_INVALID_ARGUMENT;
    }
    SplitGemmParams g;
    g.A = A; g.B = B; g.C = C; g.M = M; g.N = N; g.Kd = Kd; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.alpha = alpha;
    g.ep_w = nullptr; g.ep_t = nullptr; g.ep_scale = 0.f; g.ep_csum = nullptr;
    g.terms = (terms == 3) ? 3 : 6;
    const int64_t n_kb = (Kd + sg::KB - 1) / sg::KB;
    if (k_splits < 1) k_splits = 1;
    if (k_splits > n_kb) k_splits = (int)n_kb;
    g.kb_per_split = (int)((n_kb + k_splits - 1) / k_splits);
    k_splits = (int)((n_kb + g.kb_per_split - 1) / g.kb_per_split);
    g.atomic = (atomic || k_splits > 1) ? 1 : 0;
    const unsigned gm = (unsigned)((M + sg::TM - 1) / sg::TM);
    if (N > 64 && Kd <= sg::KB && !g.atomic) {
        // one K-block: persistent CTAs with the B tile resident
        const unsigned gn = (unsigned)((N + 255) / 256);
        unsigned per = (unsigned)sm_count() / gn;
        if (per < 1) per = 1;
        if (per > gm) per = gm;
        const size_t smem = 3 * 256 * 128 + 6 * sg::TM * 128 + 8 * 4096 + 64 + 1024;
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_split_gemm_bres_kernel, (int)smem));
        CUtensorMap mc;
        const int use_tma = (!(ldc & 3) && !((uintptr_t)C & 15) && make_tensor_map_f32(&mc, C, M, N, ldc, 32)) ? 1 : 0;
        if (!use_tma) memset(&mc, 0, sizeof(mc));
        rq_split_gemm_bres_kernel<<<dim3(gn, per), sg::THREADS, smem, st>>>(mc, use_tma, g);
    } else if (N > 64) {
        constexpr int NT = 256;
        const size_t smem = 3 * sg::TM * 128 + 3 * NT * 128 + 64 + 1024;
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_split_gemm_kernel<NT>, (int)smem));
        dim3 grid(gm, (unsigned)((N + NT - 1) / NT), (unsigned)k_splits);
        rq_split_gemm_kernel<NT><<<grid, sg::THREADS, smem, st>>>(g);
    } else {
        constexpr int NT = 64;
        const size_t smem = 3 * sg::TM * 128 + 3 * NT * 128 + 64 + 1024;
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_split_gemm_kernel<NT>, (int)smem));
        // 74 KB per CTA: three CTAs per SM hide each other's staging latency, if the carve-out leaves them the room
        static bool carve_set = false;
        if (!carve_set) {
            cudaFuncSetAttribute((const void*)rq_split_gemm_kernel<NT>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared);
            carve_set = true;
        }
        dim3 grid(gm, 1, (unsigned)k_splits);
        rq_split_gemm_kernel<NT><<<grid, sg::THREADS, smem, st>>>(g);
    }
    RQ_LAUNCHED();
    return 0;
}

// C[m, n] = scale * Wm[m, n] * ((A B^T)[m, n] - t[m]) and csum[n] += sum_m C[m, n]: the dd = -W (dW - t) / T step of the
// Gumbel backward inside the epilogue of dW = de C^T (Kd <= 64, N > 64: the B-resident kernel).  Wm has C's stride.
int split_gemm_softmax_jacobian(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd, int64_t lda,
                                int64_t ldb, int64_t ldc, const float* Wm, const float* t, float scale, float* csum,
                                cudaStream_t st) {
    if (M <= 0 || N <= 0 || Kd <= 0) return 0;
    if (N <= 64 || Kd > sg::KB || (lda & 3) || (ldb & 3) || (ldc & 3) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) ||
        ((uintptr_t)Wm & 15)) {
        set_error("split_gemm_softmax_jacobian: needs N > 64, Kd <= 64 and 16-byte aligned rows");
        return RQ_ERR_INVALID_ARGUMENT;
    }
    SplitGemmParams g;
    g.A = A; g.B = B; g.C = C; g.M = M; g.N = N; g.Kd = Kd; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.alpha = 1.f; g.terms = 6; g.atomic = 0; g.kb_per_split = 1;
    g.ep_w = Wm; g.ep_t = t; g.ep_scale = scale; g.ep_csum = csum;
    const unsigned gm = (unsigned)((M + sg::TM - 1) / sg::TM), gn = (unsigned)((N + 255) / 256);
    unsigned per = (unsigned)sm_count() / gn;
    if (per < 1) per = 1;
    if (per > gm) per = gm;
    const size_t smem = 3 * 256 * 128 + 6 * sg::TM * 128 + 8 * 4096 + 64 + 1024;
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_split_gemm_bres_kernel, (int)smem));
    CUtensorMap mc;
    const int use_tma = (!((uintptr_t)C & 15) && make_tensor_map_f32(&mc, C, M, N, ldc, 32)) ? 1 : 0;
    if (!use_tma) memset(&mc, 0, sizeof(mc));
    rq_split_gemm_bres_kernel<<<dim3(gn, per), sg::THREADS, smem, st>>>(mc, use_tma, g);
    RQ_LAUNCHED();
    return 0;
}

// C[M, N] (+)= alpha * At[Kd, M]^T * Bt[Kd, N]: both operands given with the contraction index as their ROW index (the
// shape of the Gumbel backward's codebook gradients, W^T de and dd^T x over the batch).  N <= 64; ldat, ldbt multiples of 4.
int split_gemm_tn(const float* At, const float* Bt, float* C, int64_t M, int64_t N, int64_t Kd, int64_t ldat, int64_t ldbt,
                  int64_t ldc, float alpha, int atomic, int k_splits, cudaStream_t st, int terms) {
    if (M <= 0 || N <= 0 || Kd <= 0) return 0;
    if (N > 64 || (ldat & 3) || (ldbt & 3) || ((uintptr_t)At & 15) || ((uintptr_t)Bt & 15)) {
        set_error("split_gemm_tn: N <= 64 and 16-byte aligned operand rows required");
        return RQ_ERR_INVALID_ARGUMENT;
    }
    SplitGemmParams g;
    g.A = At; g.B = Bt; g.C = C; g.M = M; g.N = N; g.Kd = Kd; g.lda = ldat; g.ldb = ldbt; g.ldc = ldc;
    g.alpha = alpha;
    g.ep_w = nullptr; g.ep_t = nullptr; g.ep_scale = 0.f; g.ep_csum = nullptr;
    g.terms = (terms == 3) ? 3 : 6;
    const int64_t n_kb = (Kd + sg::KB - 1) / sg::KB;
    if (k_splits < 1) k_splits = 1;
    if (k_splits > n_kb) k_splits = (int)n_kb;
    g.kb_per_split = (int)((n_kb + k_splits - 1) / k_splits);
    k_splits = (int)((n_kb + g.kb_per_split - 1) / g.kb_per_split);
    g.atomic = (atomic || k_splits > 1) ? 1 : 0;
    constexpr int NT = 64;
    const size_t smem = 3 * sg::TM * 128 + 3 * NT * 128 + 64 + 1024;
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_split_gemm_kernel<NT, true>, (int)smem));
    static bool carve_set = false;
    if (!carve_set) {
        cudaFuncSetAttribute((const void*)rq_split_gemm_kernel<NT, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
        carve_set = true;
    }
    dim3 grid((unsigned)((M + sg::TM - 1) / sg::TM), 1, (unsigned)k_splits);
    rq_split_gemm_kernel<NT, true><<<grid, sg::THREADS, smem, st>>>(g);
    RQ_LAUNCHED();
    return 0;
}

}  // namespace rq

// the same with both operands transposed in memory (k_splits < 0: two-plane form)
extern "C" int rq_debug_split_gemm_tn(const float* At, const float* Bt, float* C, int64_t M, int64_t N, int64_t Kd,
                                      int64_t ldat, int64_t ldbt, int64_t ldc, float alpha, int atomic, int k_splits,
                                      rq_stream_t stream) {
    return rq::split_gemm_tn(At, Bt, C, M, N, Kd, ldat, ldbt, ldc, alpha, atomic, k_splits < 0 ? -k_splits : k_splits,
                             (cudaStream_t)stream, k_splits < 0 ? 3 : 6);
}

// Test hook: the split-precision GEMM on its own (tests/test_gpu_gumbel.py compares it with float64).
extern "C" int rq_debug_split_gemm(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t Kd,
                                   int64_t lda, int64_t ldb, int64_t ldc, float alpha, int atomic, int k_splits,
                                   rq_stream_t stream) {
    // k_splits < 0 selects the two-plane (three-product) form with |k_splits| splits
    return rq::split_gemm(A, B, C, M, N, Kd, lda, ldb, ldc, alpha, atomic, k_splits < 0 ? -k_splits : k_splits,
                          (cudaStream_t)stream, k_splits < 0 ? 3 : 6);
}
