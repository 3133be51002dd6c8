// One layer of the bias-free SiLU MLP that feeds the quantiser (reference: modules/encoder.py:17-32, called from
// modules/rqvae.py:117-118 `encode`): Y[M, N] = act(X[M, Kd] * W[N, Kd]^T), X / W / Y fp32 row-major, with the
// activation (SiLU, or the final L2 normalisation, modules/normalize.py:9-13) applied while the tile is still in
// registers.  The product runs on the 5th-gen tensor cores as kind::tf32 — the precision the reference itself selects
// (`torch.set_float32_matmul_precision('high')`, modules/rqvae.py:19) — so the fp32 operands go from HBM to shared
// memory by TMA and into the MMA untouched: no conversion pass, no staging through registers.
//
// Persistent CTAs of ten warps:
//   warp 0 (one lane)  TMA producer: 256 x 32 fp32 box of X and NT x 32 box of W per K-block into a ring of stages
//                      (128-byte swizzle = the UMMA K-major SW128 layout; rows/columns outside the matrix arrive as 0);
//   warp 1 (one lane)  issues 2 x 4 tcgen05.mma (K = 8 each; the two 128-row halves of X against the same W stage) per
//                      K-block into a pair of TMEM accumulators, commits the stage back to the producer and, after the
//                      last K-block, the pair to the epilogue.  A 256-row tile doubles the flops per byte staged, which
//                      is what the ring depth needs: 128-row tiles left the tensor pipe 32 % busy waiting for data;
//   warps 2-9          epilogue, one warp per TMEM lane quadrant and half: tcgen05.ld 32 columns at a time, activation,
//                      a swizzled 4 KB staging tile per warp and a TMA store of it, then hand the pair back.  Tiles up to 128 columns wide keep
//                      two accumulator pairs in TMEM, so the next tile's MMAs run under this tile's epilogue.
// Tiles are numbered with the N index fastest, so the CTAs that share an X tile run at the same time and the second
// read of it is an L2 hit.
//
// Algorithmic bytes per row of one layer: 4 (Kd + N); the whole 768-512-256-128-d encoder moves 4 * (1280 + 768 + 384 +
// 128 + d) bytes per item against 2 * (768*512 + 512*256 + 256*128 + 128*d) flops.
#include <cuda.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_tc_ptx.cuh"

namespace rq {

namespace lt {
constexpr int TM = 256;            // rows per tile: two UMMA M = 128 halves that share every W stage
constexpr int KB = 32;             // fp32 contraction elements per K-block: one 128-byte swizzled row
constexpr int THREADS = 320;          // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quadrant)
constexpr int MAX_STAGES = 8;
constexpr int A_STAGE = TM * KB * 4;                                   // 32 KB: one TMA box, halves 16 KB apart
constexpr int OUT_STAGE = 8 * 32 * 128;                                // the epilogue warps' staging tiles
__host__ __device__ constexpr int b_stage_bytes(int NT) { return (NT * KB * 4 + 1023) & ~1023; }

struct Params {
    float* C;
    int64_t M, ldc;
    int N, Kd, NT, n_tiles_n, stages, act, round_out, nbuf, halves;
    int64_t tiles;
    int* timeout_flag;
};
struct Bars {
    uint64_t full[MAX_STAGES], empty[MAX_STAGES], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
    volatile int abort;
};

// v * sigmoid(v) as two MUFU ops (ex2, rcp: 2^-22-grade, far below the TF32 product error) and no branches: the first
// version's correctly rounded reciprocal cost 19 instructions per element and made the epilogue warps the bottleneck.
__device__ __forceinline__ float silu(float v) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return v * r;
}
// nearest value with a 10-bit significand: what the next layer's kind::tf32 MMA would otherwise obtain by truncation
__device__ __forceinline__ float round_tf32(float v) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
    return __uint_as_float(u);
}

__global__ void __launch_bounds__(THREADS, 1)
rq_linear_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                      const __grid_constant__ CUtensorMap map_c, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int b_stage = b_stage_bytes(p.NT);
    const int a_stage = p.halves * (A_STAGE / 2), tile_rows = p.halves * 128;
    const int stage_bytes = a_stage + b_stage;
    uint8_t* out_stage = smem + (size_t)p.stages * stage_bytes;                    // 8 epilogue warps x 4 KB
    Bars* bars = reinterpret_cast<Bars*>(out_stage + OUT_STAGE);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nkb = (p.Kd + KB - 1) / KB;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc_full[b], 1);
            ptx::mbar_init(&bars->acc_empty[b], 4 * p.halves);   // one arrival per working epilogue warp
        }
        bars->abort = 0;
        ptx::fence_mbar_init();
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_w);
        ptx::prefetch_tensormap(&map_c);
    }
    if (warp == 1) ptx::tmem_alloc(&bars->tmem_base, 512);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = bars->tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int64_t tile = blockIdx.x; tile < p.tiles && !bars->abort; tile += gridDim.x) {
                const int m0 = (int)(tile / p.n_tiles_n) * tile_rows, n0 = (int)(tile % p.n_tiles_n) * p.NT;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    if (!ptx::mbar_wait(&bars->empty[s], ph ^ 1)) { bars->abort = 1; break; }
                    uint8_t* sa = smem + (size_t)s * stage_bytes;
                    ptx::mbar_arrive_expect_tx(&bars->full[s], (uint32_t)(a_stage + p.NT * KB * 4));
                    ptx::tma_load_2d(sa, &map_a, kb * KB, m0, &bars->full[s]);
                    ptx::tma_load_2d(sa + a_stage, &map_w, kb * KB, n0, &bars->full[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = ptx::make_idesc_tf32_f32(128, p.NT);
            uint32_t it = 0, tcount = 0;
            for (int64_t tile = blockIdx.x; tile < p.tiles && !bars->abort; tile += gridDim.x, ++tcount) {
                const uint32_t buf = tcount % p.nbuf, aph = (tcount / p.nbuf) & 1;
                if (!ptx::mbar_wait(&bars->acc_empty[buf], aph ^ 1)) { bars->abort = 1; break; }
                const uint32_t acc = tmem + buf * p.halves * p.NT;
                ptx::tc_fence_after_sync();
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    if (!ptx::mbar_wait(&bars->full[s], ph)) { bars->abort = 1; break; }
                    ptx::tc_fence_after_sync();
                    const uint32_t a0 = ptx::smem_u32(smem + (size_t)s * stage_bytes), b0 = a0 + a_stage;
#pragma unroll
                    for (int k4 = 0; k4 < KB / 8; ++k4) {
                        const uint64_t bd = ptx::make_smem_desc(b0 + k4 * 32, 16, 1024, ptx::kLayoutSw128);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (h >= p.halves) break;
                            const uint64_t ad = ptx::make_smem_desc(a0 + h * (A_STAGE / 2) + k4 * 32, 16, 1024, ptx::kLayoutSw128);
                            ptx::umma_tf32(acc + h * p.NT, ad, bd, idesc, (kb | k4) != 0);
                        }
                    }
                    ptx::umma_commit(&bars->empty[s]);
                }
                ptx::umma_commit(&bars->acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;                                  // the TMEM lane quadrant this warp may read
        const int half = (warp - 2) >> 2;                        // which 128-row half (accumulator) of the tile
        uint8_t* my_out = out_stage + (warp - 2) * 4096;
        uint32_t tcount = 0;
        for (int64_t tile = blockIdx.x; tile < p.tiles && !bars->abort && half < p.halves; tile += gridDim.x, ++tcount) {
            const uint32_t buf = tcount % p.nbuf, aph = (tcount / p.nbuf) & 1;
            const int n0 = (int)(tile % p.n_tiles_n) * p.NT;
            if (!ptx::mbar_wait(&bars->acc_full[buf], aph)) { bars->abort = 1; break; }
            ptx::tc_fence_after_sync();
            const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + buf * p.halves * p.NT + half * p.NT;
            const int ncols = min(p.NT, p.N - n0);
            const int row0 = (int)(tile / p.n_tiles_n) * tile_rows + half * 128 + q * 32;   // this warp's first row
            // one 32-row x 32-column chunk: registers -> this warp's 4 KB staging tile (128-byte rows, 16-byte pieces
            // XOR-swizzled by the row so that the 32 lanes' v4 stores are conflict-free) -> one TMA store, which writes
            // whole 128-byte lines and clips at the matrix edge.  (Thread-per-row st.global hit 32 lines per instruction
            // and left the epilogue warps waiting on the store queue.)
            auto put_chunk = [&](const float (&v)[32], int c0) {
                if (lane == 0) ptx::bulk_wait_group_read0();                 // the previous chunk has left the staging tile
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(my_out + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    ptx::tma_store_2d(&map_c, n0 + c0, row0, my_out);
                    ptx::bulk_commit_group();
                }
            };
            if (p.act == RQ_ACT_L2NORM) {                        // the whole row is in this tile (N <= 64, host-checked)
                uint32_t r0[32], r1[32];
                ptx::tmem_ld_32x32b_x32(tbase, r0);
                if (p.NT > 32) ptx::tmem_ld_32x32b_x32(tbase + 32, r1);
                else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r1[j] = 0u;
                }
                ptx::tmem_ld_wait_for2(r0, r1);
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->acc_empty[buf]);
                float ss = 0.f;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < ncols) ss = fmaf(__uint_as_float(r0[j]), __uint_as_float(r0[j]), ss);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (32 + j < ncols) ss = fmaf(__uint_as_float(r1[j]), __uint_as_float(r1[j]), ss);
                const float inv = 1.f / fmaxf(sqrtf(ss), 1e-12f);         // F.normalize(p=2, eps=1e-12)
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r0[j]) * inv;
                put_chunk(v, 0);
                if (ncols > 32) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r1[j]) * inv;
                    put_chunk(v, 32);
                }
            } else {
                for (int c0 = 0; c0 < p.NT; c0 += 32) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(tbase + c0, r);
                    ptx::tmem_ld_wait_for(r);
                    if (c0 + 32 >= p.NT) {                       // this warp's last read of the accumulator: give it back
                        ptx::tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(&bars->acc_empty[buf]);
                    }
                    if (c0 >= ncols) continue;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        float t = __uint_as_float(r[j]);
                        if (p.act == RQ_ACT_SILU) t = silu(t);
                        if (p.round_out) t = round_tf32(t);
                        v[j] = t;
                    }
                    put_chunk(v, c0);
                }
            }
        }
        if (lane == 0) ptx::bulk_wait_group0();                  // every store of this warp has landed
    }
    __syncwarp();
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after_sync();
        ptx::tmem_dealloc(tmem, 512);
    }
    if (tid == 0 && bars->abort && p.timeout_flag) *p.timeout_flag = 1;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}
// [rows, cols] fp32 row-major (leading dimension ld floats), boxes of box_rows x 32 floats, 128-byte swizzle
static bool make_map(CUtensorMap* m, const float* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    if (ld % 4) return false;
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)KB, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static int* g_timeout_flag = nullptr;
}  // namespace lt

bool linear_tf32_supported(int64_t M, int N, int Kd, int act_flags) {
    const int act = act_flags & ~RQ_ACT_TF32_OUT;
    if (M < 1 || N < 8 || Kd < 4 || (N & 3) || (Kd & 3)) return false;          // 16-byte rows for TMA and the stores
    if (act == RQ_ACT_L2NORM && (N > 64 || (act_flags & RQ_ACT_TF32_OUT))) return false;
    if (act < RQ_ACT_NONE || act > RQ_ACT_L2NORM) return false;
    return true;
}

// Y = act(X W^T); engine status code.
int linear_act_tf32(const float* X, const float* W, float* Y, int64_t M, int N, int Kd, int act_flags, cudaStream_t st) {
    using namespace lt;
    const int act = act_flags;
    RQ_CHECK_ARG(linear_tf32_supported(M, N, Kd, act), RQ_ERR_UNSUPPORTED_SHAPE,
                 "encoder layer: rows=%lld in=%d out=%d act=%d is outside the tf32 kernel (in, out multiples of 4; L2 norm "
                 "needs out <= 64)", (long long)M, Kd, N, act);
    RQ_CHECK_ARG(!((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(Y)) & 15),
                 RQ_ERR_INVALID_ARGUMENT, "encoder layer: x, weight and y must be 16-byte aligned");
    Params p{};
    // 256-wide tiles when the layer is MMA-heavy (one accumulator pair fills TMEM); otherwise <= 128 so that two pairs fit
    // and the epilogue of one tile runs under the MMAs of the next
    p.NT = N >= 512 ? 256 : (N >= 128 ? 128 : ((N + 15) & ~15));
#ifdef RQ_EXPERIMENTS
    if (const char* e = getenv("RQ_LIN_NT")) p.NT = std::min(atoi(e), (N + 15) & ~15);
#endif
    p.halves = 2;                            // 128-row tiles measured 15-60 % slower at every encoder shape
    p.nbuf = p.halves * p.NT <= 256 ? 2 : 1;
    p.n_tiles_n = (N + p.NT - 1) / p.NT;
    p.C = Y; p.M = M; p.ldc = N; p.N = N; p.Kd = Kd; p.act = act & ~RQ_ACT_TF32_OUT; p.round_out = (act & RQ_ACT_TF32_OUT) ? 1 : 0;
    const int tile_rows = 128 * p.halves;
    p.tiles = ((M + tile_rows - 1) / tile_rows) * p.n_tiles_n;
    const int stage_bytes = p.halves * (A_STAGE / 2) + b_stage_bytes(p.NT);
    p.stages = std::min<int>(MAX_STAGES, (224 * 1024 - OUT_STAGE) / stage_bytes);
    const size_t smem = (size_t)p.stages * stage_bytes + OUT_STAGE + sizeof(Bars) + 1024;
    CUtensorMap ma, mw, mc;
    RQ_CHECK_ARG(make_map(&ma, X, M, Kd, Kd, tile_rows) && make_map(&mw, W, N, Kd, Kd, p.NT) && make_map(&mc, Y, M, N, N, 32),
                 (int)cudaErrorInvalidValue,
                 "encoder layer: cuTensorMapEncodeTiled failed (rows=%lld in=%d out=%d)", (long long)M, Kd, N);
    RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_linear_tf32_kernel, (int)smem));
    if (!g_timeout_flag) {
        RQ_CUDA(cudaMalloc(&g_timeout_flag, sizeof(int)));
        RQ_CUDA(cudaMemset(g_timeout_flag, 0, sizeof(int)));
    }
    p.timeout_flag = g_timeout_flag;
    int grid = (int)std::min<int64_t>(p.tiles, sm_count());
#ifdef RQ_EXPERIMENTS
    if (const char* e = getenv("RQ_LIN_GRID")) grid = std::min(grid, atoi(e));
    if (const char* e = getenv("RQ_LIN_STAGES")) p.stages = std::min(p.stages, atoi(e));
#endif
    rq_linear_tf32_kernel<<<grid, THREADS, smem, st>>>(ma, mw, mc, p);
    RQ_LAUNCHED();
    return 0;
}
int linear_tf32_timeout_flag() {
    int v = 0;
    if (lt::g_timeout_flag) cudaMemcpy(&v, lt::g_timeout_flag, sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}

}  // namespace rq

// C-ABI (include/rq_engine.h)
extern "C" int rq_encoder_layer_supported(int64_t rows, int in_features, int out_features, int act) {
    return rq::linear_tf32_supported(rows, out_features, in_features, act) ? 1 : 0;
}
extern "C" int rq_encoder_layer(const float* x, const float* weight, float* y, int64_t rows, int in_features,
                                int out_features, int act, rq_stream_t stream) {
    RQ_CHECK_ARG(x && weight && y, RQ_ERR_NULL_POINTER, "rq_encoder_layer: null buffer");
    if (rows == 0) return 0;
    return rq::linear_act_tf32(x, weight, y, rows, out_features, in_features, act, (cudaStream_t)stream);
}
