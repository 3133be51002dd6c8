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

#include <type_traits>

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
    int st_mode;                       // 0: TMA store of the staging tile, 1: coalesced st.global from it
    int noload;                        // RQ_EXPERIMENTS what-if: stage contents are loaded once, then only the barriers cycle
    int64_t tiles;
    unsigned long long* dbg;           // RQ_EXPERIMENTS: cycle counters of block 0
};
struct Bars {
    uint64_t full[MAX_STAGES], empty[MAX_STAGES], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
    volatile int abort;
};

// raised when a barrier wait gave up (rq_debug_tc_timeout_flag); a file-scope word, so a launch allocates nothing and
// can be captured into a CUDA graph from its first call on
__device__ int g_linear_timeout = 0;

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

// PAIR: a cluster of two CTAs works on a 512-row x 256-column tile with cta_group::2 MMAs (M = 256: each CTA supplies
// 128 rows of X per MMA and HALF of the W stage, so a K-block costs each SM 32 + 16 KB of shared-memory fill instead of
// 32 + 32 KB - the single-CTA kernel is bound by exactly that fill rate on the 768 -> 512 layer).  Both CTAs run their own
// TMA producer (completion bytes counted on the leader's barrier) and their own epilogue on their own TMEM; only the
// leader issues MMAs, and its commits arrive on the barriers of both CTAs.
template <bool PAIR>
__global__ void __launch_bounds__(THREADS, 1)
rq_linear_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                      const __grid_constant__ CUtensorMap map_c, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;
    const int b_rows = PAIR ? p.NT / 2 : p.NT;                                     // W rows this CTA stages per K-block
    const int b_stage = b_stage_bytes(b_rows);
    const int a_stage = p.halves * (A_STAGE / 2), tile_rows = p.halves * 128;      // X rows this CTA owns per tile
    const int64_t first = PAIR ? blockIdx.x >> 1 : blockIdx.x, step = PAIR ? gridDim.x >> 1 : gridDim.x;
    const int stage_bytes = a_stage + b_stage;
    uint8_t* out_stage = smem + (size_t)p.stages * stage_bytes;                    // 8 epilogue warps x 4 KB
    Bars* bars = reinterpret_cast<Bars*>(out_stage + OUT_STAGE);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nkb = (p.Kd + KB - 1) / KB;
#ifdef RQ_EXPERIMENTS
    const long long t_start = clock64();
#endif

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc_full[b], 1);
            ptx::mbar_init(&bars->acc_empty[b], 8 * (PAIR ? 2 : 1));   // one arrival per epilogue warp
        }
        bars->abort = 0;
        ptx::fence_mbar_init();
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_w);
        ptx::prefetch_tensormap(&map_c);
    }
    if (warp == 1) {
        if (PAIR) ptx::tmem_alloc_pair(&bars->tmem_base, 512);
        else ptx::tmem_alloc(&bars->tmem_base, 512);
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (PAIR) ptx::cluster_sync();                       // the peer's barriers exist before anything is sent to them
    ptx::tc_fence_after_sync();
    const uint32_t tmem = bars->tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t full0 = PAIR ? ptx::mapa_shared(&bars->full[0], 0) : 0u;       // the leader's full[] barriers
            for (int64_t tile = first; tile < p.tiles && !bars->abort; tile += step) {
                const int m0 = (int)(tile / p.n_tiles_n) * tile_rows * (PAIR ? 2 : 1) + (int)rank * tile_rows;
                const int n0 = (int)(tile % p.n_tiles_n) * p.NT + (int)rank * (PAIR ? b_rows : 0);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
                    if (!ptx::mbar_wait(&bars->empty[s], ph ^ 1)) { bars->abort = 1; break; }
                    uint8_t* sa = smem + (size_t)s * stage_bytes;
                    const uint32_t bytes = (uint32_t)(a_stage + b_rows * KB * 4);
#ifdef RQ_EXPERIMENTS
                    if (p.noload == 1 && it >= (uint32_t)p.stages) {
                        if (rank == 0) ptx::mbar_arrive(&bars->full[s]);
                        continue;
                    }
                    if (!PAIR && p.noload >= 2 && it >= (uint32_t)p.stages) {      // 2: W only is loaded, 3: X only
                        if (p.noload == 2) {
                            ptx::mbar_arrive_expect_tx(&bars->full[s], (uint32_t)(b_rows * KB * 4));
                            ptx::tma_load_2d(sa + a_stage, &map_w, kb * KB, n0, &bars->full[s]);
                        } else {
                            ptx::mbar_arrive_expect_tx(&bars->full[s], (uint32_t)a_stage);
                            ptx::tma_load_2d(sa, &map_a, kb * KB, m0, &bars->full[s]);
                        }
                        continue;
                    }
#endif
                    if (PAIR) {
                        if (rank == 0) ptx::mbar_arrive_expect_tx(&bars->full[s], 2 * bytes);
                        ptx::tma_load_2d_pair(sa, &map_a, kb * KB, m0, full0 + s * 8);
                        ptx::tma_load_2d_pair(sa + a_stage, &map_w, kb * KB, n0, full0 + s * 8);
                    } else {
                        ptx::mbar_arrive_expect_tx(&bars->full[s], bytes);
                        ptx::tma_load_2d(sa, &map_a, kb * KB, m0, &bars->full[s]);
                        ptx::tma_load_2d(sa + a_stage, &map_w, kb * KB, n0, &bars->full[s]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = ptx::make_idesc_tf32_f32(PAIR ? 256 : 128, p.NT);
            uint32_t it = 0, tcount = 0;
            for (int64_t tile = first; tile < p.tiles && !bars->abort; tile += step, ++tcount) {
                const uint32_t buf = tcount % p.nbuf, aph = (tcount / p.nbuf) & 1;
#ifdef RQ_EXPERIMENTS
                const long long w0 = clock64();
#endif
                if (!ptx::mbar_wait(&bars->acc_empty[buf], aph ^ 1)) { bars->abort = 1; break; }
#ifdef RQ_EXPERIMENTS
                if (p.dbg && blockIdx.x == 0) p.dbg[2] += clock64() - w0;
#endif
                const uint32_t acc = tmem + buf * p.halves * p.NT;
                ptx::tc_fence_after_sync();
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % p.stages;
                    const uint32_t ph = (it / p.stages) & 1;
#ifdef RQ_EXPERIMENTS
                    const long long w1 = clock64();
#endif
                    if (!ptx::mbar_wait(&bars->full[s], ph)) { bars->abort = 1; break; }
#ifdef RQ_EXPERIMENTS
                    if (p.dbg && blockIdx.x == 0) p.dbg[3] += clock64() - w1;
#endif
                    ptx::tc_fence_after_sync();
                    const uint32_t a0 = ptx::smem_u32(smem + (size_t)s * stage_bytes), b0 = a0 + a_stage;
#pragma unroll
                    for (int k4 = 0; k4 < KB / 8; ++k4) {
                        const uint64_t bd = ptx::make_smem_desc(b0 + k4 * 32, 16, 1024, ptx::kLayoutSw128);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (h >= p.halves) break;
                            const uint64_t ad = ptx::make_smem_desc(a0 + h * (A_STAGE / 2) + k4 * 32, 16, 1024, ptx::kLayoutSw128);
                            if (PAIR) ptx::umma_tf32_pair(acc + h * p.NT, ad, bd, idesc, (kb | k4) != 0);
                            else ptx::umma_tf32(acc + h * p.NT, ad, bd, idesc, (kb | k4) != 0);
                        }
                    }
                    if (PAIR) ptx::umma_commit_pair(&bars->empty[s], 3);
                    else ptx::umma_commit(&bars->empty[s]);
                }
                if (PAIR) ptx::umma_commit_pair(&bars->acc_full[buf], 3);
                else ptx::umma_commit(&bars->acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;                                  // the TMEM lane quadrant this warp may read
        // two halves per CTA: warps 2-5 drain accumulator 0, warps 6-9 accumulator 1.  One half: both warps of a lane
        // quadrant work on the same accumulator and take alternate 32-column chunks.
        const int half = p.halves == 2 ? (warp - 2) >> 2 : 0;
        const int cfirst = p.halves == 2 ? 0 : ((warp - 2) >> 2) * 32, cstep = p.halves == 2 ? 32 : 64;
        uint8_t* my_out = out_stage + (warp - 2) * 4096;
        uint32_t tcount = 0;
        const uint32_t acc_empty0 = PAIR ? ptx::mapa_shared(&bars->acc_empty[0], 0) : 0u;   // the leader's acc_empty[]
        auto release_acc = [&](uint32_t b) {
            if (PAIR) ptx::mbar_arrive_cluster(acc_empty0 + b * 8);
            else ptx::mbar_arrive(&bars->acc_empty[b]);
        };
        for (int64_t tile = first; tile < p.tiles && !bars->abort; tile += step, ++tcount) {
            const uint32_t buf = tcount % p.nbuf, aph = (tcount / p.nbuf) & 1;
            const int n0 = (int)(tile % p.n_tiles_n) * p.NT;
#ifdef RQ_EXPERIMENTS
            const long long e0 = clock64();
#endif
            if (!ptx::mbar_wait(&bars->acc_full[buf], aph)) { bars->abort = 1; break; }
#ifdef RQ_EXPERIMENTS
            const long long e1 = clock64();
            if (p.dbg && blockIdx.x == 0 && warp == 2 && lane == 0) { p.dbg[5] += e1 - e0; }
#endif
            ptx::tc_fence_after_sync();
            const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + buf * p.halves * p.NT + half * p.NT;
            const int ncols = min(p.NT, p.N - n0);
            const int row0 = (int)(tile / p.n_tiles_n) * tile_rows * (PAIR ? 2 : 1) + (int)rank * tile_rows + half * 128 +
                             q * 32;                                                  // this warp's first row
            // one 32-row x 32-column chunk: registers -> this warp's 4 KB staging tile (128-byte rows, 16-byte pieces
            // XOR-swizzled by the row so that the 32 lanes' v4 stores are conflict-free) -> one TMA store, which writes
            // whole 128-byte lines and clips at the matrix edge.  (Thread-per-row st.global hit 32 lines per instruction
            // and left the epilogue warps waiting on the store queue.)
            auto put_chunk = [&](const float (&v)[32], int c0) {
#ifdef RQ_EXPERIMENTS
                if (p.st_mode == 2) {                                        // what-if: no store path at all
                    if (v[0] == 123.456f) p.C[0] = v[1];
                    return;
                }
                if (p.st_mode == 3) {                                        // what-if: staging writes only
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4*>(my_out + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    return;
                }
#endif
                if (p.st_mode == 0) {
                    if (lane == 0) ptx::bulk_wait_group_read0();             // the previous chunk has left the staging tile
                }
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(my_out + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                if (p.st_mode == 0) {
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_2d(&map_c, n0 + c0, row0, my_out);
                        ptx::bulk_commit_group();
                    }
                } else {
                    // read the tile back transposed: 8 lanes cover one 128-byte row, a store instruction 4 whole rows
                    __syncwarp();
                    const int pc = lane & 7, col = n0 + c0 + pc * 4;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int r = (lane >> 3) + 4 * i;
                        const float4 t = *reinterpret_cast<const float4*>(my_out + r * 128 + ((pc ^ (r & 7)) << 4));
                        if (row0 + r < p.M && col + 3 < p.N) ptx::st_global_v4(p.C + (int64_t)(row0 + r) * p.ldc + col, t);
                    }
                }
            };
            if (cfirst >= p.NT || (p.act == RQ_ACT_L2NORM && cfirst)) {     // nothing for this warp in this tile
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) release_acc(buf);
            } else if (p.act == RQ_ACT_L2NORM) {                 // the whole row is in this tile (N <= 64, host-checked)
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
                if (lane == 0) release_acc(buf);
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
                // the activation is a compile-time constant inside the chunk loop: with a run-time test per element the
                // 32 values of a chunk were separated by 64 uniform branches and the chunk took ~1,600 cycles
                auto run_chunks = [&](auto act_c, auto round_c) {
                    constexpr int ACT = decltype(act_c)::value;
                    constexpr bool ROUND = decltype(round_c)::value;
                    for (int c0 = cfirst; c0 < p.NT; c0 += cstep) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32b_x32(tbase + c0, r);
                        ptx::tmem_ld_wait_for(r);
                        if (c0 + cstep >= p.NT) {                // this warp's last read of the accumulator: give it back
                            ptx::tc_fence_before_sync();
                            __syncwarp();
                            if (lane == 0) release_acc(buf);
                        }
                        if (c0 >= ncols) continue;
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float t = __uint_as_float(r[j]);
                            if (ACT == RQ_ACT_SILU) t = silu(t);
                            if (ROUND) t = round_tf32(t);
                            v[j] = t;
                        }
                        put_chunk(v, c0);
                    }
                };
                using std::integral_constant;
                if (p.act == RQ_ACT_SILU) {
                    if (p.round_out) run_chunks(integral_constant<int, RQ_ACT_SILU>{}, integral_constant<bool, true>{});
                    else run_chunks(integral_constant<int, RQ_ACT_SILU>{}, integral_constant<bool, false>{});
                } else {
                    if (p.round_out) run_chunks(integral_constant<int, RQ_ACT_NONE>{}, integral_constant<bool, true>{});
                    else run_chunks(integral_constant<int, RQ_ACT_NONE>{}, integral_constant<bool, false>{});
                }
            }
#ifdef RQ_EXPERIMENTS
            if (p.dbg && blockIdx.x == 0 && warp == 2 && lane == 0) p.dbg[0] += clock64() - e1;
#endif
        }
        if (lane == 0) ptx::bulk_wait_group0();                  // every store of this warp has landed
#ifdef RQ_EXPERIMENTS
        if (p.dbg && blockIdx.x == 0 && warp == 2 && lane == 0) { p.dbg[1] = tcount; p.dbg[4] = clock64() - t_start; }
#endif
    }
    __syncwarp();
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (PAIR) ptx::cluster_sync();                       // neither CTA leaves while the other may still signal it
    if (warp == 1) {
        ptx::tc_fence_after_sync();
        if (PAIR) ptx::tmem_dealloc_pair(tmem, 512);
        else ptx::tmem_dealloc(tmem, 512);
    }
    if (tid == 0 && bars->abort) g_linear_timeout = 1;
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
}  // namespace lt

// Shared with the Gumbel GEMMs (rq_gemm_split.cu): a 2-D fp32 tensor map with box_rows x 32-float boxes, 128-byte swizzle.
// `map` points to a CUtensorMap (declared void* in rq_kernels.h, which does not include <cuda.h>).
bool make_tensor_map_f32(void* map, const float* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    return lt::make_map(reinterpret_cast<CUtensorMap*>(map), base, rows, cols, ld, box_rows);
}

bool linear_tf32_supported(int64_t M, int N, int Kd, int act_flags) {
    const int act = act_flags & ~(RQ_ACT_TF32_OUT | RQ_ACT_CTA_PAIR);
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
    // 256-wide tiles whenever the layer has 256 columns: a kind::tf32 MMA costs ~120 cycles at N = 128 as at N = 256
    // (measured: 192 MMAs per tile take 23k cycles either way), so halving N halves the tensor rate; that outweighs
    // the serialised epilogue (one accumulator pair fills TMEM).  Narrower layers are HBM-bound: <= 128 columns, two
    // accumulator pairs, the epilogue of one tile under the MMAs of the next.
    p.NT = N >= 256 ? 256 : (N >= 128 ? 128 : ((N + 15) & ~15));
#ifdef RQ_EXPERIMENTS
    if (const char* e = getenv("RQ_LIN_NT")) p.NT = std::min(atoi(e), (N + 15) & ~15);
#endif
    p.halves = 2;                            // single CTA: 128-row tiles measured 15-60 % slower at every encoder shape
    // CTA pairs (cta_group::2) on request, for the 256-wide tiles: measured equal to the single-CTA kernel on the
    // 768 -> 512 layer (566 vs 569 TFLOP/s; with the loads switched off both reach 620), so it is not the default
    const bool pair = (act & RQ_ACT_CTA_PAIR) && p.NT == 256 && (N % 256) == 0 && M >= 512;
    // a pair gives each CTA one 128-row x 256-column accumulator per tile: two of them fit in TMEM, so the epilogue of
    // one tile runs under the MMAs of the next while the W stage is still shared by two 128-row halves (one per CTA)
    if (pair) p.halves = 1;
    p.nbuf = p.halves * p.NT <= 256 ? 2 : 1;
    p.n_tiles_n = (N + p.NT - 1) / p.NT;
    p.C = Y; p.M = M; p.ldc = N; p.N = N; p.Kd = Kd; p.act = act & ~(RQ_ACT_TF32_OUT | RQ_ACT_CTA_PAIR); p.round_out = (act & RQ_ACT_TF32_OUT) ? 1 : 0;
    const int tile_rows = 128 * p.halves;                       // X rows per CTA and tile
    const int64_t rows_per_tile = tile_rows * (pair ? 2 : 1);
    p.tiles = ((M + rows_per_tile - 1) / rows_per_tile) * p.n_tiles_n;
    const int b_rows = pair ? p.NT / 2 : p.NT;
    const int stage_bytes = p.halves * (A_STAGE / 2) + b_stage_bytes(b_rows);
    p.stages = std::min<int>(MAX_STAGES, (224 * 1024 - OUT_STAGE) / stage_bytes);
#ifdef RQ_EXPERIMENTS
    if (const char* e = getenv("RQ_LIN_STAGES")) p.stages = std::min(p.stages, atoi(e));
    if (const char* e = getenv("RQ_LIN_NOLOAD")) p.noload = atoi(e);
    if (const char* e = getenv("RQ_LIN_STMODE")) p.st_mode = atoi(e);
#endif
    const size_t smem = (size_t)p.stages * stage_bytes + OUT_STAGE + sizeof(Bars) + 1024;
    CUtensorMap ma, mw, mc;
    RQ_CHECK_ARG(make_map(&ma, X, M, Kd, Kd, tile_rows) && make_map(&mw, W, N, Kd, Kd, b_rows) && make_map(&mc, Y, M, N, N, 32),
                 (int)cudaErrorInvalidValue,
                 "encoder layer: cuTensorMapEncodeTiled failed (rows=%lld in=%d out=%d)", (long long)M, Kd, N);
#ifdef RQ_EXPERIMENTS
    static unsigned long long* dbg = nullptr;
    if (getenv("RQ_LIN_STATS")) {
        if (!dbg) cudaMalloc(&dbg, 64);
        cudaMemsetAsync(dbg, 0, 64, st);
        p.dbg = dbg;
    }
#endif
    if (pair) {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_linear_tf32_kernel<true>, (int)smem));
        cudaLaunchConfig_t cfg{};
        cudaLaunchAttribute attr{};
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
        cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st; cfg.attrs = &attr; cfg.numAttrs = 1;
        cfg.gridDim = dim3(2 * (sm_count() / 2));
        int clusters = 0;                                       // how many pairs the device can hold at once
        RQ_CUDA(cudaOccupancyMaxActiveClusters(&clusters, rq_linear_tf32_kernel<true>, &cfg));
        RQ_CHECK_ARG(clusters > 0, (int)cudaErrorLaunchOutOfResources, "encoder layer: no CTA pair fits on this device");
        cfg.gridDim = dim3(2 * (unsigned)std::min<int64_t>(p.tiles, clusters));
#ifdef RQ_EXPERIMENTS
        if (getenv("RQ_LIN_VERBOSE")) fprintf(stderr, "encoder layer: %d CTA pairs resident, %lld tiles, %d stages\n", clusters, (long long)p.tiles, p.stages);
#endif
        RQ_CUDA(cudaLaunchKernelEx(&cfg, rq_linear_tf32_kernel<true>, ma, mw, mc, p));
    } else {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_linear_tf32_kernel<false>, (int)smem));
        int grid = (int)std::min<int64_t>(p.tiles, sm_count());
#ifdef RQ_EXPERIMENTS
        if (const char* e = getenv("RQ_LIN_GRID")) grid = std::min(grid, atoi(e));
#endif
        rq_linear_tf32_kernel<false><<<grid, THREADS, smem, st>>>(ma, mw, mc, p);
    }
    RQ_LAUNCHED();
#ifdef RQ_EXPERIMENTS
    if (p.dbg) {
        unsigned long long h[8];
        cudaMemcpy(h, p.dbg, 64, cudaMemcpyDeviceToHost);
        fprintf(stderr, "block 0: %llu tiles, %llu cycles total; per tile: epilogue active %llu, epilogue waits for the accumulator %llu, "
                "MMA thread waits for the epilogue %llu, for data %llu\n", h[1], h[4], h[0] / (h[1] ? h[1] : 1), h[5] / (h[1] ? h[1] : 1),
                h[2] / (h[1] ? h[1] : 1), h[3] / (h[1] ? h[1] : 1));
    }
#endif
    return 0;
}
int linear_tf32_timeout_flag() {
    int v = 0;
    cudaMemcpyFromSymbol(&v, lt::g_linear_timeout, sizeof(int));
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
