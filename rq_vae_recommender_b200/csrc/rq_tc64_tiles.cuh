// Tile geometry and device helpers of the D = 64 tcgen05 kernels (rq_quantize_tc64.cu: K <= 256, levels resident;
// rq_quantize_bigk.cu: K > 256, code chunks streamed through a shared-memory ring).
#pragma once
#include <cuda_fp16.h>

#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_tc_ptx.cuh"

namespace rq {

namespace tc64 {
constexpr int D = 64;
constexpr int KC = 256;                          // codes per level (UMMA N)
constexpr int TILE_M = 128;                      // rows per tile (UMMA M)
constexpr int MAX_L = 3;
constexpr int NWG = 2;
constexpr int THREADS = 128 * NWG;
constexpr int B_TILE_BYTES = KC * 128;           // [256 codes][64 fp16], SW128, K-major
constexpr int A_BLOCK_BYTES = TILE_M * 128;      // one 64-element K-block of 128 rows, SW128
constexpr int A_TILE_BYTES = 2 * A_BLOCK_BYTES;  // block 0 = hi(t r), block 1 = lo(t r)
constexpr int CC_TILE_BYTES = KC * 32;           // [256 codes][16 fp16], no swizzle
constexpr int ONES_TILE_BYTES = TILE_M * 32;

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
    s.misc = off; off += 128;
    s.total = off;
    return s;
}
__device__ __forceinline__ uint32_t sw128_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}
__device__ __forceinline__ uint32_t nosw_off(int r, int c) {
    return (uint32_t)((r >> 3) * 256 + c * 128 + (r & 7) * 16);
}
// kind::f16 instruction descriptor, FP16 x FP16 -> FP32, K-major operands
__host__ __device__ constexpr uint32_t make_idesc_f16_f32(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {      // element 0 = a in the low half
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
// (hi, lo) fp16x2 words of the pair (a, b): hi = rn(a), lo = rn(a - hi)
__device__ __forceinline__ void split_pair_h(float a, float b, uint32_t& hi, uint32_t& lo) {
    __half2 h = __floats2half2_rn(a, b);
    float2 hf = __half22float2(h);
    hi = *reinterpret_cast<uint32_t*>(&h);
    lo = pack_h2(a - hf.x, b - hf.y);
}
__device__ __forceinline__ float h_round(float v) { return __half2float(__float2half_rn(v)); }

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
}  // namespace tc64

}  // namespace rq
