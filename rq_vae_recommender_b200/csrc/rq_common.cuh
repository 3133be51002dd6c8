// Shared helpers for the rq-engine CUDA translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/rq_engine.h"

namespace rq {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;

#define RQ_CHECK_ARG(cond, code, ...)     \
    do {                                  \
        if (!(cond)) {                    \
            ::rq::set_error(__VA_ARGS__); \
            return (code);                \
        }                                 \
    } while (0)

#define RQ_CUDA(expr)                                                                      \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            ::rq::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                            __LINE__);                                                     \
            return (int)_e;                                                                \
        }                                                                                  \
    } while (0)

// Call after every kernel launch: counts it and surfaces launch-configuration errors.
#define RQ_LAUNCHED()                                                                 \
    do {                                                                              \
        ::rq::g_launches.fetch_add(1, std::memory_order_relaxed);                     \
        cudaError_t _e = cudaPeekAtLastError();                                       \
        if (_e != cudaSuccess) {                                                      \
            ::rq::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), \
                            __FILE__, __LINE__);                                      \
            return (int)_e;                                                           \
        }                                                                             \
    } while (0)

int sm_count();
bool pdl_enabled();                                               // RQ_NO_PDL=1 (read once) switches PDL off
cudaError_t ensure_max_dynamic_smem(const void* fn, int bytes);   // per (function, device), set once

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------
// Kernels of the training step are launched with cudaLaunchAttributeProgrammaticStreamSerialization: the next kernel's
// CTAs are scheduled as this kernel's CTAs retire (not when the whole grid has drained) and run their set-up — barrier
// init, TMEM allocation — up to pdl_wait(), which returns once the preceding kernel has completed and its writes are
// visible.  Measured on B200 (tools/launch_gap.cu): 4.2 -> 1.1 us of per-launch overhead for a 148 x 512-thread kernel
// with 226 KB of shared memory and a TMEM allocation.  Both instructions are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Launches `kern` with the PDL attribute (capturable into a CUDA graph as a programmatic dependency edge).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- small device helpers -------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// MUFU-based reciprocal / square root (<= 1-2 ulp): scalar coefficients of the estimators only, never distances
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_sqrt(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// ---- packed fp32x2 arithmetic (sm_100: FFMA2 / FADD2 — two IEEE fp32 operations per issue slot) --------------
// The row kernels are issue-bound; a register pair (even-aligned, which is how LDS.128 / LDG.128 results land)
// goes through one instruction instead of two.  Per-lane results are bit-identical to fmaf / + / -.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float hsum2(f32x2 v) {
    float a, b;
    unpack2(v, a, b);
    return a + b;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    // REDG.E.ADD.F32x4 — one L2 reduction per 16 bytes (sm_90+)
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Philox4x32-10 (Salmon et al. 2011).  Counter = (c0,c1,c2,c3), key = (k0,k1).
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}
// 24-bit uniform in [0,1), the granularity torch.rand uses for fp32
__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }

// gumbel(0,1) from a uniform draw — distributions/gumbel.py:9-10 (eps = 1e-20)
__device__ __forceinline__ float gumbel_from_u(float u) { return -logf(-logf(u + 1e-20f) + 1e-20f); }

}  // namespace rq
