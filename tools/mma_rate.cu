// Microbenchmark (profiling only): cycles per tcgen05.mma (kind::f16, bf16 x bf16 -> fp32, M = 128) for the operand
// shapes of the d=32 quantiser, with and without concurrent shared-memory traffic from the other warps.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_rate tools/mma_rate.cu && tools/mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../rq_vae_recommender_b200/csrc/rq_tc_ptx.cuh"

using namespace rq;

struct Cfg {
    int n;           // UMMA N
    int n_mma;       // MMAs per "tile-level" (accumulate chain)
    int reps;        // tile-levels
    int noise;       // 0 none, 1 LDS.128 loops in the other warps, 2 STS.128 loops
    int two_acc;     // alternate between two accumulators
    int b_stride;    // byte advance of the B descriptor between MMAs of a chain (32 = K slices of one SW128 tile)
};

__global__ void __launch_bounds__(512, 1) mma_rate_kernel(Cfg c, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;                  // 16 KB
    uint8_t* sB = smem + 16384;          // 3 x 32 KB
    uint8_t* sN = smem + 16384 + 98304;  // 64 KB noise area
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 98304 + 65536);
    uint32_t* tbase = reinterpret_cast<uint32_t*>(bar + 2);
    volatile int* stop = reinterpret_cast<volatile int*>(bar + 4);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (16384 + 98304 + 65536) / 4; i += 512) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (tid == 0) {
        ptx::mbar_init(bar, 1);
        ptx::fence_mbar_init();
        *stop = 0;
    }
    if (warp == 1) ptx::tmem_alloc(tbase, 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tm = *tbase;
    if (tid == 0) {
        const uint32_t idesc = ptx::make_idesc_bf16_f32(128, c.n);
        const uint32_t a0 = ptx::smem_u32(sA), b0 = ptx::smem_u32(sB);
        long long t0 = clock64();
        for (int r = 0; r < c.reps; ++r) {
            const uint32_t acc = tm + ((c.two_acc && (r & 1)) ? 256u : 0u);
            for (int i = 0; i < c.n_mma; ++i) {
                const uint64_t ad = ptx::make_smem_desc(a0 + (i & 3) * 32, 16, 1024, ptx::kLayoutSw128);
                const uint64_t bd = ptx::make_smem_desc(b0 + (r % 3) * 32768 + (i & 3) * c.b_stride, 16, 1024, ptx::kLayoutSw128);
                ptx::umma_bf16(acc, ad, bd, idesc, i > 0);
            }
        }
        long long t1 = clock64();
        ptx::umma_commit(bar);
        ptx::mbar_wait(bar, 0);
        long long t2 = clock64();
        *stop = 1;
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } else if (warp >= 4 && c.noise) {
        // the other 12 warps hammer shared memory until the MMA chain is done
        float4 acc4 = make_float4(0, 0, 0, 0);
        int it = 0;
        while (!*stop) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4* p = reinterpret_cast<float4*>(sN + ((tid * 16 + j * 8192 + it * 16) & 65535 & ~15));
                if (c.noise == 1) { float4 v = *p; acc4.x += v.x; acc4.y += v.y; }
                else *p = acc4;
            }
            ++it;
        }
        if (acc4.x == 123.f) out[7] = 1;
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tm, 512);
}

int main() {
    long long* d;
    cudaMalloc(&d, 64);
    const int smem = 16384 + 98304 + 65536 + 1024 + 256;
    cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const Cfg cfgs[] = {
        {256, 7, 12, 0, 0, 32}, {256, 7, 12, 0, 1, 32}, {256, 7, 12, 1, 1, 32}, {256, 7, 12, 2, 1, 32},
        {128, 7, 12, 0, 0, 32}, {256, 1, 84, 0, 0, 32}, {256, 28, 3, 0, 0, 32}, {256, 7, 12, 0, 0, 0},
        {64, 7, 12, 0, 0, 32},  {256, 7, 1, 0, 0, 32},
    };
    for (const Cfg& c : cfgs) {
        for (int grid : {1, 148}) {
            long long h[2] = {0, 0};
            for (int rep = 0; rep < 2; ++rep) {
                mma_rate_kernel<<<grid, 512, smem>>>(c, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
            const int total = c.n_mma * c.reps;
            printf("N=%3d mma/chain=%2d chains=%2d noise=%d two_acc=%d bstride=%2d grid=%3d: issue %6lld cyc, done %6lld cyc = %.1f cyc/MMA\n",
                   c.n, c.n_mma, c.reps, c.noise, c.two_acc, c.b_stride, grid, h[0], h[1], (double)h[1] / total);
        }
    }
    return 0;
}
