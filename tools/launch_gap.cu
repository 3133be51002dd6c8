// Microbenchmark (profiling only): per-launch cost of back-to-back launches of a 148 x 512-thread persistent-style
// kernel as a function of its shared-memory size / TMEM allocation / programmatic dependent launch.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/launch_gap tools/launch_gap.cu && tools/launch_gap
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../rq_vae_recommender_b200/csrc/rq_tc_ptx.cuh"

using namespace rq;

template <bool TMEM, bool PDL>
__global__ void __launch_bounds__(512, 1) k_launch(int spin_cycles, long long* sink) {
    extern __shared__ uint8_t smem[];
    __shared__ uint32_t tbase;
    if (PDL) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (TMEM) {
        if ((threadIdx.x >> 5) == 1) ptx::tmem_alloc(&tbase, 512);
        ptx::tc_fence_before_sync();
        __syncthreads();
        ptx::tc_fence_after_sync();
    }
    if (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long t0 = clock64();
    while (clock64() - t0 < spin_cycles) {}
    if (TMEM) {
        ptx::tc_fence_before_sync();
        __syncthreads();
        if ((threadIdx.x >> 5) == 1) ptx::tmem_dealloc(tbase, 512);
    }
    if (sink && threadIdx.x == 0 && blockIdx.x == 0) sink[0] = clock64();
}

template <bool TMEM, bool PDL>
float run_graph(int smem, int spin, int n) {
    auto kern = k_launch<TMEM, PDL>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148);
    cfg.blockDim = dim3(512);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = PDL ? 1 : 0;
    cudaStream_t st;
    cudaStreamCreate(&st);
    cfg.stream = st;
#define CKG(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e_)); fflush(stdout); return -1.f; } } while (0)
    cudaGraph_t g = nullptr;
    cudaGraphExec_t ge = nullptr;
    CKG(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < n; ++i) CKG(cudaLaunchKernelEx(&cfg, kern, spin, (long long*)nullptr));
    CKG(cudaStreamEndCapture(st, &g));
    if (!g) { printf("no graph\n"); return -1.f; }
    CKG(cudaGraphInstantiate(&ge, g, 0));
    CKG(cudaGraphLaunch(ge, st));
    CKG(cudaStreamSynchronize(st));
    cudaEvent_t a, b;
    CKG(cudaEventCreate(&a));
    CKG(cudaEventCreate(&b));
    CKG(cudaEventRecord(a, st));
    CKG(cudaGraphLaunch(ge, st));
    CKG(cudaEventRecord(b, st));
    CKG(cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaGraphExecDestroy(ge);
    cudaGraphDestroy(g);
    cudaStreamDestroy(st);
    return ms * 1e3f / n;
}

template <bool TMEM, bool PDL>
float run(int smem, int spin, int n) {
    auto kern = k_launch<TMEM, PDL>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148);
    cfg.blockDim = dim3(512);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = PDL ? 1 : 0;
    cudaStream_t st;
    cudaStreamCreate(&st);
    cfg.stream = st;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e_)); fflush(stdout); return -1.f; } } while (0)
    // plain stream launches (the launches are asynchronous: the host runs ahead, the device sees them back to back)
    for (int i = 0; i < 20; ++i) CK(cudaLaunchKernelEx(&cfg, kern, spin, (long long*)nullptr));
    CK(cudaStreamSynchronize(st));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    CK(cudaEventRecord(a, st));
    for (int i = 0; i < n; ++i) CK(cudaLaunchKernelEx(&cfg, kern, spin, (long long*)nullptr));
    CK(cudaEventRecord(b, st));
    CK(cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaStreamDestroy(st);
    return ms * 1e3f / n;
}

int main() {
    const int n = 200;
    for (int spin : {0, 20000}) {
        fflush(stdout); printf("spin %d cycles (%.1f us at 1.965 GHz)\n", spin, spin / 1965.0);
        printf("  smem 0      no tmem        : %.2f us/launch\n", run<false, false>(0, spin, n));
        printf("  smem 226 KB no tmem        : %.2f us/launch\n", run<false, false>(226 * 1024, spin, n));
        printf("  smem 226 KB tmem           : %.2f us/launch\n", run<true, false>(226 * 1024, spin, n));
        printf("  smem 226 KB tmem      PDL  : %.2f us/launch\n", run<true, true>(226 * 1024, spin, n));
        printf("  smem 0      no tmem   PDL  : %.2f us/launch\n", run<false, true>(0, spin, n));
        fflush(stdout);
        printf("  graph of %d: smem 226 KB tmem      : %.2f us/launch\n", n, run_graph<true, false>(226 * 1024, spin, n));
        printf("  graph of %d: smem 226 KB tmem PDL  : %.2f us/launch\n", n, run_graph<true, true>(226 * 1024, spin, n));
        printf("  graph of 1 x many: tmem PDL         : ");
        { float t = 0; for (int i = 0; i < 1; ++i) t = run_graph<true, true>(226 * 1024, spin, 1); printf("%.2f us (single-node graph, one launch)\n", t); }
    }
    return 0;
}
