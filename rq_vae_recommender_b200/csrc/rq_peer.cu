// One-shot all-reduce of a small fp32 buffer over peer-mapped (symmetric) memory: NVLink loads, no NCCL.
//
// The path's exchange steps are tiny and latency-bound: the [L,K,D] codebook gradient of a training step
// (98,304 B at 3x256x32, 1 MB at 4x1024x64; the reference leaves it to DDP, train_rqvae.py:115-117,153) and the
// [K, D+1] sums / counts of a k-means iteration.  Every rank's buffer lives in a symmetric allocation whose peer
// addresses all ranks know (torch.distributed._symmetric_memory hands out the pointer tables).  One launch, the
// same on every rank and capturable into the step's CUDA graph right behind the backward kernel:
//
//   barrier 1   CTA b of every rank adds 1 to counter b in every rank's signal pad (system-scope release) and
//               waits until its own counter b reaches (2E+1) * world: every rank's producer kernel has finished
//   reduce      each thread sums its 16-byte chunks over the ranks in rank order (the same order everywhere, so
//               all ranks end with bit-identical values) with cache-bypassing loads through the peer pointers
//   barrier 2   same counters, (2E+2) * world: nobody still reads -> results are written (in place is allowed)
//
// E = number of earlier launches on this allocation, kept per CTA in device memory, so replaying a captured launch
// needs no host-side state.  Waits are bounded (2^23 polls of a system-scope load, a few seconds — the maximum rank
// skew the exchange tolerates): a CTA whose barrier times out does NOT publish a sum of possibly stale buffers, it
// fills its part of the output with NaN and raises the per-device error flag (rq_debug_peer_timeout_flag; PeerAllReduce
// .check() turns it into an exception), so a rank that fell behind can never yield silently different gradients.
#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {

// A light launch: it has to find room beside the persistent quantiser grids (which fill the register file of every
// SM they run on; parallel.reserve_sms_for_collectives keeps a few SMs free), so <= 32 registers per thread and two
// 1024-thread CTAs per SM; one or two 16-byte chunks per thread keep every NVLink load of the launch in flight at once.
constexpr int kPeerThreads = 1024;
constexpr int kPeerMaxCtas = 32;
constexpr int kPeerChunksPerThread = 2;          // 16-byte chunks held in registers across barrier 2
constexpr int kPeerMaxWorld = 16;
constexpr int kPeerPadWord0 = 2048;              // first uint32 of the signal pad used here (torch's own barriers use the front)

struct PeerArParams {
    const float* const* bufs;     // device array [world]: every rank's buffer (symmetric allocation base)
    uint32_t* const* pads;        // device array [world]: every rank's signal pad
    int rank, world;
    int64_t offset_floats;        // where the reduced region starts inside the allocation
    int64_t n4;                   // 16-byte chunks to reduce
    float* out;                   // local destination (may be this rank's own region)
    float scale;
    uint32_t* epoch;              // device-local [kPeerMaxCtas] launch counters
    int* err;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_peer_v4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(kPeerThreads, 2) rq_peer_allreduce_kernel(PeerArParams P) {
    const int b = blockIdx.x, t = threadIdx.x;
    const uint32_t E = P.epoch[b];
    uint32_t* my_ctr = P.pads[P.rank] + kPeerPadWord0 + b;
    __shared__ int s_ok;
    if (t == 0) s_ok = 1;
    auto barrier = [&](uint32_t target) {
        __syncthreads();                                   // this CTA's earlier accesses are done
        if (t < P.world) red_release_sys_add(P.pads[t] + kPeerPadWord0 + b, 1u);
        if (t == 0) {
            bool ok = false;
#pragma unroll 1
            for (uint32_t i = 0; i < (1u << 23); ++i) {
                if ((int32_t)(ld_acquire_sys(my_ctr) - target) >= 0) { ok = true; break; }
            }
            if (!ok) {
                s_ok = 0;
                if (P.err) atomicExch(P.err, 1);
            }
        }
        __syncthreads();
        return s_ok != 0;
    };
    bool ok = barrier((2u * E + 1u) * (uint32_t)P.world);
    float4 acc[kPeerChunksPerThread];
    const int64_t stride = (int64_t)gridDim.x * kPeerThreads;
#pragma unroll
    for (int u = 0; u < kPeerChunksPerThread; ++u) {
        const int64_t i = (int64_t)b * kPeerThreads + t + u * stride;
        acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < P.n4) {
            const int64_t off = P.offset_floats + 4 * i;
            int r = 0;
            for (; r + 4 <= P.world; r += 4) {             // four ranks' loads in flight, summed in rank order
                const float4 v0 = ld_peer_v4(P.bufs[r] + off), v1 = ld_peer_v4(P.bufs[r + 1] + off);
                const float4 v2 = ld_peer_v4(P.bufs[r + 2] + off), v3 = ld_peer_v4(P.bufs[r + 3] + off);
                acc[u].x = (((acc[u].x + v0.x) + v1.x) + v2.x) + v3.x;
                acc[u].y = (((acc[u].y + v0.y) + v1.y) + v2.y) + v3.y;
                acc[u].z = (((acc[u].z + v0.z) + v1.z) + v2.z) + v3.z;
                acc[u].w = (((acc[u].w + v0.w) + v1.w) + v2.w) + v3.w;
            }
            for (; r < P.world; ++r) {
                const float4 v = ld_peer_v4(P.bufs[r] + off);
                acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
            }
        }
    }
    ok = barrier((2u * E + 2u) * (uint32_t)P.world) && ok;
    const float poison = ok ? 1.f : __int_as_float(0x7fc00000);     // a timed-out barrier: NaN, never a stale sum
#pragma unroll
    for (int u = 0; u < kPeerChunksPerThread; ++u) {
        const int64_t i = (int64_t)b * kPeerThreads + t + u * stride;
        if (i < P.n4)
            reinterpret_cast<float4*>(P.out)[i] =
                make_float4(acc[u].x * P.scale * poison, acc[u].y * P.scale * poison, acc[u].z * P.scale * poison,
                            acc[u].w * P.scale * poison);
    }
    if (t == 0) P.epoch[b] = E + 1u;
}

static int* d_peer_err[64] = {};          // one error word per device

static int* peer_err_word() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (!d_peer_err[dev]) {
        if (cudaMalloc((void**)&d_peer_err[dev], sizeof(int)) != cudaSuccess) return nullptr;
        cudaMemset(d_peer_err[dev], 0, sizeof(int));
    }
    return d_peer_err[dev];
}

}  // namespace rq

extern "C" int64_t rq_peer_allreduce_max_floats(void) {
    return (int64_t)rq::kPeerMaxCtas * rq::kPeerThreads * rq::kPeerChunksPerThread * 4;
}
extern "C" int rq_peer_allreduce_state_words(void) { return rq::kPeerMaxCtas; }
extern "C" int rq_peer_allreduce_pad_bytes(void) { return (rq::kPeerPadWord0 + rq::kPeerMaxCtas) * 4; }

extern "C" int rq_peer_allreduce_f32(const void* peer_bufs_dev, const void* peer_pads_dev, int rank, int world,
                                     int64_t offset_floats, int64_t n_floats, float* out, float scale,
                                     uint32_t* epoch_dev, rq_stream_t stream) {
    using namespace rq;
    cudaStream_t st = (cudaStream_t)stream;
    RQ_CHECK_ARG(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, RQ_ERR_INVALID_ARGUMENT,
                 "rq_peer_allreduce_f32: need 1 <= world <= %d and 0 <= rank < world", kPeerMaxWorld);
    RQ_CHECK_ARG(n_floats >= 0 && (n_floats & 3) == 0 && (offset_floats & 3) == 0 && offset_floats >= 0,
                 RQ_ERR_INVALID_ARGUMENT, "rq_peer_allreduce_f32: n_floats and offset_floats must be multiples of 4");
    RQ_CHECK_ARG(n_floats <= rq_peer_allreduce_max_floats(), RQ_ERR_UNSUPPORTED_SHAPE,
                 "rq_peer_allreduce_f32: at most %lld floats per launch", (long long)rq_peer_allreduce_max_floats());
    RQ_CHECK_ARG(peer_bufs_dev && peer_pads_dev && out && epoch_dev, RQ_ERR_NULL_POINTER, "rq_peer_allreduce_f32: null pointer");
    RQ_CHECK_ARG(((uintptr_t)out & 15) == 0, RQ_ERR_INVALID_ARGUMENT, "rq_peer_allreduce_f32: out must be 16-byte aligned");
    if (n_floats == 0) return 0;
    int* err_word = peer_err_word();
    RQ_CHECK_ARG(err_word != nullptr, RQ_ERR_NO_DEVICE, "rq_peer_allreduce_f32: no error word for the current device");
    PeerArParams P;
    P.bufs = reinterpret_cast<const float* const*>(peer_bufs_dev);
    P.pads = reinterpret_cast<uint32_t* const*>(peer_pads_dev);
    P.rank = rank; P.world = world; P.offset_floats = offset_floats; P.n4 = n_floats / 4;
    P.out = out; P.scale = scale; P.epoch = epoch_dev; P.err = err_word;
    // every rank must launch the same grid (the barrier counters are per CTA): a function of n only
    int64_t ctas = (P.n4 + kPeerThreads - 1) / kPeerThreads;
    if (ctas > kPeerMaxCtas) ctas = kPeerMaxCtas;
    rq_peer_allreduce_kernel<<<(unsigned)ctas, kPeerThreads, 0, st>>>(P);
    RQ_LAUNCHED();
    return 0;
}

extern "C" int rq_debug_peer_timeout_flag(void) {
    int v = 0, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && rq::d_peer_err[dev])
        cudaMemcpy(&v, rq::d_peer_err[dev], sizeof(int), cudaMemcpyDeviceToHost);
    return v;
}
