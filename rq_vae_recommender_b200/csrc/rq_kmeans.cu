// K-means assignment + accumulation (init/kmeans.py:40-58 of the reference), exact SIMT family.
//
// The reference materialises a [B,K,D] broadcast of (x-c)^2 (655 MB at B=20k) and then runs a
// Python loop over the K clusters with a host sync each.  Here one thread owns a point, scans
// the centroids staged in shared memory in the reference's direct form sum_d (x_d - c_d)^2,
// and scatters its row into per-cluster sums with 16-byte L2 reductions; a second tiny kernel
// turns (sums, counts) into centroids, the max shift (kmeans.py:67) and the empty-cluster list.
#include "rq_common.cuh"
#include "rq_kernels.h"
#include "rq_rowmath.cuh"

namespace rq {

constexpr int kKmThreads = 256;

template <int D>
__global__ void __launch_bounds__(kKmThreads) rq_kmeans_assign_simt_kernel(KmeansParams p) {
    extern __shared__ __align__(16) float smem[];
    float* sC = smem;
    const int tid = threadIdx.x;
    const int64_t row = (int64_t)blockIdx.x * kKmThreads + tid;
    const bool active = row < p.B;
    float r[D];
    if (active) load_row<D>(p.x + row * D, r);
    else {
#pragma unroll
        for (int j = 0; j < D; ++j) r[j] = 0.f;
    }
    float best = INFINITY;
    int bi = 0;
    for (int k0 = 0; k0 < p.K; k0 += p.KC) {
        const int kc = min(p.KC, p.K - k0);
        __syncthreads();
        for (int i = tid; i < kc * (D / 4); i += kKmThreads)
            reinterpret_cast<float4*>(sC)[i] = __ldg(reinterpret_cast<const float4*>(p.centroids + (int64_t)k0 * D) + i);
        __syncthreads();
        for (int k = 0; k < kc; ++k) {
            const float4* c = reinterpret_cast<const float4*>(sC + k * D);
            float a = 0.f;
#pragma unroll
            for (int j = 0; j < D / 4; ++j) {
                float4 v = c[j];
                float t0 = r[4 * j] - v.x, t1 = r[4 * j + 1] - v.y, t2 = r[4 * j + 2] - v.z, t3 = r[4 * j + 3] - v.w;
                a = fmaf(t0, t0, a); a = fmaf(t1, t1, a); a = fmaf(t2, t2, a); a = fmaf(t3, t3, a);
            }
            if (a < best) { best = a; bi = k0 + k; }
        }
    }
    int changed = 0;
    if (active) {
        changed = (!p.has_prev) || (p.assign[row] != (int64_t)bi);
        p.assign[row] = bi;
        float* dst = p.sums + (int64_t)bi * D;
#pragma unroll
        for (int j = 0; j < D; j += 4) red_add_v4(dst + j, r[j], r[j + 1], r[j + 2], r[j + 3]);
        atomicAdd(p.counts + bi, 1ull);
    }
    unsigned m = __ballot_sync(0xffffffffu, changed);
    if ((tid & 31) == 0 && m) atomicAdd(p.n_changed, (unsigned long long)__popc(m));
}

template <int D>
static int launch_km(const KmeansParams& p0, cudaStream_t st) {
    KmeansParams p = p0;
    int kc = 16384 / D;
    if (kc > p.K) kc = p.K;
    p.KC = kc;
    size_t smem = (size_t)kc * D * sizeof(float);
    auto kern = rq_kmeans_assign_simt_kernel<D>;
    RQ_CUDA(ensure_max_dynamic_smem((const void*)kern, (int)smem));
    dim3 grid((unsigned)((p.B + kKmThreads - 1) / kKmThreads));
    kern<<<grid, kKmThreads, smem, st>>>(p);
    RQ_LAUNCHED();
    return 0;
}

int kmeans_assign_accumulate_simt(const KmeansParams& p, int D, cudaStream_t st) {
    RQ_CUDA(cudaMemsetAsync(p.sums, 0, (size_t)p.K * D * sizeof(float), st));
    RQ_CUDA(cudaMemsetAsync(p.counts, 0, (size_t)p.K * sizeof(unsigned long long), st));
    RQ_CUDA(cudaMemsetAsync(p.n_changed, 0, sizeof(unsigned long long), st));
    if (p.B == 0) return 0;
    switch (D) {
        case 4: return launch_km<4>(p, st);
        case 8: return launch_km<8>(p, st);
        case 16: return launch_km<16>(p, st);
        case 32: return launch_km<32>(p, st);
        case 64: return launch_km<64>(p, st);
        case 128: return launch_km<128>(p, st);
        default:
            set_error("unsupported embed_dim D=%d (supported: 4, 8, 16, 32, 64, 128)", D);
            return RQ_ERR_UNSUPPORTED_SHAPE;
    }
}

// ---- tensor-core assignment + coalesced accumulation --------------------------------------------------------
// The assignment IS the fused quantiser's forward with L = 1 in ids-only mode (rq_quantize_tc.cu: TMA-staged
// centroids, tcgen05 distances, fp32 re-scoring of near ties), writing straight into `assign` and counting
// changed rows.  The accumulation is a second HBM-streaming kernel in the coalesced pattern: D/4 lanes per
// point, one 16-byte L2 reduction (REDG.128) per lane into sums[assign], one count atomic per point.
template <int D>
__global__ void __launch_bounds__(256) rq_kmeans_accumulate_kernel(const float* __restrict__ x,
                                                                   const int64_t* __restrict__ assign, int64_t B,
                                                                   float* __restrict__ sums,
                                                                   unsigned long long* __restrict__ counts) {
    constexpr int LPR = D / 4;
    const int64_t n_chunks = B * LPR;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / LPR;
        const int sub = (int)(i - row * LPR);
        const int64_t k = __ldg(assign + row);
        float4 v;
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "l"(x + i * 4));
        red_add_v4(sums + k * D + sub * 4, v.x, v.y, v.z, v.w);
        if (sub == 0) atomicAdd(counts + k, 1ull);
    }
}

// Accumulation with CTA-private partial sums: K*D fp32 sums + K counts live in shared memory, rows stream
// through in the coalesced pattern (D/4 lanes per row, one 16-byte load each) and are added with shared-memory
// atomics; each CTA flushes its partials once with 16-byte L2 reductions.  The global-reduction kernel above sends
// B*D*4 bytes of REDG traffic at K hot lines (39k row-adds per line at 10M x 256); here it is (#CTAs * K * D * 4).
// Component order is rotated by lane/8 so that the 32 lanes of one atomic hit 32 different banks whatever the ids.
template <int D>
__global__ void __launch_bounds__(512) rq_kmeans_accumulate_smem_kernel(const float* __restrict__ x,
                                                                        const int64_t* __restrict__ assign, int64_t B,
                                                                        int K, float* __restrict__ sums,
                                                                        unsigned long long* __restrict__ counts) {
    extern __shared__ __align__(16) float s_sums[];
    unsigned* s_cnt = reinterpret_cast<unsigned*>(s_sums + (size_t)K * D);
    for (int i = threadIdx.x; i < K * D + K; i += blockDim.x) s_sums[i] = 0.f;       // +0.0f == 0u bitwise
    __syncthreads();
    constexpr int LPR = D / 4;
    constexpr int U = 4;
    const int64_t n_chunks = B * LPR;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int rot = (threadIdx.x >> 3) & 3;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n_chunks; i0 += U * stride) {
        float4 v[U];
        int k[U];
        int sub[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + u * stride;
            k[u] = -1;
            if (i < n_chunks) {
                const int64_t row = i / LPR;
                sub[u] = (int)(i - row * LPR);
                k[u] = (int)__ldg(assign + row);
                asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                             : "l"(x + i * 4));
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (k[u] < 0) continue;
            float* dst = s_sums + (size_t)k[u] * D + sub[u] * 4;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int j = (t + rot) & 3;
                const float val = (j == 0) ? v[u].x : (j == 1) ? v[u].y : (j == 2) ? v[u].z : v[u].w;
                atomicAdd(dst + j, val);
            }
            if (sub[u] == 0) atomicAdd(s_cnt + k[u], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K * D / 4; i += blockDim.x) {
        const float4 t = reinterpret_cast<const float4*>(s_sums)[i];
        if (t.x != 0.f || t.y != 0.f || t.z != 0.f || t.w != 0.f) red_add_v4(sums + 4 * i, t.x, t.y, t.z, t.w);
    }
    for (int kk = threadIdx.x; kk < K; kk += blockDim.x)
        if (s_cnt[kk]) atomicAdd(counts + kk, (unsigned long long)s_cnt[kk]);
}

// The same accumulation on the integer unit.  atomicAdd(float*) on shared memory is a compare-and-swap loop on sm_100a
// (ATOMS.CAST.SPIN); the 32-bit integer add is native (ATOMS.ADD).  A component v is therefore added as the fixed-point
// number f = rint(v * S), S = 2^22 / range, in two pieces (f >> 11 and f & 0x7ff, each good for 2^20 rows per CTA
// before an int32 could overflow), and the CTA's partials leave as one float reduction per (cluster, component).
// `range` is a power of two >= 16 max|centroid component|, computed by every CTA from the centroids; a component outside
// (-range, range) — an outlier far from every centroid, Inf or NaN — is added straight to the global sum in floating
// point, so the result never depends on the choice of range, only the speed does.  |f / S - v| <= 2^-23 range, unbiased.
template <int D>
__global__ void __launch_bounds__(512, 2) rq_kmeans_accumulate_fx_kernel(const float* __restrict__ x,
                                                                      const int64_t* __restrict__ assign, int64_t B,
                                                                      int K, const float* __restrict__ centroids,
                                                                      float* __restrict__ sums,
                                                                      unsigned long long* __restrict__ counts) {
    extern __shared__ __align__(16) int s_fx[];
    int* s_hi = s_fx;
    int* s_lo = s_fx + (size_t)K * D;
    unsigned* s_cnt = reinterpret_cast<unsigned*>(s_fx + 2 * (size_t)K * D);
    __shared__ float s_max[16];
    for (int i = threadIdx.x; i < 2 * K * D + K; i += blockDim.x) s_fx[i] = 0;
    float cm = 0.f;
    for (int i = threadIdx.x; i < K * D; i += blockDim.x) cm = fmaxf(cm, fabsf(__ldg(centroids + i)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = cm;
    __syncthreads();
    cm = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) cm = fmaxf(cm, s_max[w]);
    // power of two in [16 cm, 32 cm); exponent clamped so that S and 1/S stay normal numbers (NaN / Inf centroids: cm
    // propagates as the largest finite exponent, every component then takes the floating-point path or a tiny f)
    int e = (int)((__float_as_uint(cm) >> 23) & 0xffu) + 5;
    e = e < 30 ? 30 : (e > 220 ? 220 : e);
    const float range = __uint_as_float((unsigned)e << 23);
    const float S = 4194304.f / range;
    // A warp takes whole rows, lane l the components l, l + 32, ...: every load instruction reads 128 contiguous bytes
    // and the 32 lanes of a reduction hit 32 different banks whatever the id (word k D + 32 c + l), with no per-lane
    // component shuffling.  The id is one broadcast load per row.  U rows are in flight per warp.  The loop is bound by
    // instruction issue, so the body is branch-free (32-bit shared addresses, predicated count, outliers collected in
    // one flag per U rows and handled after the fact).
    constexpr int C = D / 32;
    constexpr int U = 8;
    const int lane = threadIdx.x & 31;
    const uint32_t hi_base = (uint32_t)__cvta_generic_to_shared(s_hi) + lane * 4;
    const uint32_t lo_base = (uint32_t)__cvta_generic_to_shared(s_lo) + lane * 4;
    const uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(s_cnt);
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    auto fx_add = [&](int k, int c, float val) -> bool {
        const uint32_t off = (uint32_t)(k * D + 32 * c) * 4u;
        const float sc = val * S;
        const bool ok = fabsf(sc) < 4194304.f;
        // rint(sc) for |sc| < 2^22: the low mantissa bits of sc + 1.5 * 2^23
        const int f = ok ? (__float_as_int(sc + 12582912.f) - 0x4B400000) : 0;
        asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(hi_base + off), "r"(f >> 11) : "memory");
        asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(lo_base + off), "r"(f & 0x7ff) : "memory");
        return !ok;
    };
    auto count_one = [&](int k) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %2, 0;\n\t@p red.shared.add.u32 [%0], %1;\n\t}"
                     ::"r"(cnt_base + (uint32_t)k * 4u), "r"(1u), "r"((uint32_t)lane) : "memory");
    };
    const int64_t n_full = B / U;                         // blocks of U whole rows
    auto load_block = [&](int64_t blk, int (&k)[U], float (&v)[U][C]) {
        const int64_t r0 = blk * U;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            k[u] = (int)__ldg(reinterpret_cast<const int*>(assign + r0 + u));       // low word of the int64 id
#pragma unroll
            for (int c = 0; c < C; ++c)
                asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v[u][c]) : "l"(x + (r0 + u) * D + 32 * c + lane));
        }
    };
    auto add_block = [&](const int (&k)[U], const float (&v)[U][C]) {
        bool bad = false;
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
            for (int c = 0; c < C; ++c) bad |= fx_add(k[u], c, v[u][c]);
            count_one(k[u]);
        }
        if (__any_sync(0xffffffffu, bad)) {              // rare: components outside the fixed-point range
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int c = 0; c < C; ++c)
                    if (!(fabsf(v[u][c] * S) < 4194304.f)) atomicAdd(sums + k[u] * D + 32 * c + lane, v[u][c]);
        }
    };
    // two blocks in flight per warp: the loads of one are outstanding while the other is added
    {
        int kA[U], kB[U];
        float vA[U][C], vB[U][C];
        int64_t blk = warp0;
        if (blk < n_full) load_block(blk, kA, vA);
        while (blk < n_full) {
            const int64_t blkB = blk + n_warps;
            const bool hasB = blkB < n_full;
            if (hasB) load_block(blkB, kB, vB);
            add_block(kA, vA);
            blk = blkB + n_warps;
            if (blk < n_full) load_block(blk, kA, vA);
            if (hasB) add_block(kB, vB);
        }
    }
    // the last B % U rows: one warp of the grid
    if (warp0 == n_warps - 1) {
        for (int64_t row = n_full * U; row < B; ++row) {
            const int k = (int)__ldg(reinterpret_cast<const int*>(assign + row));
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const float val = __ldg(x + row * D + 32 * c + lane);
                if (fx_add(k, c, val)) atomicAdd(sums + k * D + 32 * c + lane, val);
            }
            count_one(k);
        }
    }
    __syncthreads();
    const double inv = 1.0 / (double)S;
    for (int i = threadIdx.x; i < K * D; i += blockDim.x) {
        const long long tot = (long long)s_hi[i] * 2048 + (long long)s_lo[i];
        if (tot != 0) atomicAdd(sums + i, (float)((double)tot * inv));
    }
    for (int kk = threadIdx.x; kk < K; kk += blockDim.x)
        if (s_cnt[kk]) atomicAdd(counts + kk, (unsigned long long)s_cnt[kk]);
}

// sums / counts must be zeroed by the caller.  Picks the shared-memory kernel when K*D partials fit.
template <int D>
static int launch_kmeans_accumulate(const KmeansParams& p, cudaStream_t st) {
    const size_t smem = (size_t)p.K * D * 4 + (size_t)p.K * 4;
    const int64_t n_chunks = p.B * (D / 4);
    const size_t smem_fx = 2 * (size_t)p.K * D * 4 + (size_t)p.K * 4;
    // fixed-point partials: an int32 piece holds 2^20 rows per CTA (see the kernel), the grid below gives each CTA B / grid
    if (smem_fx <= 200 * 1024) {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_kmeans_accumulate_fx_kernel<D>, 200 * 1024));
        int per_sm = (int)((224 * 1024) / (smem_fx + 1024));
        if (per_sm > 3) per_sm = 3;
        int64_t blocks = (n_chunks + 512 * 4 - 1) / (512 * 4);
        const int64_t cap = (int64_t)sm_count() * per_sm;
        if (blocks > cap) blocks = cap;
        const int64_t min_blocks = (p.B >> 20) + 1;           // <= 2^20 rows per CTA
        if (blocks < min_blocks) blocks = min_blocks;
        rq_kmeans_accumulate_fx_kernel<D><<<(unsigned)blocks, 512, smem_fx, st>>>(p.x, p.assign, p.B, p.K, p.centroids,
                                                                                 p.sums, p.counts);
    } else if (smem <= 100 * 1024) {
        RQ_CUDA(ensure_max_dynamic_smem((const void*)rq_kmeans_accumulate_smem_kernel<D>, 100 * 1024));
        int per_sm = (int)((220 * 1024) / (smem + 1024));
        if (per_sm > 4) per_sm = 4;
        int64_t blocks = (n_chunks + 512 * 4 - 1) / (512 * 4);
        const int64_t cap = (int64_t)sm_count() * per_sm;
        if (blocks > cap) blocks = cap;
        rq_kmeans_accumulate_smem_kernel<D><<<(unsigned)blocks, 512, smem, st>>>(p.x, p.assign, p.B, p.K, p.sums, p.counts);
    } else {
        int64_t blocks = (n_chunks + 255) / 256;
        const int64_t cap = (int64_t)sm_count() * 16;
        if (blocks > cap) blocks = cap;
        rq_kmeans_accumulate_kernel<D><<<(unsigned)blocks, 256, 0, st>>>(p.x, p.assign, p.B, p.sums, p.counts);
    }
    RQ_LAUNCHED();
    return 0;
}

int kmeans_assign_accumulate_tc(const KmeansParams& p, int D, cudaStream_t st) {
    RQ_CUDA(cudaMemsetAsync(p.sums, 0, (size_t)p.K * D * sizeof(float), st));
    RQ_CUDA(cudaMemsetAsync(p.counts, 0, (size_t)p.K * sizeof(unsigned long long), st));
    RQ_CUDA(cudaMemsetAsync(p.n_changed, 0, sizeof(unsigned long long), st));
    if (p.B == 0) return 0;
    FwdParams f;
    f.x = p.x; f.cb = p.centroids; f.B = p.B; f.K = p.K; f.L = 1; f.mode = RQ_MODE_EVAL; f.dist = RQ_DIST_L2;
    f.beta = 0.f; f.temp = 1.f; f.gu = nullptr; f.seed = 0; f.offset = 0;
    f.emb = nullptr; f.res = nullptr; f.ids = p.assign; f.loss = nullptr; f.KC = 0;
    f.n_changed = p.n_changed; f.has_prev = p.has_prev;
    // D = 32, K <= 256: the accumulation rides in the assignment launch (the rows are in shared memory when their id is
    // known); only the last B % 128 rows, which that launch hands to the exact family, go through the stand-alone kernel
    const bool fuse = (D == 32) && p.K <= 256 && p.B >= 128 && (p.B >> 20) < (int64_t)sm_count();
    if (fuse) {
        f.km_sums = p.sums;
        f.km_counts = p.counts;
    }
    int rc = quantize_forward_tc(f, D, st);
    if (rc) return rc;
    if (fuse) {
        const int64_t done = (p.B / 128) * 128;
        if (done == p.B) return 0;
        KmeansParams t = p;
        t.x = p.x + done * D;
        t.assign = p.assign + done;
        t.B = p.B - done;
        return launch_kmeans_accumulate<32>(t, st);
    }
    if (D == 32) return launch_kmeans_accumulate<32>(p, st);
    if (D == 64) return launch_kmeans_accumulate<64>(p, st);
    set_error("kmeans tensor-core path: D=%d not served", D);
    return RQ_ERR_UNSUPPORTED_SHAPE;
}

// One CTA; thread k owns cluster k (strided).  c_new = sums/count (kmeans.py:57) or unchanged
// for an empty cluster; shift_k = ||c_new - c_old||_2 (kmeans.py:67).
__global__ void rq_kmeans_update_kernel(const float* __restrict__ sums, const int64_t* __restrict__ counts,
                                        float* __restrict__ centroids, int D, int K, float* max_shift,
                                        int64_t* n_empty, int64_t* empty_ids) {
    __shared__ float s_max[32];
    __shared__ int s_nempty;
    if (threadIdx.x == 0) s_nempty = 0;
    __syncthreads();
    float mx = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        int64_t n = counts[k];
        if (n == 0) {
            atomicAdd(&s_nempty, 1);
            continue;
        }
        float inv = (float)n;
        float sh = 0.f;
        for (int j = 0; j < D; ++j) {
            float c = sums[(int64_t)k * D + j] / inv;
            float t = c - centroids[(int64_t)k * D + j];
            sh = fmaf(t, t, sh);
            centroids[(int64_t)k * D + j] = c;
        }
        mx = fmaxf(mx, sqrtf(sh));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = 0.f;
        for (int w = 0; w < (blockDim.x + 31) / 32; ++w) m = fmaxf(m, s_max[w]);
        *max_shift = m;
        *n_empty = s_nempty;
        // ascending list of empty clusters (K is small; serial scan by one thread)
        int c = 0;
        for (int k = 0; k < K; ++k)
            if (counts[k] == 0) empty_ids[c++] = k;
    }
}

int kmeans_update(const float* sums, const int64_t* counts, float* centroids, int D, int K, float* max_shift,
                  int64_t* n_empty, int64_t* empty_ids, cudaStream_t st) {
    rq_kmeans_update_kernel<<<1, 256, 0, st>>>(sums, counts, centroids, D, K, max_shift, n_empty, empty_ids);
    RQ_LAUNCHED();
    return 0;
}

}  // namespace rq
