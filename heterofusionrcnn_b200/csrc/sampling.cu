// sampling.cu — farthest point sampling and gather_point for sm_100a.
//
// Replaces sampling/tf_sampling_g.cu:105-181 of the reference (see include/hfops.h).
//
// FPS design (B200-first, not a port):
//   * One scene = one thread-block CLUSTER of 1/2/4/8 CTAs.  Each CTA keeps its slice of
//     the cloud entirely on chip: xyz and the running min-distance of every point live in
//     REGISTERS (PTS points per thread), plus an SoA copy of xyz in shared memory that is
//     only used to look up the winner's coordinates.  Nothing is re-read from HBM during
//     the m-1 sequential iterations (the reference re-reads/re-writes a global `temp`
//     array every iteration).
//   * Per iteration the arg-max runs as: per-thread scan -> redux.sync (warp) -> one
//     __syncthreads -> redux.sync (CTA) -> one st.async push of {key, xyz} to every CTA of
//     the cluster through DSMEM, completing on the receiver's mbarrier (no cluster-wide
//     barrier in the loop) -> every CTA picks the max of the <=8 candidates locally.
//   * The reference's tie-break (max distance, then lowest index mod 512, then lowest
//     index; tf_sampling_g.cu:146-163) is reproduced under this different thread->point
//     mapping by reducing on an explicit 64-bit composite key.
#include "common.cuh"

namespace hf {
// fps_bucket.cu: the bucket-pruned kernel (default for clouds it covers)
int fps_bucket_dispatch(int b, int n, int m, const float *inp, int *out, cudaStream_t s, int w_req, int j_req);
// fps_multi.cu: the multi-pick kernel (default for clouds of up to 65536 points)
int fps_multi_dispatch(int b, int n, int m, const float *inp, int *out, float *oxyz, cudaStream_t s, int cluster_req, int warps_req);
int fps_pyramid_dispatch(int b, int levels, int n, const int *m, const float *inp, int *const *idx, float *const *xyz,
                         cudaStream_t s);
int fps_sorted_dispatch(int b, int n, int m, const float *inp, const void *index, int *out, float *oxyz, cudaStream_t s,
                        int warps_req, int mode_req);

namespace {

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                     : "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
// Remote (DSMEM) store that signals the receiver's mbarrier with the bytes written.
__device__ __forceinline__ void st_async_v4(uint32_t raddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                            uint32_t rbar) {
    asm volatile(
        "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(
            raddr),
        "r"(a), "r"(b), "r"(c), "r"(d), "r"(rbar)
        : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t raddr, uint32_t a, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
                 "r"(a), "r"(rbar)
                 : "memory");
}

// Low word of the composite key: larger = preferred.  Encodes (k mod 512, k) ascending.
__device__ __forceinline__ uint32_t tie_word(uint32_t k) { return ~(((k & 511u) << 23) | (k >> 9)); }
__device__ __forceinline__ uint32_t tie_decode(uint32_t lo) {
    uint32_t inv = ~lo;
    return ((inv & 0x7FFFFFu) << 9) | (inv >> 23);
}

constexpr int kMsgBytes = 20;  // {lo, hi, x, y} + {z}

// Per iteration (the sequential critical path, m-1 of them):
//   1. every thread updates the running min-distance of its PTS points and keeps only the
//      MAX VALUE (8 instructions per point; no index bookkeeping);
//   2. warp max (one redux.sync) -> shared -> ONE __syncthreads -> CTA max (one redux.sync);
//   3. only the warp that owns the CTA maximum recovers WHICH point it is (the reference's
//      tie-break as an explicit key) and pushes {key, xyz} to every CTA of the cluster with
//      st.async (DSMEM), completing on the receivers' mbarriers; all other warps go straight
//      to the wait.  If several warps hold the same maximum (duplicate points) a second
//      barrier decides between them — rare, and uniform across the CTA;
//   4. every warp picks the cluster maximum of the <= 8 received candidates with two
//      redux.sync and reads the winner's coordinates as a shared-memory broadcast.
// Measured alternative (profiles/r1_fps_sweep_v3_polling_vs_mbarrier.txt): plain remote
// stores of self-tagged 64-bit words with receivers spinning on ld.volatile.shared was
// 1.6-2.2x SLOWER per iteration than st.async + mbarrier (the HW-sleeping try_wait keeps
// the issue slots free for the warps that still compute), so it was dropped.
template <int CLUSTER, int THREADS, int PTS>
__global__ void __launch_bounds__(THREADS, 1)
    fps_cluster_kernel(int n, int m, const float *__restrict__ inp, int *__restrict__ out) {
    constexpr int CAP = THREADS * PTS;  // points held by one CTA (a power of two)
    constexpr int NW = THREADS / 32;
    static_assert((CAP & (CAP - 1)) == 0, "CAP must be a power of two");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *sx = reinterpret_cast<float *>(smem_raw);
    float *sy = sx + CAP;
    float *sz = sy + CAP;
    __shared__ uint32_t wmax[2][32];
    __shared__ uint32_t cand[2];
    __shared__ __align__(16) uint32_t slots[2][8][8];
    __shared__ __align__(8) uint64_t bars[2];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t rank = CLUSTER > 1 ? cluster_ctarank() : 0u;
    const int scene = blockIdx.x / CLUSTER;
    const float *data = inp + (size_t)scene * n * 3;
    int *idxs = out + (size_t)scene * m;
    const uint32_t base = rank * CAP;

    float x[PTS], y[PTS], z[PTS], td[PTS];
#pragma unroll
    for (int p = 0; p < PTS; ++p) {
        uint32_t k = base + p * THREADS + t;
        if (k < (uint32_t)n) {
            x[p] = data[(size_t)k * 3 + 0];
            y[p] = data[(size_t)k * 3 + 1];
            z[p] = data[(size_t)k * 3 + 2];
            td[p] = 1e38f;  // tf_sampling_g.cu:118
        } else {
            x[p] = y[p] = z[p] = 0.f;
            td[p] = -1.f;  // padding never wins: its key is (0, 0), below every real point's
        }
        sx[p * THREADS + t] = x[p];
        sy[p * THREADS + t] = y[p];
        sz[p * THREADS + t] = z[p];
    }
    if (t == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&bars[0], CLUSTER * kMsgBytes);
        mbar_arrive_expect_tx(&bars[1], CLUSTER * kMsgBytes);
        cand[0] = cand[1] = 0u;
    }
    if (CLUSTER > 1) cluster_sync_all();
    else __syncthreads();

    // start at index 0 (tf_sampling_g.cu:114-116)
    float cx = data[0], cy = data[1], cz = data[2];
    if (rank == 0 && t == 0) idxs[0] = 0;

    const int iters = m - 1;
    for (int it = 0; it < iters; ++it) {
        const int buf = it & 1;
        // ---- 1. update ----
        float mx = -1.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            float d = sqdist(x[p] - cx, y[p] - cy, z[p] - cz);
            td[p] = fminf(d, td[p]);
            mx = fmaxf(mx, td[p]);
        }
        // distances are >= +0, so their bit patterns order like unsigned ints; padding -> 0
        const uint32_t hi = __float_as_uint(fmaxf(mx, 0.f));
        // ---- 2. CTA max ----
        const uint32_t wm = __reduce_max_sync(0xffffffffu, hi);
        if (lane == 0) wmax[buf][warp] = wm;
        __syncthreads();
        if (t == 0) cand[buf ^ 1] = 0u;  // next iteration's tie buffer (its last readers are behind this barrier)
        const uint32_t v = lane < NW ? wmax[buf][lane] : 0u;
        const uint32_t cm = __reduce_max_sync(0xffffffffu, v);
        const unsigned owners = __ballot_sync(0xffffffffu, lane < NW && v == cm);
        // ---- 3. owner warp: recover the point, push ----
        bool push = false;
        uint32_t wl = 0u;
        if (wm == cm) {  // warp-uniform
            uint32_t lo = 0u;
            if (hi == cm && mx >= 0.f) {
#pragma unroll
                for (int p = 0; p < PTS; ++p) {
                    uint32_t w = tie_word(base + p * THREADS + t);
                    if (td[p] == mx && w > lo) lo = w;
                }
            }
            wl = __reduce_max_sync(0xffffffffu, lo);
            push = true;
        }
        if (owners & (owners - 1u)) {  // several warps hold the CTA maximum: CTA-uniform slow path
            if (push && lane == 0) atomicMax(&cand[buf], wl);
            __syncthreads();
            const uint32_t c = cand[buf];
            // c == 0: no real point in this CTA (all padding) -> warp 0 reports the empty key
            push = push && c == wl && (c != 0u || warp == 0);
        }
        if (push && lane < CLUSTER) {
            float wx = 0.f, wy = 0.f, wz = 0.f;
            if (wl != 0u) {
                uint32_t li = tie_decode(wl) - base;
                wx = sx[li];
                wy = sy[li];
                wz = sz[li];
            }
            uint32_t dst = mapa_u32(smem_u32(&slots[buf][rank][0]), (uint32_t)lane);
            uint32_t dbar = mapa_u32(smem_u32(&bars[buf]), (uint32_t)lane);
            st_async_v4(dst, wl, wl != 0u ? cm : 0u, __float_as_uint(wx), __float_as_uint(wy), dbar);
            st_async_b32(dst + 16, __float_as_uint(wz), dbar);
        }
        // ---- 4. cluster max ----
        mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
        if (t == 0 && it + 2 < iters) mbar_arrive_expect_tx(&bars[buf], CLUSTER * kMsgBytes);
        uint32_t win;
        if (CLUSTER == 1) {
            win = tie_decode(slots[buf][0][0]);
            cx = __uint_as_float(slots[buf][0][2]);
            cy = __uint_as_float(slots[buf][0][3]);
            cz = __uint_as_float(slots[buf][0][4]);
        } else {
            uint32_t klo = 0u, khi = 0u;
            if (lane < CLUSTER) {
                klo = slots[buf][lane][0];
                khi = slots[buf][lane][1];
            }
            const uint32_t gh = __reduce_max_sync(0xffffffffu, khi);
            const uint32_t gl = __reduce_max_sync(0xffffffffu, khi == gh ? klo : 0u);
            win = tie_decode(gl);
            const uint32_t bq = win / CAP;
            cx = __uint_as_float(slots[buf][bq][2]);
            cy = __uint_as_float(slots[buf][bq][3]);
            cz = __uint_as_float(slots[buf][bq][4]);
        }
        if (rank == 0 && t == 0) idxs[it + 1] = (int)win;
    }
    if (CLUSTER > 1) cluster_sync_all();  // no CTA may exit while peers can still push into it
}

// Fallback for clouds larger than the on-chip capacity (n > 8 * 512 * 16): one CTA per
// scene slot, running distances in a caller-provided global scratch (32*n floats, the
// reference's own temp contract, tf_sampling.cpp:115).
__global__ void __launch_bounds__(1024, 1)
    fps_global_kernel(int b, int n, int m, const float *__restrict__ inp, float *__restrict__ temp,
                      int *__restrict__ out) {
    __shared__ unsigned long long wkey[2][32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    float *td = temp + (size_t)blockIdx.x * n;
    for (int scene = blockIdx.x; scene < b; scene += gridDim.x) {
        const float *data = inp + (size_t)scene * n * 3;
        int *idxs = out + (size_t)scene * m;
        for (int k = t; k < n; k += 1024) td[k] = 1e38f;
        if (t == 0) idxs[0] = 0;
        float cx = data[0], cy = data[1], cz = data[2];
        for (int it = 0; it < m - 1; ++it) {
            float best = -1.f;
            uint32_t bk = 0;
            bool has = false;
            for (uint32_t k = t; k < (uint32_t)n; k += 1024) {
                float d = sqdist(data[(size_t)k * 3] - cx, data[(size_t)k * 3 + 1] - cy, data[(size_t)k * 3 + 2] - cz);
                float d2 = fminf(d, td[k]);
                td[k] = d2;
                if (d2 > best) {
                    best = d2;
                    bk = k;
                    has = true;
                }
            }
            uint32_t hi = has ? __float_as_uint(best) : 0u, lo = has ? tie_word(bk) : 0u;
            uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
            uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
            if (lane == 0) wkey[it & 1][warp] = ((unsigned long long)mh << 32) | ml;
            __syncthreads();
            unsigned long long wk = wkey[it & 1][lane];
            uint32_t h2 = (uint32_t)(wk >> 32), l2 = (uint32_t)wk;
            uint32_t ch = __reduce_max_sync(0xffffffffu, h2);
            uint32_t cl = __reduce_max_sync(0xffffffffu, h2 == ch ? l2 : 0u);
            uint32_t win = tie_decode(cl);
            cx = data[(size_t)win * 3];
            cy = data[(size_t)win * 3 + 1];
            cz = data[(size_t)win * 3 + 2];
            if (t == 0) idxs[it + 1] = (int)win;
        }
        __syncthreads();
    }
}

// One thread per output row (3 floats).  tf_sampling_g.cu:172-181.
__global__ void gather_point_kernel(long long total, int n, int m, const float *__restrict__ inp,
                                    const int *__restrict__ idx, float *__restrict__ out) {
    long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= total) return;
    long long i = g / m;
    int a = idx[g];
    const float *src = inp + ((size_t)i * n + a) * 3;
    float v0 = src[0], v1 = src[1], v2 = src[2];
    float *dst = out + (size_t)g * 3;
    dst[0] = v0;
    dst[1] = v1;
    dst[2] = v2;
}

thread_local int g_cluster_override = 0;  // per calling thread: a tuning override never changes what another thread launches

template <int CLUSTER, int THREADS, int PTS>
int launch_fps(int b, int n, int m, const float *inp, int *out, cudaStream_t stream) {
    auto kern = fps_cluster_kernel<CLUSTER, THREADS, PTS>;
    size_t smem = (size_t)3 * THREADS * PTS * sizeof(float);
    if (smem > 40 * 1024) HF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)b * CLUSTER);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HF_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, n, m, inp, out));
    HF_LAUNCH_OK("fps_cluster_kernel");
    return HF_OK;
}

thread_local int g_threads_override = 0;
thread_local int g_algo = 0, g_bucket_w = 0, g_bucket_j = 0;  // hf_fps_set_variant

template <int CLUSTER>
int dispatch_pts(int per_cta, int b, int n, int m, const float *inp, int *out, cudaStream_t s) {
    // 256-thread CTAs halve the per-iteration reduction overhead (8 warps instead of 16);
    // 512-thread CTAs hide more latency.  Heuristic: 256 threads while a thread holds <= 8 points.
    const bool t256 = g_threads_override == 256 || (g_threads_override == 0 && per_cta <= 2048);
    if (t256 && per_cta <= 4096) {
        if (per_cta <= 256) return launch_fps<CLUSTER, 256, 1>(b, n, m, inp, out, s);
        if (per_cta <= 512) return launch_fps<CLUSTER, 256, 2>(b, n, m, inp, out, s);
        if (per_cta <= 1024) return launch_fps<CLUSTER, 256, 4>(b, n, m, inp, out, s);
        if (per_cta <= 2048) return launch_fps<CLUSTER, 256, 8>(b, n, m, inp, out, s);
        return launch_fps<CLUSTER, 256, 16>(b, n, m, inp, out, s);
    }
    if (per_cta <= 512) return launch_fps<CLUSTER, 512, 1>(b, n, m, inp, out, s);
    if (per_cta <= 1024) return launch_fps<CLUSTER, 512, 2>(b, n, m, inp, out, s);
    if (per_cta <= 2048) return launch_fps<CLUSTER, 512, 4>(b, n, m, inp, out, s);
    if (per_cta <= 4096) return launch_fps<CLUSTER, 512, 8>(b, n, m, inp, out, s);
    if (per_cta <= 8192) return launch_fps<CLUSTER, 512, 16>(b, n, m, inp, out, s);
    return HF_ERR_UNSUPPORTED;
}

}  // namespace
}  // namespace hf

extern "C" void hf_fps_set_cluster(int cluster) { hf::g_cluster_override = cluster; }
extern "C" void hf_fps_set_threads(int threads) { hf::g_threads_override = threads; }
extern "C" void hf_fps_set_variant(int algo, int warps, int buckets_per_warp) {
    hf::g_algo = algo;
    hf::g_bucket_w = warps;
    hf::g_bucket_j = buckets_per_warp;
}
extern "C" void hf_fps_get_variant(int *algo, int *warps, int *buckets_per_warp) {
    if (algo) *algo = hf::g_algo;
    if (warps) *warps = hf::g_bucket_w;
    if (buckets_per_warp) *buckets_per_warp = hf::g_bucket_j;
}

namespace hf {
namespace {
int gather_point_launch(int b, int n, int m, const float *inp, const int *idx, float *out, cudaStream_t s) {
    long long total = (long long)b * m;
    if (total == 0) return HF_OK;
    int threads = 256;
    long long blocks = (total + threads - 1) / threads;
    gather_point_kernel<<<(unsigned)blocks, threads, 0, s>>>(total, n, m, inp, idx, out);
    HF_LAUNCH_OK("gather_point_kernel");
    return HF_OK;
}

// FPS of one level; oxyz (b,m,3) != null also receives the picked points (= gather_point(inp, out)): the
// multi-pick kernel writes them next to the indices, the other kernels are followed by a gather launch.
int fps_any(int b, int n, int m, const float *inp, float *temp, int *out, float *oxyz, hf_stream_t stream) {
    HF_REQUIRE(b >= 0 && n >= 1, "FarthestPointSample expects (batch_size,num_points,3) inp shape");
    HF_REQUIRE(m > 0, "FarthestPointSample expects positive npoint");
    HF_REQUIRE(inp && out, "FarthestPointSample: null pointer");
    if (b == 0) return HF_OK;
    cudaStream_t s = (cudaStream_t)stream;

    // The bucket-pruned kernel (fps_bucket.cu) is exact and touches ~1% of the cloud per
    // iteration, but its per-iteration dependent chain (test -> bucket -> warp arg-max ->
    // resolve -> exchange) currently costs 1.4k cycles against 1.2k for the plain kernel at
    // n = 16384 (profiles/r1_fps_bucket_notes.md), so it is opt-in (hf_fps_set_variant(2,..)).
    if (g_algo == 2) {
        int rc = fps_bucket_dispatch(b, n, m, inp, out, s, g_bucket_w, g_bucket_j);
        if (rc == HF_ERR_UNSUPPORTED)
            set_error("FarthestPointSample: bucket kernel has no shape for n=%d warps=%d buckets/warp=%d", n, g_bucket_w, g_bucket_j);
        if (rc == HF_OK && oxyz) rc = gather_point_launch(b, n, m, inp, out, oxyz, s);
        return rc;
    }

    // Default (algo 0; 3 = forced): the multi-pick kernel of fps_multi.cu, several picks per
    // DSMEM exchange.  algo 1 = the one-pick-per-exchange cluster kernel below, which also
    // serves clouds the multi-pick kernel has no shape for.
    if (g_algo == 0 || g_algo == 3) {
        int rc = fps_multi_dispatch(b, n, m, inp, out, oxyz, s, g_cluster_override, g_algo == 3 ? g_bucket_w : 0);
        if (rc != HF_ERR_UNSUPPORTED) return rc;
        if (g_algo == 3) {
            set_error("FarthestPointSample: multi-pick kernel has no shape for n=%d", n);
            return rc;
        }
    }

    // Cluster size: as many CTAs per scene as the GPU can co-schedule for this batch, but
    // not so many that a CTA holds < 1024 points (then the DSMEM exchange costs more than
    // the distance updates it saves).  A cloud of <= 4096 points stays in one CTA.
    int cluster = 1;
    if (g_cluster_override == 1 || g_cluster_override == 2 || g_cluster_override == 4 || g_cluster_override == 8) {
        cluster = g_cluster_override;
    } else if (n > 4096) {
        int sms = num_sms();
        cluster = 8;
        while (cluster > 1 && (long long)b * cluster > sms) cluster >>= 1;
        while (cluster > 1 && n / cluster < 1024) cluster >>= 1;
    }
    while (cluster < 8 && (n + cluster - 1) / cluster > 8192) cluster <<= 1;
    int per = (n + cluster - 1) / cluster;
    if (per > 8192) {
        HF_REQUIRE(temp != nullptr,
                   "FarthestPointSample: n=%d exceeds the on-chip capacity (65536); pass the reference's 32*n float temp", n);
        int grid = b < 32 ? b : 32;
        fps_global_kernel<<<grid, 1024, 0, s>>>(b, n, m, inp, temp, out);
        HF_LAUNCH_OK("fps_global_kernel");
        return oxyz ? gather_point_launch(b, n, m, inp, out, oxyz, s) : HF_OK;
    }
    int rc;
    switch (cluster) {
        case 1: rc = dispatch_pts<1>(per, b, n, m, inp, out, s); break;
        case 2: rc = dispatch_pts<2>(per, b, n, m, inp, out, s); break;
        case 4: rc = dispatch_pts<4>(per, b, n, m, inp, out, s); break;
        default: rc = dispatch_pts<8>(per, b, n, m, inp, out, s); break;
    }
    if (rc == HF_OK && oxyz) rc = gather_point_launch(b, n, m, inp, out, oxyz, s);
    return rc;
}
}  // namespace
}  // namespace hf

extern "C" int hf_farthest_point_sample(int b, int n, int m, const float *inp, float *temp, int *out,
                                        hf_stream_t stream) {
    return hf::fps_any(b, n, m, inp, temp, out, nullptr, stream);
}

extern "C" int hf_farthest_point_sample_xyz(int b, int n, int m, const float *inp, float *temp, int *out,
                                            float *new_xyz, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(new_xyz != nullptr, "FarthestPointSample: new_xyz is null");
    return fps_any(b, n, m, inp, temp, out, new_xyz, stream);
}

extern "C" int hf_farthest_point_sample_indexed(int b, int n, int m, const float *inp, const void *index, int *out,
                                                float *new_xyz, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1, "FarthestPointSample expects (batch_size,num_points,3) inp shape");
    HF_REQUIRE(m > 0, "FarthestPointSample expects positive npoint");
    HF_REQUIRE(inp && out, "FarthestPointSample: null pointer");
    if (b == 0) return HF_OK;
    if (index && (reinterpret_cast<uintptr_t>(index) & 15) == 0 && (g_algo == 0 || g_algo == 3)) {
        const int rc = fps_sorted_dispatch(b, n, m, inp, index, out, new_xyz, (cudaStream_t)stream, g_algo == 3 ? g_bucket_w : 0,
                                           g_algo == 3 ? g_bucket_j : 0);
        if (rc != HF_ERR_UNSUPPORTED) return rc;
    }
    return fps_any(b, n, m, inp, nullptr, out, new_xyz, stream);
}

extern "C" int hf_fps_pyramid(int b, int levels, int n, const int *m, const float *inp, int *const *idx,
                              float *const *new_xyz, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1, "FarthestPointSample expects (batch_size,num_points,3) inp shape");
    HF_REQUIRE(levels >= 1 && levels <= HF_FPS_PYRAMID_MAX_LEVELS && m && idx && new_xyz, "fps_pyramid: 1..%d levels",
               HF_FPS_PYRAMID_MAX_LEVELS);
    HF_REQUIRE(inp != nullptr, "FarthestPointSample: null pointer");
    for (int j = 0; j < levels; ++j) {
        HF_REQUIRE(m[j] > 0, "FarthestPointSample expects positive npoint");
        HF_REQUIRE(idx[j] && new_xyz[j], "fps_pyramid: null output pointer at level %d", j);
    }
    if (b == 0) return HF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (g_algo == 0 || g_algo == 3) {
        const int rc = fps_pyramid_dispatch(b, levels, n, m, inp, idx, new_xyz, s);
        if (rc != HF_ERR_UNSUPPORTED) return rc;
    }
    // no single-launch shape (or another FPS variant was asked for): level by level, each on the previous level's picks
    const float *cloud = inp;
    int pts = n;
    for (int j = 0; j < levels; ++j) {
        HF_REQUIRE(pts <= 65536, "fps_pyramid: clouds beyond 65536 points need the temp buffer of hf_farthest_point_sample");
        const int rc = fps_any(b, pts, m[j], cloud, nullptr, idx[j], new_xyz[j], stream);
        if (rc != HF_OK) return rc;
        cloud = new_xyz[j];
        pts = m[j];
    }
    return HF_OK;
}

extern "C" int hf_gather_point(int b, int n, int m, const float *inp, const int *idx, float *out,
                               hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1 && m >= 0, "GatherPoint expects (batch_size,num_points,3) inp shape");
    HF_REQUIRE(inp && idx && out, "GatherPoint: null pointer");
    return gather_point_launch(b, n, m, inp, idx, out, (cudaStream_t)stream);
}
