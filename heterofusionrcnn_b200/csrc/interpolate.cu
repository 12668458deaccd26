// interpolate.cu — three_nn and three_interpolate for sm_100a.
//
// Replaces interpolate/tf_interpolate_g.cu:22-124 of the reference (see include/hfops.h).
//
// three_nn: one thread per unknown point; the known points are streamed through shared
// memory in TMA-staged tiles (xyz_tiles.cuh) and read as broadcasts.  The reference keeps
// its three best distances as double but computes every distance in float
// (tf_interpolate_g.cu:43,49); float -> double is exact, so keeping them as float with a
// +inf start gives the same comparisons and the same stored values ((float)1e40 == +inf).
//
// three_interpolate: HBM-bound 3-tap weighted row gather.  The channel-last kernel is the
// one the Python API uses (it removes the two full transposes the reference wraps around
// its channel-first op, tf_interpolate.py:35-37): one warp per output row, 128-bit loads
// and streaming stores along c.  The channel-first kernel keeps the native launcher's
// layout for callers that bind at that level.
#include "common.cuh"
#include "xyz_tiles.cuh"

namespace hf {
namespace {

constexpr int kNnThreads = 256;
constexpr int kNnTile = 1024;

__global__ void __launch_bounds__(kNnThreads)
    three_nn_kernel(int n, int m, const float *__restrict__ unknown, const float *__restrict__ known,
                    float *__restrict__ dist2, int *__restrict__ idx, float *__restrict__ weight) {
    __shared__ __align__(16) float stage0[kNnTile * 3];
    __shared__ __align__(16) float stage1[kNnTile * 3];
    __shared__ __align__(8) uint64_t bars[2];

    const int scene = blockIdx.y;
    const int j = blockIdx.x * kNnThreads + threadIdx.x;
    const bool active = j < n;
    const float *u = unknown + ((size_t)scene * n + (active ? j : 0)) * 3;
    const float ux = u[0], uy = u[1], uz = u[2];

    float best1 = INFINITY, best2 = INFINITY, best3 = INFINITY;
    int besti1 = 0, besti2 = 0, besti3 = 0;

    XyzTilePipe<kNnTile> pipe;
    pipe.init(stage0, stage1, bars, known + (size_t)scene * m * 3, m);
    pipe.prefetch(0);
    for (int i = 0; i < pipe.ntiles; ++i) {
        pipe.prefetch(i + 1);
        const float *s = pipe.acquire(i);
        const int c = pipe.count(i);
        const int kbase = i * kNnTile;
        const float4 *s4 = reinterpret_cast<const float4 *>(s);
        const int groups = c >> 2;
        auto insert = [&](float d, int k) {
            // strict '<' cascade of tf_interpolate_g.cu:50-61: equal distances keep the lower index
            if (d < best1) {
                best3 = best2; besti3 = besti2;
                best2 = best1; besti2 = besti1;
                best1 = d; besti1 = k;
            } else if (d < best2) {
                best3 = best2; besti3 = besti2;
                best2 = d; besti2 = k;
            } else if (d < best3) {
                best3 = d; besti3 = k;
            }
        };
#pragma unroll 2
        for (int g = 0; g < groups; ++g) {
            float4 a = s4[3 * g], bb = s4[3 * g + 1], cc = s4[3 * g + 2];
            float d0 = sqdist(ux - a.x, uy - a.y, uz - a.z);
            float d1 = sqdist(ux - a.w, uy - bb.x, uz - bb.y);
            float d2 = sqdist(ux - bb.z, uy - bb.w, uz - cc.x);
            float d3 = sqdist(ux - cc.y, uy - cc.z, uz - cc.w);
            if (fminf(fminf(d0, d1), fminf(d2, d3)) < best3) {
                const int k = kbase + 4 * g;
                insert(d0, k);
                insert(d1, k + 1);
                insert(d2, k + 2);
                insert(d3, k + 3);
            }
        }
        for (int k = groups << 2; k < c; ++k) insert(sqdist(ux - s[3 * k], uy - s[3 * k + 1], uz - s[3 * k + 2]), kbase + k);
        __syncthreads();  // stage i&1 is free for tile i+2
    }
    if (active) {
        float *dd = dist2 + ((size_t)scene * n + j) * 3;
        int *ii = idx + ((size_t)scene * n + j) * 3;
        dd[0] = best1; dd[1] = best2; dd[2] = best3;
        ii[0] = besti1; ii[1] = besti2; ii[2] = besti3;
        if (weight) fp_weights3(best1, best2, best3, weight + ((size_t)scene * n + j) * 3);
    }
}

// Value order fixed to what nvcc emits for tf_interpolate_g.cu:109 on sm_100a:
// fma(w2,p2, fma(w0,p0, w1*p1)).
__device__ __forceinline__ float tap3(float w0, float p0, float w1, float p1, float w2, float p2) {
    return __fmaf_rn(w2, p2, __fmaf_rn(w0, p0, __fmul_rn(w1, p1)));
}

constexpr int kTiThreads = 256;

// channel-last, c % 4 == 0: points (b,m,c4) float4 -> out (b,n,c4).  grid = (blocks per scene, b).
// A warp takes U consecutive rows per round: their 3U indices and weights are one coalesced load by lanes 0..3U-1,
// issued one round AHEAD (the gathers depend on them: without the prefetch every round is two dependent memory round
// trips), and the 3U gathered 128-bit loads of a channel chunk are all in flight before the first FMA.
// ncu (profiles/r2_three_interpolate_ncu.md): the kernel is bound by memory LATENCY per round, not by L2 or DRAM
// throughput (both < 40 %), so what counts is how soon a warp is back at its next batch of loads: the scene is a grid
// dimension and every index is 32-bit inside it (no 64-bit division or multiply per row; round 1 spent 80 instructions
// per row, this kernel ~30), and U = 2 rows per warp at 4 CTAs/SM overlaps the phases of twice as many warps as U = 4
// at 2 CTAs/SM (47.6 vs 51.8 us at c = 128, 37.3 vs 39.9 us at c = 512 for 16 scans).
// Measured alternatives that lost and are not in the tree: cp.async (LDGSTS) staging of the rows in shared memory,
// 4 rounds in flight per warp: 127 us (c = 128) / 56 us (c = 512); rows visited along the Hilbert curve of the
// unknown cloud for L1 reuse: 64 us / 42 us (same file).
constexpr int kTiRows = 2;
template <int U>
__global__ void __launch_bounds__(kTiThreads, 8 / U)
    three_interpolate_nlc_vec4_kernel(int m, int c4, int n, const float4 *__restrict__ points,
                                      const int *__restrict__ idx, const float *__restrict__ weight,
                                      float4 *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int scene = blockIdx.y;
    const int warp = (blockIdx.x * kTiThreads + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * kTiThreads) >> 5;
    points += (size_t)scene * m * c4;
    idx += (size_t)scene * n * 3;
    weight += (size_t)scene * n * 3;
    out += (size_t)scene * n * c4;
    const int e_all = n * 3;
    // lane < 3U: float4 offset of the tap's source row in the scene's points, and its weight
    auto fetch = [&](int r0, unsigned &off, float &ww) {
        const int e = r0 * 3 + lane;
        off = 0u;
        ww = 0.f;
        if (lane < 3 * U && e < e_all) {
            off = (unsigned)__ldg(&idx[e]) * (unsigned)c4;
            ww = __ldg(&weight[e]);
        }
    };
    int r0 = warp * U;
    unsigned cur_o, nxt_o = 0u;
    float cur_w, nxt_w = 0.f;
    fetch(r0, cur_o, cur_w);
    for (; r0 < n; r0 += nwarps * U) {
        fetch(r0 + nwarps * U, nxt_o, nxt_w);
        unsigned p0[U], p1[U], p2[U];
        float w0[U], w1[U], w2[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            p0[u] = __shfl_sync(0xffffffffu, cur_o, 3 * u) + lane;
            p1[u] = __shfl_sync(0xffffffffu, cur_o, 3 * u + 1) + lane;
            p2[u] = __shfl_sync(0xffffffffu, cur_o, 3 * u + 2) + lane;
            w0[u] = __shfl_sync(0xffffffffu, cur_w, 3 * u);
            w1[u] = __shfl_sync(0xffffffffu, cur_w, 3 * u + 1);
            w2[u] = __shfl_sync(0xffffffffu, cur_w, 3 * u + 2);
        }
        float4 *o = out + (size_t)r0 * c4;
        for (int v = lane, vo = 0; v < c4; v += 32, vo += 32) {
            float4 a[U], b[U], c[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {  // rows past the end carry source row 0: valid memory, never stored
                a[u] = __ldg(points + (p0[u] + vo));
                b[u] = __ldg(points + (p1[u] + vo));
                c[u] = __ldg(points + (p2[u] + vo));
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (r0 + u < n) {
                    float4 r;
                    r.x = tap3(w0[u], a[u].x, w1[u], b[u].x, w2[u], c[u].x);
                    r.y = tap3(w0[u], a[u].y, w1[u], b[u].y, w2[u], c[u].y);
                    r.z = tap3(w0[u], a[u].z, w1[u], b[u].z, w2[u], c[u].z);
                    r.w = tap3(w0[u], a[u].w, w1[u], b[u].w, w2[u], c[u].w);
                    stg_cs_f4(o + u * c4 + v, r);
                }
            }
        }
        cur_o = nxt_o;
        cur_w = nxt_w;
    }
}

// channel-last, any c.
__global__ void __launch_bounds__(kTiThreads)
    three_interpolate_nlc_scalar_kernel(int m, int c, int n, const float *__restrict__ points,
                                        const int *__restrict__ idx, const float *__restrict__ weight,
                                        float *__restrict__ out) {
    const int scene = blockIdx.y;
    const float *pts = points + (size_t)scene * m * c;
    const int *ind = idx + (size_t)scene * n * 3;
    const float *wgt = weight + (size_t)scene * n * 3;
    float *o = out + (size_t)scene * n * c;
    const long long total = (long long)n * c;
    for (long long e = (long long)blockIdx.x * kTiThreads + threadIdx.x; e < total;
         e += (long long)gridDim.x * kTiThreads) {
        int r = (int)(e / c);
        int l = (int)(e - (long long)r * c);
        o[e] = tap3(wgt[r * 3], __ldg(&pts[(size_t)ind[r * 3] * c + l]), wgt[r * 3 + 1],
                    __ldg(&pts[(size_t)ind[r * 3 + 1] * c + l]), wgt[r * 3 + 2],
                    __ldg(&pts[(size_t)ind[r * 3 + 2] * c + l]));
    }
}

// channel-first (native launcher layout): points (b,c,m) -> out (b,c,n).
// grid = (ceil(n/256), c, b); consecutive threads = consecutive n => coalesced stores and
// idx/weight loads; the gathers along m hit one channel row (m*4 bytes, L1/L2 resident).
__global__ void __launch_bounds__(256)
    three_interpolate_cf_kernel(int c, int m, int n, const float *__restrict__ points,
                                const int *__restrict__ idx, const float *__restrict__ weight,
                                float *__restrict__ out) {
    const int scene = blockIdx.z, ch = blockIdx.y;
    const int j = blockIdx.x * 256 + threadIdx.x;
    if (j >= n) return;
    const float *pts = points + ((size_t)scene * c + ch) * m;
    const int *ind = idx + ((size_t)scene * n + j) * 3;
    const float *w = weight + ((size_t)scene * n + j) * 3;
    out[((size_t)scene * c + ch) * n + j] = tap3(w[0], __ldg(&pts[ind[0]]), w[1], __ldg(&pts[ind[1]]), w[2], __ldg(&pts[ind[2]]));
}

}  // namespace
}  // namespace hf

namespace hf {
int three_nn_every_pair(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                        float *weight, cudaStream_t s) {
    HF_REQUIRE(b >= 0 && n >= 0, "ThreeNN expects (b,n,3) xyz1 shape.");
    HF_REQUIRE(m >= 0, "ThreeNN expects (b,m,3) xyz2 shape.");
    HF_REQUIRE(unknown && known && dist2 && idx, "ThreeNN: null pointer");
    HF_REQUIRE(b <= 65535, "ThreeNN: batch_size > 65535 not supported");
    if (b == 0 || n == 0) return HF_OK;
    dim3 grid((n + kNnThreads - 1) / kNnThreads, b);
    three_nn_kernel<<<grid, kNnThreads, 0, s>>>(n, m, unknown, known, dist2, idx, weight);
    HF_LAUNCH_OK("three_nn_kernel");
    return HF_OK;
}
}  // namespace hf

extern "C" int hf_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                           hf_stream_t stream) {
    return hf::three_nn_every_pair(b, n, m, unknown, known, dist2, idx, nullptr, (cudaStream_t)stream);
}

extern "C" int hf_three_interpolate_nlc(int b, int m, int c, int n, const float *points, const int *idx,
                                        const float *weight, float *out, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && m >= 1 && c >= 1 && n >= 0, "ThreeInterpolate expects (b,m,c) points shape");
    HF_REQUIRE(points && idx && weight && out, "ThreeInterpolate: null pointer");
    HF_REQUIRE(b <= 65535, "ThreeInterpolate: batch_size > 65535 not supported");
    HF_REQUIRE((long long)n * 3 < (1ll << 31), "ThreeInterpolate: n too large");
    if (b == 0 || n == 0) return HF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int sms = num_sms();
    bool vec = (c % 4 == 0) && ((reinterpret_cast<uintptr_t>(points) & 15) == 0) &&
               ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    // 32-bit offsets inside a scene: m * c4 float4 of points
    if (vec && (long long)m * (c / 4) < (1ll << 31)) {
        auto kern = three_interpolate_nlc_vec4_kernel<kTiRows>;
        int per_sm = 0;  // occupancy on the current device: a cheap runtime query, not cached (re-entrant, any device)
        HF_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kTiThreads, 0));
        if (per_sm < 1) per_sm = 1;
        // exactly as many CTAs as the GPU holds at once: these launches last 15-50 us, a partial last wave is a visible
        // fraction of that
        long long blocks = ((long long)n + (long long)kTiRows * (kTiThreads / 32) - 1) / ((long long)kTiRows * (kTiThreads / 32));
        const long long cap = ((long long)sms * per_sm) / b;  // rounded down: never a second, nearly empty wave
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        kern<<<dim3((unsigned)blocks, b), kTiThreads, 0, s>>>(m, c / 4, n, reinterpret_cast<const float4 *>(points), idx, weight,
                                                             reinterpret_cast<float4 *>(out));
        HF_LAUNCH_OK("three_interpolate_nlc_vec4_kernel");
    } else {
        long long total = (long long)n * c;
        long long blocks = (total + kTiThreads - 1) / kTiThreads;
        long long cap = ((long long)sms * 8 + b - 1) / b;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        three_interpolate_nlc_scalar_kernel<<<dim3((unsigned)blocks, b), kTiThreads, 0, s>>>(m, c, n, points, idx, weight, out);
        HF_LAUNCH_OK("three_interpolate_nlc_scalar_kernel");
    }
    return HF_OK;
}

extern "C" int hf_three_interpolate(int b, int c, int m, int n, const float *points, const int *idx,
                                    const float *weight, float *out, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && m >= 1 && c >= 1 && n >= 0, "ThreeInterpolate expects (b,c,m) points shape");
    HF_REQUIRE(points && idx && weight && out, "ThreeInterpolate: null pointer");
    HF_REQUIRE(b <= 65535 && c <= 65535, "ThreeInterpolate: b and c must be <= 65535");
    if (b == 0 || n == 0) return HF_OK;
    dim3 grid((n + 255) / 256, c, b);
    three_interpolate_cf_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(c, m, n, points, idx, weight, out);
    HF_LAUNCH_OK("three_interpolate_cf_kernel");
    return HF_OK;
}
