// grouping.cu — query_ball_point and group_point for sm_100a.
//
// Replaces grouping/tf_grouping_g.cu:3-57 of the reference (see include/hfops.h).
//
// query_ball_point: one thread per query scans the dataset points in ascending index
// (which is what "first nsample in the ball" means, tf_grouping_g.cu:15-33); the dataset
// is streamed through shared memory in TMA-staged tiles and read as broadcasts, four
// candidates per three 128-bit loads.  The reference's test
//     max(sqrtf(d2), 1e-20f) < radius
// is evaluated as d2 < thr2 with thr2 computed on the host as the smallest float whose
// correctly-rounded sqrtf is >= radius — sqrtf is monotonic and correctly rounded on
// both sides, so the two predicates accept exactly the same d2 and the inner loop has no
// square root.  Hit lists are assembled per CTA in shared memory and written coalesced.
//
// group_point: pure row gather, HBM bound.  One warp per output row walks the channels
// with 128-bit loads/stores (channel-last, as the reference API already is), four rows in
// flight per warp; c not divisible by 4 (xyz: 3, intensity: 1) takes a scalar path.
#include <math.h>

#include "common.cuh"
#include "xyz_tiles.cuh"

namespace hf {
namespace {

constexpr int kBqThreads = 128;
constexpr int kBqTile = 1024;

// ROWBUF: hit rows are staged in dynamic shared memory [threads][nsample+1] and written
// coalesced at the end; otherwise (very large nsample) they go straight to global.
template <bool ROWBUF>
__global__ void __launch_bounds__(kBqThreads)
    query_ball_kernel(int n, int m, float thr2, int nsample, const float *__restrict__ xyz1,
                      const float *__restrict__ xyz2, int *__restrict__ idx, int *__restrict__ pts_cnt) {
    __shared__ __align__(16) float stage0[kBqTile * 3];
    __shared__ __align__(16) float stage1[kBqTile * 3];
    __shared__ __align__(8) uint64_t bars[2];
    extern __shared__ int rowbuf[];  // ROWBUF: kBqThreads * (nsample + 1)

    const int t = threadIdx.x;
    const int scene = blockIdx.y;
    const int j = blockIdx.x * kBqThreads + t;
    const bool active = j < m;
    const float *q = xyz2 + ((size_t)scene * m + (active ? j : 0)) * 3;
    const float x2 = q[0], y2 = q[1], z2 = q[2];
    int *grow = idx + ((size_t)scene * m + (active ? j : 0)) * nsample;
    int *srow = rowbuf + t * (nsample + 1);
    int cnt = active ? 0 : nsample;

    XyzTilePipe<kBqTile> pipe;
    pipe.init(stage0, stage1, bars, xyz1 + (size_t)scene * n * 3, n);
    pipe.prefetch(0);
    for (int i = 0; i < pipe.ntiles; ++i) {
        pipe.prefetch(i + 1);
        const float *s = pipe.acquire(i);
        const int c = pipe.count(i);
        const int kbase = i * kBqTile;
        if (cnt < nsample) {
            const float4 *s4 = reinterpret_cast<const float4 *>(s);
            const int groups = c >> 2;
            int g = 0;
#pragma unroll 2
            for (; g < groups; ++g) {
                float4 a = s4[3 * g], bb = s4[3 * g + 1], cc = s4[3 * g + 2];
                float d0 = sqdist(x2 - a.x, y2 - a.y, z2 - a.z);
                float d1 = sqdist(x2 - a.w, y2 - bb.x, z2 - bb.y);
                float d2 = sqdist(x2 - bb.z, y2 - bb.w, z2 - cc.x);
                float d3 = sqdist(x2 - cc.y, y2 - cc.z, z2 - cc.w);
                if (fminf(fminf(d0, d1), fminf(d2, d3)) < thr2) {
                    float dd[4] = {d0, d1, d2, d3};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (dd[u] < thr2 && cnt < nsample) {
                            if (ROWBUF) srow[cnt] = kbase + 4 * g + u;
                            else grow[cnt] = kbase + 4 * g + u;
                            ++cnt;
                        }
                    }
                    if (cnt >= nsample) break;
                }
            }
            if (cnt < nsample) {
                for (int k = groups << 2; k < c; ++k) {
                    float d = sqdist(x2 - s[3 * k], y2 - s[3 * k + 1], z2 - s[3 * k + 2]);
                    if (d < thr2 && cnt < nsample) {
                        if (ROWBUF) srow[cnt] = kbase + k;
                        else grow[cnt] = kbase + k;
                        ++cnt;
                    }
                }
            }
        }
        // all queries of the CTA full -> nothing left to find (also frees the stage)
        if (__syncthreads_and(cnt >= nsample)) break;
    }
    pipe.drain();

    if (active) pts_cnt[(size_t)scene * m + j] = cnt;
    if (ROWBUF) {
        srow[nsample] = active ? cnt : 0;
        __syncthreads();
        const int rows = min(kBqThreads, m - blockIdx.x * kBqThreads);
        int *gbase = idx + ((size_t)scene * m + (size_t)blockIdx.x * kBqThreads) * nsample;
        const int total = rows * nsample;
        for (int e = t; e < total; e += kBqThreads) {
            int r = e / nsample, l = e - r * nsample;
            const int *row = rowbuf + r * (nsample + 1);
            int rc = row[nsample];
            // unused slots repeat the first hit (:26-29); a row with no hit is written as 0
            gbase[e] = l < rc ? row[l] : (rc > 0 ? row[0] : 0);
        }
    } else if (active) {
        int first = cnt > 0 ? grow[0] : 0;
        for (int l = cnt; l < nsample; ++l) grow[l] = first;
    }
}

// ---- group_point ------------------------------------------------------------------
constexpr int kGpThreads = 256;
constexpr int kGpRowsPerWarp = 4;

// c % 4 == 0, 16-byte aligned rows.  grid = (blocks, b).
// XYZ: the same pass also gathers the (b,n,3) coordinates of the rows into out_xyz (b,rows,3) — sample_and_group calls
// group_point twice with the same idx (hf/core/feature_extractors/pointnet_util.py:49,56), and the 12-byte rows of the
// coordinate call are latency-sized work that does not deserve its own launch: lanes 0..11 of the warp copy the 4 x 3
// floats of its four rows (48 contiguous bytes of out_xyz).
template <bool XYZ>
__global__ void __launch_bounds__(kGpThreads)
    group_point_vec4_kernel(int n, int c4, int rows, const float4 *__restrict__ points,
                            const int *__restrict__ idx, float4 *__restrict__ out, const float *__restrict__ xyz,
                            float *__restrict__ out_xyz) {
    const int scene = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * kGpThreads + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * kGpThreads) >> 5;
    const float4 *pts = points + (size_t)scene * n * c4;
    const int *ind = idx + (size_t)scene * rows;
    float4 *o = out + (size_t)scene * rows * c4;
    for (int r0 = warp * kGpRowsPerWarp; r0 < rows; r0 += nwarps * kGpRowsPerWarp) {
        int src[kGpRowsPerWarp];
#pragma unroll
        for (int u = 0; u < kGpRowsPerWarp; ++u) src[u] = (r0 + u < rows) ? ind[r0 + u] : -1;
        if (XYZ && lane < 3 * kGpRowsPerWarp) {
            const int u = lane / 3, j = lane - 3 * u;
            int su = src[0];
#pragma unroll
            for (int q = 1; q < kGpRowsPerWarp; ++q)
                if (u == q) su = src[q];
            if (su >= 0)
                out_xyz[((size_t)scene * rows + r0 + u) * 3 + j] = __ldg(&xyz[((size_t)scene * n + su) * 3 + j]);
        }
        for (int v = lane; v < c4; v += 32) {
            float4 val[kGpRowsPerWarp];
#pragma unroll
            for (int u = 0; u < kGpRowsPerWarp; ++u)
                if (src[u] >= 0) val[u] = __ldg(&pts[(size_t)src[u] * c4 + v]);
#pragma unroll
            for (int u = 0; u < kGpRowsPerWarp; ++u)
                if (src[u] >= 0) stg_cs_f4(&o[(size_t)(r0 + u) * c4 + v], val[u]);
        }
    }
}

// TMA row mover (c % 4 == 0): the op is a pure copy, so no byte has to pass through a register.  Every WARP runs its
// own ring of kGbStages shared-memory stages.  A stage holds R consecutive output rows: each lane issues the 1-D bulk
// copies (cp.async.bulk global -> shared, one per gathered source row) of its rows, all completing on the stage's
// mbarrier; the R rows are contiguous in `out`, so ONE bulk copy shared -> global writes the whole stage.  Loads of two
// stages are in flight while the third is being written: the bytes in flight per SM are bounded by its shared memory
// (4 warps x 3 stages x <= 16 KB), not by registers, and the kernel issues ~2 instructions per row, which leaves the
// issue slots of the SM to whatever else runs beside it (the sampling clusters).
constexpr int kGbThreads = 128;
constexpr int kGbStages = 3;
constexpr int kGbStageBytes = 16 * 1024;

__global__ void __launch_bounds__(kGbThreads)
    group_point_bulk_kernel(int n, int rowbytes, int R, long long rows_all, int rows_per_scene,
                            const unsigned char *__restrict__ points, const int *__restrict__ idx,
                            unsigned char *__restrict__ out) {
    extern __shared__ __align__(128) unsigned char gb_smem[];
    __shared__ __align__(8) uint64_t bars[kGbThreads / 32][kGbStages];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *ring = gb_smem + (size_t)warp * kGbStages * kGbStageBytes;
    if (lane == 0) {
        for (int s = 0; s < kGbStages; ++s) mbar_init(&bars[warp][s], 1);
        mbar_fence_init();
    }
    __syncwarp();
    const long long groups = (rows_all + R - 1) / R;
    const long long gw = (long long)blockIdx.x * (kGbThreads / 32) + warp, nw = (long long)gridDim.x * (kGbThreads / 32);
    // a warp's k-th group is g(k) = gw + k * nw
    auto issue = [&](long long k) {
        const long long g = gw + k * nw;
        if (g >= groups) return;
        const int s = (int)(k % kGbStages);
        const long long r0 = g * R;
        const int cnt = (int)((rows_all - r0) < R ? (rows_all - r0) : R);
        if (lane == 0) mbar_arrive_expect_tx(&bars[warp][s], (uint32_t)cnt * (uint32_t)rowbytes);
        __syncwarp();
        for (int u = lane; u < cnt; u += 32) {
            const long long row = r0 + u;
            const long long scene = row / rows_per_scene;
            const unsigned char *src = points + ((size_t)scene * n + (size_t)__ldg(&idx[row])) * rowbytes;
            tma_load_1d(ring + (size_t)s * kGbStageBytes + (size_t)u * rowbytes, src, (uint32_t)rowbytes, &bars[warp][s]);
        }
    };
    issue(0);
    issue(1);
    for (long long k = 0;; ++k) {
        const long long g = gw + k * nw;
        if (g >= groups) break;
        const int s = (int)(k % kGbStages);
        // stage (k + 2) % 3 == (k - 1) % 3 was written out as bulk group k - 1, the NEWEST committed group: every
        // committed store must have finished reading shared memory before that stage is refilled (wait_group.read 0; a
        // wait of 1 would let exactly group k - 1 still be reading).  That store was issued a whole mbar_wait ago.
        if (lane == 0) tma_store_wait_read<0>();
        __syncwarp();
        issue(k + 2);
        mbar_wait(&bars[warp][s], (uint32_t)(k / kGbStages) & 1u);
        if (lane == 0) {
            const long long r0 = g * R;
            const int cnt = (int)((rows_all - r0) < R ? (rows_all - r0) : R);
            tma_store_1d(out + (size_t)r0 * rowbytes, ring + (size_t)s * kGbStageBytes, (uint32_t)cnt * (uint32_t)rowbytes);
            tma_store_commit();
        }
    }
    if (lane == 0) tma_store_wait_read<0>();  // shared memory must outlive the last store's read
}

// any c: flat (row, channel) mapping so stores are coalesced; gathers are 4-byte.
__global__ void __launch_bounds__(kGpThreads)
    group_point_scalar_kernel(int n, int c, int rows, const float *__restrict__ points,
                              const int *__restrict__ idx, float *__restrict__ out) {
    const int scene = blockIdx.y;
    const float *pts = points + (size_t)scene * n * c;
    const int *ind = idx + (size_t)scene * rows;
    float *o = out + (size_t)scene * rows * c;
    const long long total = (long long)rows * c;
    for (long long e = (long long)blockIdx.x * kGpThreads + threadIdx.x; e < total;
         e += (long long)gridDim.x * kGpThreads) {
        int r = (int)(e / c);
        int l = (int)(e - (long long)r * c);
        o[e] = __ldg(&pts[(size_t)ind[r] * c + l]);
    }
}

// Rows of a few floats (xyz = 3, intensity = 1): one THREAD per output row, CX + CP floats per row gathered from two
// sources that share the index (CP = 0: one source).  Consecutive threads write consecutive rows, so the stores of a
// warp are one contiguous run; the flat (element / c) mapping of the scalar kernel above costs a division per float.
template <int CX, int CP>
__global__ void __launch_bounds__(kGpThreads)
    group_point_small_kernel(int n, int rows, const float *__restrict__ a, const float *__restrict__ bsrc,
                             const int *__restrict__ idx, float *__restrict__ out_a, float *__restrict__ out_b) {
    const int scene = blockIdx.y;
    const int *ind = idx + (size_t)scene * rows;
    for (int r = blockIdx.x * kGpThreads + threadIdx.x; r < rows; r += gridDim.x * kGpThreads) {
        const size_t src = (size_t)scene * n + ind[r], dst = (size_t)scene * rows + r;
        float va[CX], vb[CP > 0 ? CP : 1];
#pragma unroll
        for (int j = 0; j < CX; ++j) va[j] = __ldg(&a[src * CX + j]);
#pragma unroll
        for (int j = 0; j < CP; ++j) vb[j] = __ldg(&bsrc[src * CP + j]);
#pragma unroll
        for (int j = 0; j < CX; ++j) out_a[dst * CX + j] = va[j];
#pragma unroll
        for (int j = 0; j < CP; ++j) out_b[dst * CP + j] = vb[j];
    }
}

// sample_and_group's tail in one pass (hf/core/feature_extractors/pointnet_util.py:49-60): out row r =
// [xyz[idx[r]] - new_xyz[r / nsample], points[idx[r]]], width 3 + c (c = 0: the translated xyz only).
// The reference runs two group_point ops, a tile + subtract and a concat: four passes over the big tensor.
// One warp per output row, four rows in flight; rows are 3 + c floats, so accesses are 32-bit, coalesced.
// XYZ_LAST: columns are [points[idx[r]], xyz[idx[r]] - new_xyz[r / nsample]] - the order pointnet_sa_module_msg
// concatenates in (pointnet_util.py:262-264), the opposite of sample_and_group's.
template <bool XYZ_LAST>
__global__ void __launch_bounds__(kGpThreads)
    group_concat_kernel(int n, int c, int rows, int nsample, const float *__restrict__ xyz, const float *__restrict__ points,
                        const int *__restrict__ idx, const float *__restrict__ new_xyz, float *__restrict__ out) {
    const int scene = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * kGpThreads + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * kGpThreads) >> 5;
    const int width = 3 + c;
    const float *x = xyz + (size_t)scene * n * 3;
    const float *pts = points ? points + (size_t)scene * n * c : nullptr;
    const int *ind = idx + (size_t)scene * rows;
    const float *ctr = new_xyz + (size_t)scene * (rows / nsample) * 3;
    float *o = out + (size_t)scene * rows * width;
    for (int r0 = warp * 4; r0 < rows; r0 += nwarps * 4) {
        int src[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) src[u] = (r0 + u < rows) ? __ldg(&ind[r0 + u]) : -1;
        for (int j = lane; j < width; j += 32) {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (src[u] < 0) continue;
                const int jx = XYZ_LAST ? j - c : j;       // column of the translated coordinates, if 0 <= jx < 3
                const int jp = XYZ_LAST ? j : j - 3;       // column of the features otherwise
                if (jx >= 0 && jx < 3) v[u] = __fsub_rn(__ldg(&x[(size_t)src[u] * 3 + jx]), __ldg(&ctr[(size_t)((r0 + u) / nsample) * 3 + jx]));
                else v[u] = __ldg(&pts[(size_t)src[u] * c + jp]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (src[u] >= 0) o[(size_t)(r0 + u) * width + j] = v[u];
        }
    }
}

}  // namespace

// Smallest float t with sqrtf(t) >= radius; d2 < t  <=>  sqrtf(d2) < radius for d2 >= 0.
// The reference additionally clamps sqrtf(d2) up to 1e-20f, which only matters for
// radius <= 1e-20f, where nothing can be accepted (returns 0).
float ball_threshold(float radius) {
    if (!(radius > 1e-20f)) return 0.f;
    double r2 = (double)radius * (double)radius;
    float t = (float)r2;
    if (isinf(t)) return t;
    while (t > 0.f && sqrtf(nextafterf(t, 0.f)) >= radius) t = nextafterf(t, 0.f);
    while (sqrtf(t) < radius) t = nextafterf(t, INFINITY);
    return t;
}

}  // namespace hf

namespace hf {
thread_local int g_gather_bulk = 0;  // hf_gather_set_mode: 1 = TMA row mover where it applies, 0 = register kernels (faster: profiles/r2_gather_bench_tma_rows.txt)
}
extern "C" void hf_gather_set_mode(int bulk) { hf::g_gather_bulk = bulk; }

extern "C" int hf_query_ball_point(int b, int n, int m, float radius, int nsample, const float *xyz1,
                                   const float *xyz2, int *idx, int *pts_cnt, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(radius > 0, "QueryBallPoint expects positive radius");
    HF_REQUIRE(nsample > 0, "QueryBallPoint expects positive nsample");
    HF_REQUIRE(b >= 0 && n >= 1 && m >= 0, "QueryBallPoint expects (batch_size, ndataset, 3) xyz1 shape.");
    HF_REQUIRE(xyz1 && xyz2 && idx && pts_cnt, "QueryBallPoint: null pointer");
    HF_REQUIRE(b <= 65535, "QueryBallPoint: batch_size > 65535 not supported");
    if (b == 0 || m == 0) return HF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    float thr2 = ball_threshold(radius);
    dim3 grid((m + kBqThreads - 1) / kBqThreads, b);
    size_t rowbytes = (size_t)kBqThreads * (nsample + 1) * sizeof(int);
    if (rowbytes <= 160 * 1024) {
        auto kern = query_ball_kernel<true>;
        if (rowbytes + 25 * 1024 > 48 * 1024)
            HF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rowbytes));
        kern<<<grid, kBqThreads, rowbytes, s>>>(n, m, thr2, nsample, xyz1, xyz2, idx, pts_cnt);
    } else {
        query_ball_kernel<false><<<grid, kBqThreads, 0, s>>>(n, m, thr2, nsample, xyz1, xyz2, idx, pts_cnt);
    }
    HF_LAUNCH_OK("query_ball_kernel");
    return HF_OK;
}

namespace hf {
namespace {
// group_point of `points` (b,n,c) and, when xyz != null, of the coordinates (b,n,3) with the same idx, in one launch
// where a kernel covers the pair (c % 4 == 0 with aligned rows, or c <= 4); otherwise one launch each.
int group_point_any(int b, int n, int c, int m, int nsample, const float *points, const int *idx, float *out,
                    const float *xyz, float *out_xyz, cudaStream_t s) {
    HF_REQUIRE(b >= 0 && n >= 1 && c >= 1 && m >= 0 && nsample >= 0,
               "GroupPoint expects (batch_size, num_points, channel) points shape");
    HF_REQUIRE(points && idx && out && (!xyz || out_xyz), "GroupPoint: null pointer");
    HF_REQUIRE(b <= 65535, "GroupPoint: batch_size > 65535 not supported");
    HF_REQUIRE((long long)m * nsample < (1ll << 31), "GroupPoint: npoint*nsample must fit int32");
    const int rows = m * nsample;
    if (b == 0 || rows == 0) return HF_OK;
    const int sms = num_sms();
    const bool vec = (c % 4 == 0) && ((reinterpret_cast<uintptr_t>(points) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    const int rowbytes = c * 4;
    if (vec && g_gather_bulk && rowbytes <= kGbStageBytes && !xyz) {
        int R = kGbStageBytes / rowbytes;
        if (R > 32) R = 32;
        const long long rows_all = (long long)b * rows;
        const long long groups = (rows_all + R - 1) / R;
        const size_t smem = (size_t)(kGbThreads / 32) * kGbStages * kGbStageBytes;
        HF_CUDA_OK(cudaFuncSetAttribute(group_point_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        long long blocks = (groups + (kGbThreads / 32) - 1) / (kGbThreads / 32);
        if (blocks > sms) blocks = sms;
        group_point_bulk_kernel<<<(unsigned)blocks, kGbThreads, smem, s>>>(
            n, rowbytes, R, rows_all, rows, reinterpret_cast<const unsigned char *>(points), idx,
            reinterpret_cast<unsigned char *>(out));
        HF_LAUNCH_OK("group_point_bulk_kernel");
        return HF_OK;
    }
    if (vec) {
        int warps_needed = (rows + kGpRowsPerWarp - 1) / kGpRowsPerWarp;
        int blocks = (warps_needed + (kGpThreads / 32) - 1) / (kGpThreads / 32);
        int cap = (sms * 8 + b - 1) / b;  // ~8 resident CTAs per SM over the whole grid
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        if (xyz)
            group_point_vec4_kernel<true><<<dim3(blocks, b), kGpThreads, 0, s>>>(
                n, c / 4, rows, reinterpret_cast<const float4 *>(points), idx, reinterpret_cast<float4 *>(out), xyz, out_xyz);
        else
            group_point_vec4_kernel<false><<<dim3(blocks, b), kGpThreads, 0, s>>>(
                n, c / 4, rows, reinterpret_cast<const float4 *>(points), idx, reinterpret_cast<float4 *>(out), nullptr, nullptr);
        HF_LAUNCH_OK("group_point_vec4_kernel");
        return HF_OK;
    }
    if (c <= 4) {  // thread per row; with xyz: (3 + c) floats per row from the two sources
        int blocks = (rows + kGpThreads - 1) / kGpThreads;
        const int cap = (sms * 8 + b - 1) / b;
        if (blocks > cap) blocks = cap;
        const dim3 grid(blocks, b);
#define HF_GP_SMALL(CX, CP, A, B, OA, OB) group_point_small_kernel<CX, CP><<<grid, kGpThreads, 0, s>>>(n, rows, A, B, idx, OA, OB)
        if (xyz) {
            switch (c) {
                case 1: HF_GP_SMALL(3, 1, xyz, points, out_xyz, out); break;
                case 2: HF_GP_SMALL(3, 2, xyz, points, out_xyz, out); break;
                case 3: HF_GP_SMALL(3, 3, xyz, points, out_xyz, out); break;
                default: HF_GP_SMALL(3, 4, xyz, points, out_xyz, out); break;
            }
        } else {
            switch (c) {
                case 1: HF_GP_SMALL(1, 0, points, nullptr, out, nullptr); break;
                case 2: HF_GP_SMALL(2, 0, points, nullptr, out, nullptr); break;
                case 3: HF_GP_SMALL(3, 0, points, nullptr, out, nullptr); break;
                default: HF_GP_SMALL(4, 0, points, nullptr, out, nullptr); break;
            }
        }
#undef HF_GP_SMALL
        HF_LAUNCH_OK("group_point_small_kernel");
        return HF_OK;
    }
    {
        long long total = (long long)rows * c;
        long long blocks = (total + kGpThreads - 1) / kGpThreads;
        long long cap = ((long long)sms * 8 + b - 1) / b;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        group_point_scalar_kernel<<<dim3((unsigned)blocks, b), kGpThreads, 0, s>>>(n, c, rows, points, idx, out);
        HF_LAUNCH_OK("group_point_scalar_kernel");
    }
    if (xyz) return group_point_any(b, n, 3, m, nsample, xyz, idx, out_xyz, nullptr, nullptr, s);
    return HF_OK;
}
}  // namespace
}  // namespace hf

extern "C" int hf_group_point(int b, int n, int c, int m, int nsample, const float *points, const int *idx,
                              float *out, hf_stream_t stream) {
    return hf::group_point_any(b, n, c, m, nsample, points, idx, out, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int hf_group_point_pair(int b, int n, int c, int m, int nsample, const float *xyz, const float *points,
                                   const int *idx, float *out_xyz, float *out_points, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(xyz && out_xyz, "GroupPoint: null pointer");
    return group_point_any(b, n, c, m, nsample, points, idx, out_points, xyz, out_xyz, (cudaStream_t)stream);
}

namespace hf {
namespace {
int group_concat_any(bool xyz_last, int b, int n, int c, int m, int nsample, const float *xyz, const float *points,
                     const int *idx, const float *new_xyz, float *out, hf_stream_t stream);
}
}  // namespace hf

extern "C" int hf_group_point_concat(int b, int n, int c, int m, int nsample, const float *xyz, const float *points,
                                     const int *idx, const float *new_xyz, float *out, hf_stream_t stream) {
    return hf::group_concat_any(false, b, n, c, m, nsample, xyz, points, idx, new_xyz, out, stream);
}

extern "C" int hf_group_point_concat_msg(int b, int n, int c, int m, int nsample, const float *xyz, const float *points,
                                         const int *idx, const float *new_xyz, float *out, hf_stream_t stream) {
    return hf::group_concat_any(true, b, n, c, m, nsample, xyz, points, idx, new_xyz, out, stream);
}

namespace hf {
namespace {
int group_concat_any(bool xyz_last, int b, int n, int c, int m, int nsample, const float *xyz, const float *points,
                     const int *idx, const float *new_xyz, float *out, hf_stream_t stream) {
    HF_REQUIRE(b >= 0 && n >= 1 && c >= 0 && m >= 0 && nsample >= 1,
               "GroupPoint expects (batch_size, num_points, channel) points shape");
    HF_REQUIRE(b <= 65535, "GroupPoint: batch_size > 65535 not supported");
    HF_REQUIRE((long long)m * nsample < (1ll << 31), "GroupPoint: npoint*nsample must fit int32");
    const int rows = m * nsample;
    if (b == 0 || rows == 0) return HF_OK;
    HF_REQUIRE(xyz && idx && new_xyz && out && (c == 0 || points), "GroupPoint: null pointer");
    int blocks = (rows + 4 * (kGpThreads / 32) - 1) / (4 * (kGpThreads / 32));
    const int cap = (num_sms() * 8 + b - 1) / b;
    if (blocks > cap) blocks = cap;
    const dim3 grid(blocks, b);
    if (xyz_last && c)
        group_concat_kernel<true><<<grid, kGpThreads, 0, (cudaStream_t)stream>>>(n, c, rows, nsample, xyz, points, idx, new_xyz, out);
    else
        group_concat_kernel<false><<<grid, kGpThreads, 0, (cudaStream_t)stream>>>(n, c, rows, nsample, xyz, c ? points : nullptr,
                                                                                  idx, new_xyz, out);
    HF_LAUNCH_OK("group_concat_kernel");
    return HF_OK;
}
}  // namespace
}  // namespace hf
