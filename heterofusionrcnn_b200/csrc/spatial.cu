// spatial.cu — query_ball_point and three_nn with exact spatial pruning, for sm_100a.
//
// Same results as the every-pair kernels in grouping.cu / interpolate.cu (and therefore as
// grouping/tf_grouping_g.cu:3-36 and interpolate/tf_interpolate_g.cu:22-65 of the reference),
// a different algorithm.  The reference tests every (query, candidate) pair: 16 x 4096 x 16384
// = 1.07e9 distance evaluations per op at BASELINE config 2, of which a few thousand per query
// can possibly matter.  Here both clouds of a call are first sorted along a Hilbert curve and
// cut into BUCKETS of 32 consecutive points with their bounding boxes (cloud_index_kernel, one
// CTA per scene and cloud, radix sort in shared memory).  A warp then owns one bucket of 32
// queries and visits a candidate bucket only if
//      LB = sqdist(per-axis gap between the two boxes)  <  threshold
// (ball query: the radius threshold; three_nn: the worst third-best distance of its 32 lanes).
// Every IEEE operation involved is monotone and LB uses the SAME operation sequence as the
// point distances (common.cuh:sqdist), so LB <= fl(d(q, c)) holds for every query q of the
// warp and candidate c of the bucket: a skipped bucket cannot contain a hit, no slack needed.
//
// Order independence.  Buckets are visited in curve order, not index order, so the reference's
// "first in index order wins" rules are restated on explicit keys:
//   * ball query: all hits are collected, the nsample SMALLEST indices are kept and written in
//     ascending order (= the first nsample hits of an ascending scan, tf_grouping_g.cu:15-33);
//   * three_nn: candidates are ranked by (distance, index) lexicographically (= strict '<'
//     insertion during an ascending scan, tf_interpolate_g.cu:50-61).
//
// The caller provides the workspace (hf_spatial_workspace_bytes); clouds beyond 16384 points
// or calls too small to pay for the sort take the every-pair kernels.
#include <limits.h>
#include <math.h>

#include "common.cuh"

namespace hf {
float ball_threshold(float radius);  // grouping.cu
int three_nn_every_pair(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                        float *weight, cudaStream_t s);  // interpolate.cu

namespace {

constexpr int kIdxThreads = 1024;
constexpr int kIdxMaxE = 16;                        // sort elements per thread
constexpr int kMaxIndexed = kIdxMaxE * kIdxThreads;  // 16384 points fit the shared-memory sort of one CTA
// Larger clouds are indexed in PARTS of <= 16384 consecutive points, each sorted by its own CTA: the query
// kernels only need buckets of 32 nearby points with a box, not one curve through the whole cloud.
constexpr int kMaxCloud = 4 * kMaxIndexed;
constexpr int kHistPitch = 257;                     // hist[warp][digit], padded: conflict free by digit and by warp
constexpr int kHistWords = 32 * kHistPitch;
constexpr size_t kIdxSmem = (size_t)kMaxIndexed * 4 + (size_t)kHistWords * 4 + (size_t)kMaxIndexed * 2;  // 128 KB

__device__ __forceinline__ int sp_f2ord(float f) {
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float sp_ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }
// Position of cell (x, y, z) (8 bits each) along a 3-D Hilbert curve.  Unlike a Morton curve it
// has no jumps: 32 consecutive points always form one connected, compact run, so no bucket box
// straddles half the scene.  The curve is the one Skilling's transpose algorithm produces, run
// here as its 24-state machine: entry [state*8 + octant] = next state << 3 | curve digit, octant
// = the (x, y, z) bits of one level, most significant level first.  (Table derived by
// enumerating Skilling's curve; any locality-preserving order would be equally exact.)
__constant__ unsigned char kHilbertLut[192] = {
    8, 17, 27, 2, 39, 46, 52, 5, 56, 71, 73, 86, 91, 20, 10, 13, 48, 1, 103, 110, 115, 18, 12, 21,
    126, 129, 29, 26, 79, 80, 140, 3, 148, 43, 37, 34, 127, 128, 78, 81, 156, 45, 35, 42, 31, 6, 160, 105,
    72, 87, 139, 4, 57, 70, 50, 53, 0, 171, 111, 76, 49, 58, 102, 61, 180, 143, 83, 184, 69, 54, 66, 97,
    16, 123, 9, 74, 47, 60, 38, 77, 132, 95, 85, 14, 67, 144, 82, 33, 142, 55, 185, 96, 93, 116, 90, 11,
    188, 107, 175, 176, 101, 98, 62, 65, 164, 109, 119, 22, 99, 106, 152, 41, 174, 177, 63, 64, 117, 114, 92, 19,
    30, 125, 161, 122, 7, 172, 104, 75, 130, 25, 133, 166, 179, 136, 84, 191, 94, 15, 141, 28, 145, 32, 138, 51,
    146, 155, 149, 36, 137, 24, 190, 167, 154, 157, 147, 44, 169, 182, 120, 135, 162, 165, 121, 134, 187, 108, 168, 183,
    118, 173, 23, 124, 153, 170, 40, 59, 178, 113, 131, 88, 181, 158, 68, 151, 186, 163, 89, 112, 189, 100, 150, 159,
};
__device__ __forceinline__ uint32_t sp_hilbert24(const unsigned char *lut, uint32_t x, uint32_t y, uint32_t z) {
    uint32_t state = 0u, h = 0u;
#pragma unroll
    for (int l = 7; l >= 0; --l) {
        const uint32_t oct = (((x >> l) & 1u) << 2) | (((y >> l) & 1u) << 1) | ((z >> l) & 1u);
        const uint32_t e = lut[state * 8u + oct];
        h = (h << 3) | (e & 7u);
        state = e >> 3;
    }
    return h;
}

struct CloudDesc {
    const float *xyz;  // (b, n, 3)
    float4 *sorted;    // (b, npad) {x, y, z, original index as int bits}; padding = {NaN, NaN, NaN, -1}
    float4 *boxes;     // (b, npad/32, 2) {lo.xyz, smallest original index in the bucket}, {hi.xyz, -}
    float4 *super;     // (b, nsuper, 2) boxes of 32 consecutive buckets, clouds above kMaxIndexed points only
    int nsuper;
    int n, npad, per_thread;  // points, rounded up to 32, sort elements per thread = ceil(part_n / 1024)
    int parts, part_n;        // the cloud is sorted in `parts` runs of part_n (a multiple of 32) consecutive points
};

// grid = (b, 2): blockIdx.y selects the cloud.  dynamic smem = kIdxSmem.
__global__ void __launch_bounds__(kIdxThreads, 1) cloud_index_kernel(CloudDesc ca, CloudDesc cb, int gparts) {
    extern __shared__ __align__(16) unsigned char idx_smem[];
    uint32_t *code = reinterpret_cast<uint32_t *>(idx_smem);               // [np] Hilbert position
    uint32_t *hist = code + kIdxMaxE * kIdxThreads;                         // [32 warps][257]
    unsigned short *sidx = reinterpret_cast<unsigned short *>(hist + kHistWords);  // [np] original index
    __shared__ int red[6][32];
    __shared__ uint32_t wsum[32];
    __shared__ unsigned char hlut[192];
    const CloudDesc cd = blockIdx.y == 0 ? ca : cb;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
#ifdef HF_SPATIAL_DEBUG
    long long tk[12]; int ntk = 0;
#define TICK() do { __syncthreads(); tk[ntk++] = clock64(); } while (0)
#else
#define TICK() do {} while (0)
#endif
    TICK();
    const int scene = blockIdx.x / gparts, part = blockIdx.x % gparts;
    if (part >= cd.parts) return;
    const int base = part * cd.part_n;
    const int n = min(cd.part_n, cd.n - base);  // points of this part
    if (n <= 0) return;
    const float *data = cd.xyz + ((size_t)scene * cd.n + base) * 3;
    if (t < 192) hlut[t] = kHilbertLut[t];  // read after the barriers of the bounding box pass

    // ---- bounding box -> cubic quantisation ----
    int mn0 = INT_MAX, mn1 = INT_MAX, mn2 = INT_MAX, mx0 = INT_MIN, mx1 = INT_MIN, mx2 = INT_MIN;
    for (int i = t; i < n; i += kIdxThreads) {
        const int o0 = sp_f2ord(data[(size_t)i * 3]), o1 = sp_f2ord(data[(size_t)i * 3 + 1]),
                  o2 = sp_f2ord(data[(size_t)i * 3 + 2]);
        mn0 = min(mn0, o0); mx0 = max(mx0, o0);
        mn1 = min(mn1, o1); mx1 = max(mx1, o1);
        mn2 = min(mn2, o2); mx2 = max(mx2, o2);
    }
    mn0 = __reduce_min_sync(0xffffffffu, mn0); mx0 = __reduce_max_sync(0xffffffffu, mx0);
    mn1 = __reduce_min_sync(0xffffffffu, mn1); mx1 = __reduce_max_sync(0xffffffffu, mx1);
    mn2 = __reduce_min_sync(0xffffffffu, mn2); mx2 = __reduce_max_sync(0xffffffffu, mx2);
    if (lane == 0) {
        red[0][warp] = mn0; red[1][warp] = mn1; red[2][warp] = mn2;
        red[3][warp] = mx0; red[4][warp] = mx1; red[5][warp] = mx2;
    }
    __syncthreads();
    int l0 = red[0][lane], l1 = red[1][lane], l2 = red[2][lane], h0 = red[3][lane], h1 = red[4][lane], h2 = red[5][lane];
    l0 = __reduce_min_sync(0xffffffffu, l0); l1 = __reduce_min_sync(0xffffffffu, l1); l2 = __reduce_min_sync(0xffffffffu, l2);
    h0 = __reduce_max_sync(0xffffffffu, h0); h1 = __reduce_max_sync(0xffffffffu, h1); h2 = __reduce_max_sync(0xffffffffu, h2);
    const float lo0 = sp_ord2f(l0), lo1 = sp_ord2f(l1), lo2 = sp_ord2f(l2);
    const float ext = fmaxf(fmaxf(sp_ord2f(h0) - lo0, sp_ord2f(h1) - lo1), sp_ord2f(h2) - lo2);
    const float inv = (ext > 0.f && ext < 1e30f) ? 255.f / ext : 0.f;  // 256 cubic cells per axis

    TICK();
    // ---- (Hilbert position, index) pairs, LSD radix sort in shared memory: 3 passes of 8 bits ----
    // A warp owns a contiguous chunk of E*32 elements; its per-digit counters are private
    // (hist[warp][digit]), ranks inside a 32-element row come from ballots, so every pass is
    // stable without atomics.  Each thread keeps its E elements in registers across the scatter,
    // so the pass is in place.
    const int E = cd.per_thread;  // elements per thread, <= kIdxMaxE
    const int np = E * kIdxThreads;
#pragma unroll 4
    for (int i = t; i < np; i += kIdxThreads) {
        uint32_t c = 0x00FFFFFFu;  // padding sorts last
        if (i < n) {
            const uint32_t qx = min(__float2uint_rz((data[(size_t)i * 3] - lo0) * inv), 255u);
            const uint32_t qy = min(__float2uint_rz((data[(size_t)i * 3 + 1] - lo1) * inv), 255u);
            const uint32_t qz = min(__float2uint_rz((data[(size_t)i * 3 + 2] - lo2) * inv), 255u);
            c = sp_hilbert24(hlut, qx, qy, qz);
        }
        code[i] = c;
        sidx[i] = (unsigned short)i;
    }
    const unsigned lt_mask = (1u << lane) - 1u;
    TICK();
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = 8 * pass;
        for (int i = t; i < kHistWords; i += kIdxThreads) hist[i] = 0u;
        __syncthreads();
        uint32_t myc[kIdxMaxE], off[kIdxMaxE];
        unsigned short myi[kIdxMaxE];
        // (a) independent: digit peers of every element of this thread (off[] holds the peer mask for now)
#pragma unroll
        for (int it = 0; it < kIdxMaxE; ++it) {
            if (it < E) {  // uniform
                const int e = (warp * E + it) * 32 + lane;
                myc[it] = code[e];
                myi[it] = sidx[e];
                const uint32_t d = (myc[it] >> shift) & 255u;
                // lanes of this row with the same digit (8 ballots; match.any costs one round per distinct value)
                unsigned peers = 0xffffffffu;
#pragma unroll
                for (int bit = 0; bit < 8; ++bit) {
                    const unsigned bal = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
                    peers &= ((d >> bit) & 1u) ? bal : ~bal;
                }
                off[it] = peers;
            }
        }
        // (b) the warp's counters, row by row in order: offset = elements of this digit seen so far
#pragma unroll
        for (int it = 0; it < kIdxMaxE; ++it) {
            if (it < E) {
                const uint32_t d = (myc[it] >> shift) & 255u;
                const unsigned peers = off[it];
                const uint32_t base = hist[warp * kHistPitch + d];
                __syncwarp();
                if ((peers & lt_mask) == 0u) hist[warp * kHistPitch + d] = base + __popc(peers);
                __syncwarp();
                off[it] = base + __popc(peers & lt_mask);
            }
        }
        __syncthreads();
        {  // exclusive scan in (digit, warp) order: thread t owns digit t/4, warps (t%4)*8 .. +7
            const int sd = t >> 2, sw0 = (t & 3) * 8;
            uint32_t v[8], sum = 0u;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                v[k] = hist[(sw0 + k) * kHistPitch + sd];
                sum += v[k];
            }
            uint32_t inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += up;
            }
            if (lane == 31) wsum[warp] = inc;
            __syncthreads();
            uint32_t wtot = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, wtot, o);
                if (lane >= o) wtot += up;
            }
            uint32_t run = inc - sum + (warp > 0 ? __shfl_sync(0xffffffffu, wtot, warp - 1) : 0u);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                hist[(sw0 + k) * kHistPitch + sd] = run;
                run += v[k];
            }
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < kIdxMaxE; ++it) {
            if (it < E) {
                const uint32_t d = (myc[it] >> shift) & 255u;
                const uint32_t dst = hist[warp * kHistPitch + d] + off[it];
                code[dst] = myc[it];
                sidx[dst] = myi[it];
            }
        }
        __syncthreads();
        TICK();
    }

    // ---- sorted copy + bucket boxes (a warp writes one bucket per step) ----
    float4 *sorted = cd.sorted + (size_t)scene * cd.npad + base;
    float4 *boxes = cd.boxes + ((size_t)scene * (cd.npad / 32) + base / 32) * 2;
    const int npad = (n + 31) & ~31;
    const float qnan = __int_as_float(0x7FC00000);
    for (int i0 = t; i0 < npad; i0 += 4 * kIdxThreads) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {  // the 12 gathers of a step are in flight together
            const int i = i0 + u * kIdxThreads;
            v[u] = make_float4(qnan, qnan, qnan, __int_as_float(-1));
            if (i < n) {
                const int src = (int)sidx[i];
                v[u].x = __ldg(&data[(size_t)src * 3]);
                v[u].y = __ldg(&data[(size_t)src * 3 + 1]);
                v[u].z = __ldg(&data[(size_t)src * 3 + 2]);
                v[u].w = __int_as_float(src + base);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * kIdxThreads;
            if (i >= npad) break;  // warp uniform: npad is a multiple of 32
            const bool real = i < n;
            sorted[i] = v[u];
            // bucket box + smallest original index: 7 redux.sync on order-preserving integer images
            const int ox = sp_f2ord(v[u].x), oy = sp_f2ord(v[u].y), oz = sp_f2ord(v[u].z);
            const float bx0 = sp_ord2f(__reduce_min_sync(0xffffffffu, real ? ox : INT_MAX));
            const float by0 = sp_ord2f(__reduce_min_sync(0xffffffffu, real ? oy : INT_MAX));
            const float bz0 = sp_ord2f(__reduce_min_sync(0xffffffffu, real ? oz : INT_MAX));
            const float bx1 = sp_ord2f(__reduce_max_sync(0xffffffffu, real ? ox : INT_MIN));
            const float by1 = sp_ord2f(__reduce_max_sync(0xffffffffu, real ? oy : INT_MIN));
            const float bz1 = sp_ord2f(__reduce_max_sync(0xffffffffu, real ? oz : INT_MIN));
            const int imin = __reduce_min_sync(0xffffffffu, real ? __float_as_int(v[u].w) : INT_MAX);
            if (lane == 0) {
                boxes[(i >> 5) * 2] = make_float4(bx0, by0, bz0, __int_as_float(imin));
                boxes[(i >> 5) * 2 + 1] = make_float4(bx1, by1, bz1, 0.f);
            }
        }
    }
    TICK();
#ifdef HF_SPATIAL_DEBUG
    if (t == 0 && blockIdx.x == 0)
        printf("cloud %d n=%d: bbox %lld keys %lld pass %lld %lld %lld final %lld\n", blockIdx.y, n, tk[1] - tk[0], tk[2] - tk[1],
               tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5]);
#endif
}

// Lower bound of sqdist over all (query in box q, candidate in box c) pairs; same op sequence as sqdist.
__device__ __forceinline__ float box_gap2(const float4 &qlo, const float4 &qhi, const float4 &clo, const float4 &chi) {
    const float gx = fmaxf(0.f, fmaxf(clo.x - qhi.x, qlo.x - chi.x));
    const float gy = fmaxf(0.f, fmaxf(clo.y - qhi.y, qlo.y - chi.y));
    const float gz = fmaxf(0.f, fmaxf(clo.z - qhi.z, qlo.z - chi.z));
    return sqdist(gx, gy, gz);
}

constexpr int kSqWarps = 4;  // warps (= query buckets) per CTA of the two query kernels

// Stage candidate bucket `cb` (32 x float4) into this warp's shared-memory tile.
__device__ __forceinline__ void stage_bucket(float4 *tile, const float4 *sortedC, int cb, int lane) {
    const float4 v = __ldg(&sortedC[(size_t)cb * 32 + lane]);
    __syncwarp();  // the previous bucket's readers are done
    tile[lane] = v;
    __syncwarp();
}

// Lower bound of sqdist between one query point and every candidate in a box (same op sequence).
__device__ __forceinline__ float point_gap2(const float4 &q, const float4 &clo, const float4 &chi) {
    const float gx = fmaxf(0.f, fmaxf(clo.x - q.x, q.x - chi.x));
    const float gy = fmaxf(0.f, fmaxf(clo.y - q.y, q.y - chi.y));
    const float gz = fmaxf(0.f, fmaxf(clo.z - q.z, q.z - chi.z));
    return sqdist(gx, gy, gz);
}

// ---- ball query: CTA = 32 curve-consecutive queries (lane = one query), 4 warps ---------
// The 4 warps split the candidate buckets (groups of 32 buckets round robin), each collecting
// its own hit rows; the rows are merged at the end.  Two-level pruning: (1) lane l tests
// candidate bucket g*32+l against the CTA's query box, one ballot per 32 buckets; (2) every
// surviving bucket is tested against each lane's own query point and visited only if some lane
// can reach it.  Hit rows live in shared memory, transposed: slot s of lane l at
// rows[s * 33 + l] (conflict free for the owner and for the coalesced write-out).
// grid = (query buckets, b).  dynamic smem = 5 row buffers of nsample*33 ints + 4 tiles.
__global__ void __launch_bounds__(kSqWarps * 32)
    ball_query_indexed_kernel(int m, int nsample, float thr2, int nbq, int nbc, const float4 *__restrict__ sortedQ,
                              const float4 *__restrict__ boxesQ, const float4 *__restrict__ sortedC,
                              const float4 *__restrict__ boxesC, int *__restrict__ idx, int *__restrict__ pts_cnt,
                              unsigned long long *__restrict__ pair_counter) {
    extern __shared__ __align__(16) unsigned char sq_smem[];
    __shared__ int cnts[kSqWarps + 1][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int scene = blockIdx.y;
    const int qb = blockIdx.x;
    float4 *tile = reinterpret_cast<float4 *>(sq_smem) + warp * 32;
    int *rows_all = reinterpret_cast<int *>(sq_smem + kSqWarps * 32 * sizeof(float4));
    int *rows = rows_all + (size_t)warp * nsample * 33;
    int *merged = rows_all + (size_t)kSqWarps * nsample * 33;
    sortedQ += (size_t)scene * nbq * 32;
    boxesQ += (size_t)scene * nbq * 2;
    sortedC += (size_t)scene * nbc * 32;
    boxesC += (size_t)scene * nbc * 2;

    const float4 q = sortedQ[(size_t)qb * 32 + lane];
    const int qi = __float_as_int(q.w);  // original query index, -1 = padding lane
    const float4 qlo = boxesQ[qb * 2], qhi = boxesQ[qb * 2 + 1];
    int cnt = 0;            // hits stored by this warp for this lane (<= nsample)
    int mxv = -1, mxp = 0;  // once full: largest stored index and its slot

    // level 1 for this warp's groups up front (independent loads); lane j keeps the ballot of its j-th group
    const int ngroups = (nbc + 31) >> 5;                            // <= 16: nbc <= 512
    const int mygroups = (ngroups - warp + kSqWarps - 1) / kSqWarps;  // groups warp, warp+4, ...
    unsigned mymask = 0u;
#pragma unroll 4
    for (int j = 0; j < mygroups; ++j) {
        const int mine = (j * kSqWarps + warp) * 32 + lane;
        bool near = false;
        if (mine < nbc) near = box_gap2(qlo, qhi, __ldg(&boxesC[mine * 2]), __ldg(&boxesC[mine * 2 + 1])) < thr2;
        const unsigned bal = __ballot_sync(0xffffffffu, near);
        if (lane == j) mymask = bal;
    }
    int j = -1;
    unsigned todo = 0u;
    // next bucket some lane can reach (level 2), or -1
    auto next = [&]() -> int {
        for (;;) {
            while (!todo) {
                if (++j >= mygroups) return -1;
                todo = __shfl_sync(0xffffffffu, mymask, j);
            }
            const int cb = (j * kSqWarps + warp) * 32 + __ffs(todo) - 1;
            todo &= todo - 1;
            const float4 clo = __ldg(&boxesC[cb * 2]), chi = __ldg(&boxesC[cb * 2 + 1]);
            // a lane whose row is full only cares about buckets holding an index below its largest kept one
            const bool want = cnt < nsample || __float_as_int(clo.w) < mxv;
            if (__any_sync(0xffffffffu, qi >= 0 && want && point_gap2(q, clo, chi) < thr2)) return cb;
        }
    };
    int cur = next();
    float4 vcur = make_float4(0.f, 0.f, 0.f, 0.f);
    if (cur >= 0) vcur = __ldg(&sortedC[(size_t)cur * 32 + lane]);
    unsigned visited = 0;  // buckets of 32 candidates this warp tested against its 32 queries
    while (cur >= 0) {
        ++visited;
        const int nxt = next();  // its candidates are in flight while the current bucket is tested
        float4 vnxt = vcur;
        if (nxt >= 0) vnxt = __ldg(&sortedC[(size_t)nxt * 32 + lane]);
        __syncwarp();
        tile[lane] = vcur;
        __syncwarp();
#pragma unroll 8
        for (int c32 = 0; c32 < 32; ++c32) {
            const float4 c = tile[c32];
            // tf_grouping_g.cu:19-24 (x2 = query, x1 = dataset point); NaN padding never passes
            if (sqdist(q.x - c.x, q.y - c.y, q.z - c.z) < thr2) {
                const int k = __float_as_int(c.w);
                if (cnt < nsample) {
                    rows[cnt * 33 + lane] = k;
                    if (k > mxv) { mxv = k; mxp = cnt; }
                    ++cnt;
                } else if (k < mxv) {  // keep the nsample smallest indices
                    rows[mxp * 33 + lane] = k;
                    mxv = -1;
                    for (int s2 = 0; s2 < nsample; ++s2) {
                        const int v = rows[s2 * 33 + lane];
                        if (v > mxv) { mxv = v; mxp = s2; }
                    }
                }
            }
        }
        cur = nxt;
        vcur = vnxt;
    }
    if (pair_counter && lane == 0 && visited) atomicAdd(pair_counter, (unsigned long long)visited * 1024ull);
    // ascending index order (insertion sort of this lane's <= nsample hits)
    for (int a = 1; a < cnt; ++a) {
        const int v = rows[a * 33 + lane];
        int p = a - 1;
        while (p >= 0 && rows[p * 33 + lane] > v) {
            rows[(p + 1) * 33 + lane] = rows[p * 33 + lane];
            --p;
        }
        rows[(p + 1) * 33 + lane] = v;
    }
    cnts[warp][lane] = cnt;
    __syncthreads();
    if (warp == 0) {  // merge the 4 sorted rows of this lane: the nsample smallest of their union
        int p[kSqWarps], c[kSqWarps], head[kSqWarps];
        int total = 0;
#pragma unroll
        for (int w = 0; w < kSqWarps; ++w) {
            p[w] = 0;
            c[w] = cnts[w][lane];
            total += c[w];
            head[w] = c[w] > 0 ? rows_all[(size_t)w * nsample * 33 + lane] : INT_MAX;
        }
        total = min(total, nsample);
        for (int o = 0; o < total; ++o) {
            int best = 0;
#pragma unroll
            for (int w = 1; w < kSqWarps; ++w)
                if (head[w] < head[best]) best = w;
            merged[o * 33 + lane] = head[best];
#pragma unroll
            for (int w = 0; w < kSqWarps; ++w)
                if (w == best) {
                    ++p[w];
                    head[w] = p[w] < c[w] ? rows_all[((size_t)w * nsample + p[w]) * 33 + lane] : INT_MAX;
                }
        }
        cnts[kSqWarps][lane] = total;
        if (qi >= 0) pts_cnt[(size_t)scene * m + qi] = total;
    }
    __syncthreads();
    // write-out: warp w writes rows w*8 .. w*8+7, lanes = slots; unused slots repeat the first hit
    // (:26-29), a row without a hit is written as 0
    for (int r = warp * 8; r < warp * 8 + 8; ++r) {
        const int rq = __shfl_sync(0xffffffffu, qi, r);
        const int rc = cnts[kSqWarps][r];
        if (rq < 0) continue;
        int *dst = idx + ((size_t)scene * m + rq) * nsample;
        const int first = rc > 0 ? merged[r] : 0;
        for (int s2 = lane; s2 < nsample; s2 += 32) dst[s2] = s2 < rc ? merged[s2 * 33 + r] : first;
    }
}

// ---- three_nn: warp = 32 curve-consecutive unknown points; lane = one of them ----------
__global__ void __launch_bounds__(kSqWarps * 32)
    three_nn_indexed_kernel(int n, int nbq, int nbc, const float4 *__restrict__ sortedQ,
                            const float4 *__restrict__ boxesQ, const float4 *__restrict__ sortedC,
                            const float4 *__restrict__ boxesC, float *__restrict__ dist2, int *__restrict__ idx,
                            float *__restrict__ weight, unsigned long long *__restrict__ pair_counter) {
    __shared__ __align__(16) float4 tiles[kSqWarps * 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int scene = blockIdx.y;
    const int qb = blockIdx.x * kSqWarps + warp;
    if (qb >= nbq) return;
    float4 *tile = tiles + warp * 32;
    sortedQ += (size_t)scene * nbq * 32;
    boxesQ += (size_t)scene * nbq * 2;
    sortedC += (size_t)scene * nbc * 32;
    boxesC += (size_t)scene * nbc * 2;

    const float4 q = sortedQ[(size_t)qb * 32 + lane];
    const int qi = __float_as_int(q.w);
    const float4 qlo = boxesQ[qb * 2], qhi = boxesQ[qb * 2 + 1];
    // tf_interpolate_g.cu:36-38: best = 1e40 (+inf once stored as float), besti = 0
    float b1 = INFINITY, b2 = INFINITY, b3 = INFINITY;
    int i1 = 0, i2 = 0, i3 = 0;

    unsigned visited = 0;  // buckets of 32 candidates this warp tested against its 32 queries
    auto visit = [&](int cb) {
        ++visited;
        stage_bucket(tile, sortedC, cb, lane);
#pragma unroll 8
        for (int j = 0; j < 32; ++j) {
            const float4 c = tile[j];
            const float d = sqdist(q.x - c.x, q.y - c.y, q.z - c.z);  // :43-49 (unknown - known)
            const int k = __float_as_int(c.w);
            // strict '<' insertion of an ascending scan == rank by (d, k); NaN padding never passes
            if (d <= b3 && (d < b3 || k < i3)) {  // one compare rejects almost every candidate
                if (d < b1 || (d == b1 && k < i1)) {
                    b3 = b2; i3 = i2; b2 = b1; i2 = i1; b1 = d; i1 = k;
                } else if (d < b2 || (d == b2 && k < i2)) {
                    b3 = b2; i3 = i2; b2 = d; i2 = k;
                } else {
                    b3 = d; i3 = k;
                }
            }
        }
    };
    // worst third-best distance of the warp's real lanes, as ordered bits (distances are >= 0)
    auto threshold = [&]() { return __reduce_max_sync(0xffffffffu, qi >= 0 ? __float_as_uint(b3) : 0u); };

    // pass A: the candidate bucket nearest to this warp's box seeds the bests (and warms L1 with the boxes)
    const int ngroups = (nbc + 31) >> 5;
    float lmin = INFINITY;
    int lwhere = 0;
#pragma unroll 4
    for (int g = 0; g < ngroups; ++g) {
        const int mine = g * 32 + lane;
        float lb = INFINITY;
        if (mine < nbc) lb = box_gap2(qlo, qhi, __ldg(&boxesC[mine * 2]), __ldg(&boxesC[mine * 2 + 1]));
        if (lb < lmin) { lmin = lb; lwhere = mine; }
    }
    const uint32_t gmin = __reduce_min_sync(0xffffffffu, __float_as_uint(lmin));
    const unsigned who = __ballot_sync(0xffffffffu, __float_as_uint(lmin) == gmin);
    const int seed = __shfl_sync(0xffffffffu, lwhere, __ffs(who) - 1);
    visit(seed);
    uint32_t thr = threshold();
    // pass B: level 1 (bucket box vs warp box, against the warp's threshold), then level 2 (bucket box vs
    // each lane's point, against that lane's own third best).  '<=': equal distances can still win on index.
    for (int g = 0; g < ngroups; ++g) {
        const int mine = g * 32 + lane;
        float lb = INFINITY;
        if (mine < nbc && mine != seed) lb = box_gap2(qlo, qhi, __ldg(&boxesC[mine * 2]), __ldg(&boxesC[mine * 2 + 1]));
        unsigned todo = __ballot_sync(0xffffffffu, mine < nbc && mine != seed && __float_as_uint(lb) <= thr);
        while (todo) {
            const int l = __ffs(todo) - 1;
            todo &= todo - 1;
            const uint32_t bound = __float_as_uint(__shfl_sync(0xffffffffu, lb, l));
            if (bound > thr) continue;  // the threshold has tightened since the ballot
            const int cb = g * 32 + l;
            const float4 clo = __ldg(&boxesC[cb * 2]), chi = __ldg(&boxesC[cb * 2 + 1]);
            if (!__any_sync(0xffffffffu, qi >= 0 && point_gap2(q, clo, chi) <= b3)) continue;
            visit(cb);
            thr = threshold();
        }
    }
    if (pair_counter && lane == 0) atomicAdd(pair_counter, (unsigned long long)visited * 1024ull);
    if (qi >= 0) {
        float *dd = dist2 + ((size_t)scene * n + qi) * 3;
        int *ii = idx + ((size_t)scene * n + qi) * 3;
        dd[0] = b1; dd[1] = b2; dd[2] = b3;
        ii[0] = i1; ii[1] = i2; ii[2] = i3;
        if (weight) fp_weights3(b1, b2, b3, weight + ((size_t)scene * n + qi) * 3);
    }
}

// ---- knn_point: the k nearest candidates of every query, ascending (distance, index) -----------------
// One WARP per query.  The running k best are one sorted list spread over the lanes (rank r = register
// r / 32 of lane r % 32) as 64-bit keys {distance bits, index}: unsigned order of the key is the
// (distance, index) order the op is defined by (distances are >= 0; NaN padding maps to the maximum key).
// A bucket of 32 candidates costs one distance per lane and one ballot against the current k-th best;
// only if some candidate beats it is the bucket sorted (bitonic, shuffles) and merged into the list.
// (Lane-per-query with an insertion list, the three_nn layout, was measured first: every insertion of one
// lane stalls the other 31, 1.4 ms for 16 x 4096 queries, k = 16.)
// INDEXED: candidate buckets are visited best-first by their exact box lower bound and the search stops at
// the first bucket that cannot reach the k-th best.  Otherwise every bucket of the cloud is tested in order.
__device__ __forceinline__ unsigned long long shfl_xor_u64(unsigned long long v, int m) {
    return ((unsigned long long)__shfl_xor_sync(0xffffffffu, (unsigned)(v >> 32), m) << 32) |
           __shfl_xor_sync(0xffffffffu, (unsigned)v, m);
}
__device__ __forceinline__ unsigned long long shfl_u64(unsigned long long v, int src) {
    return ((unsigned long long)__shfl_sync(0xffffffffu, (unsigned)(v >> 32), src) << 32) |
           __shfl_sync(0xffffffffu, (unsigned)v, src);
}
__device__ __forceinline__ unsigned long long bitonic_merge32(unsigned long long v, int lane) {  // bitonic -> ascending
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const unsigned long long o = shfl_xor_u64(v, j);
        v = ((lane & j) == 0) ? (v < o ? v : o) : (v < o ? o : v);
    }
    return v;
}
__device__ __forceinline__ unsigned long long bitonic_sort32(unsigned long long v, int lane) {
#pragma unroll
    for (int k2 = 2; k2 <= 32; k2 <<= 1) {
#pragma unroll
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            const unsigned long long o = shfl_xor_u64(v, j);
            const bool keep_min = ((lane & j) == 0) == ((lane & k2) == 0);
            v = keep_min ? (v < o ? v : o) : (v < o ? o : v);
        }
    }
    return v;
}

template <int R, bool INDEXED>
__global__ void __launch_bounds__(kSqWarps * 32)
    knn_kernel(int n, int m, int k, int nq, int nbc, const float *__restrict__ xyzQ, const float *__restrict__ xyzC,
               const float4 *__restrict__ sortedQ, const float4 *__restrict__ sortedC, const float4 *__restrict__ boxesC,
               float *__restrict__ val, int *__restrict__ idx, int pairs) {
    constexpr unsigned long long kNone = ~0ull;
    constexpr int kMaxLb = kMaxIndexed / 32 / 32;  // candidate buckets per lane and pass
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int scene = blockIdx.y;
    const int p = blockIdx.x * kSqWarps + warp;  // query slot (curve order when INDEXED)
    if (p >= nq) return;
    float4 q;
    if (INDEXED) {
        q = __ldg(&sortedQ[(size_t)scene * nq + p]);
        sortedC += (size_t)scene * nbc * 32;
        boxesC += (size_t)scene * nbc * 2;
    } else {
        const float *src = xyzQ + ((size_t)scene * m + p) * 3;
        q = make_float4(__ldg(src), __ldg(src + 1), __ldg(src + 2), __int_as_float(p));
        xyzC += (size_t)scene * n * 3;
    }
    const int qi = __float_as_int(q.w);
    if (qi < 0) return;  // padding of the sorted query cloud
    unsigned long long best[R];
#pragma unroll
    for (int r = 0; r < R; ++r) best[r] = kNone;
    unsigned long long worst = kNone;  // key of rank k-1
    const int wr = (k - 1) >> 5, wl = (k - 1) & 31;

    auto offer = [&](float cx, float cy, float cz, int ci) {
        const float d = sqdist(q.x - cx, q.y - cy, q.z - cz);
        unsigned long long key = (d == d && ci >= 0) ? (((unsigned long long)__float_as_uint(d) << 32) | (unsigned)ci) : kNone;
        if (!__any_sync(0xffffffffu, key < worst)) return;
        unsigned long long cand = bitonic_sort32(key, lane);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const unsigned long long rev = shfl_u64(cand, 31 - lane);
            const unsigned long long lo = best[r] < rev ? best[r] : rev;
            const unsigned long long hi = best[r] < rev ? rev : best[r];
            best[r] = bitonic_merge32(lo, lane);
            if (r + 1 < R) cand = bitonic_merge32(hi, lane);
        }
        unsigned long long w = best[0];
#pragma unroll
        for (int r = 1; r < R; ++r)
            if (r == wr) w = best[r];
        worst = shfl_u64(w, wl);
    };

    if (!INDEXED) {
        for (int j0 = 0; j0 < n; j0 += 32) {
            const int j = j0 + lane;
            float cx = 0.f, cy = 0.f, cz = 0.f;
            if (j < n) {
                cx = __ldg(&xyzC[(size_t)j * 3]);
                cy = __ldg(&xyzC[(size_t)j * 3 + 1]);
                cz = __ldg(&xyzC[(size_t)j * 3 + 2]);
            }
            offer(cx, cy, cz, j < n ? j : -1);
        }
    } else {
        // exact lower bound of every candidate bucket for this query; visited in ascending order of it.  A cloud of
        // more than kMaxIndexed points (indexed in parts, see cloud_index_kernel) is searched in passes of
        // kMaxLb * 32 buckets: within a pass every bucket whose bound can still beat the k-th best is visited, and the
        // k-th best carries over, so the result is that of one pass over all buckets.
        for (int base = 0; base < nbc; base += kMaxLb * 32) {
            float lb[kMaxLb];
#pragma unroll
            for (int u = 0; u < kMaxLb; ++u) {
                const int cb = base + u * 32 + lane;
                lb[u] = INFINITY;
                if (cb < nbc) lb[u] = point_gap2(q, __ldg(&boxesC[cb * 2]), __ldg(&boxesC[cb * 2 + 1]));
            }
            for (;;) {
                float lmin = lb[0];
                int lu = 0;
#pragma unroll
                for (int u = 1; u < kMaxLb; ++u)
                    if (lb[u] < lmin) {
                        lmin = lb[u];
                        lu = u;
                    }
                const unsigned gmin = __reduce_min_sync(0xffffffffu, __float_as_uint(lmin));  // lower bounds are >= 0
                if (gmin == 0x7F800000u) break;                                             // every bucket visited
                if ((unsigned)(worst >> 32) < gmin) break;  // '<': an equal distance can still win on the index
                const int src = __ffs(__ballot_sync(0xffffffffu, __float_as_uint(lmin) == gmin)) - 1;
                const int cb = base + __shfl_sync(0xffffffffu, lu, src) * 32 + src;
                if (lane == src) {
#pragma unroll
                    for (int u = 0; u < kMaxLb; ++u)
                        if (u == lu) lb[u] = INFINITY;
                }
                const float4 c = __ldg(&sortedC[(size_t)cb * 32 + lane]);
                offer(c.x, c.y, c.z, __float_as_int(c.w));
            }
        }
    }
    float *vv = val + ((size_t)scene * m + qi) * k;
    int *ii = idx + ((size_t)scene * m + qi) * k * (pairs ? 2 : 1);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int s2 = r * 32 + lane;
        if (s2 < k) {
            vv[s2] = __uint_as_float((unsigned)(best[r] >> 32));
            if (pairs) {  // (batch index, point index): what tf.gather_nd takes (hf/core/pointfly.py:206-210)
                *reinterpret_cast<int2 *>(ii + 2 * s2) = make_int2(scene, (int)(unsigned)best[r]);
            } else {
                ii[s2] = (int)(unsigned)best[r];
            }
        }
    }
}

template <int R>
int launch_knn(bool indexed, int b, int n, int m, int k, const float *xyz1, const float *xyz2, const CloudDesc *q,
               const CloudDesc *c, float *val, int *idx, int pairs, cudaStream_t s) {
    if (indexed) {
        const int nq = q->npad, nbc = c->npad / 32;
        knn_kernel<R, true><<<dim3((nq + kSqWarps - 1) / kSqWarps, b), kSqWarps * 32, 0, s>>>(
            n, m, k, nq, nbc, xyz2, xyz1, q->sorted, c->sorted, c->boxes, val, idx, pairs);
    } else {
        knn_kernel<R, false><<<dim3((m + kSqWarps - 1) / kSqWarps, b), kSqWarps * 32, 0, s>>>(
            n, m, k, m, 0, xyz2, xyz1, nullptr, nullptr, nullptr, val, idx, pairs);
    }
    HF_LAUNCH_OK("knn_kernel");
    return HF_OK;
}

thread_local int g_spatial_mode = 0;  // 0 = auto, 1 = never (every-pair kernels), 2 = whenever a shape exists

inline int round32(int v) { return (v + 31) & ~31; }
// 16 B/point + 32 B/bucket (+ 32 B per 32 buckets for clouds above kMaxIndexed points: the second box level)
inline int super_count(int n) { return round32(n) > kMaxIndexed ? (round32(n) / 32 + 31) / 32 : 0; }
inline size_t cloud_bytes(int b, int n) { return (size_t)b * ((size_t)round32(n) * 17 + (size_t)super_count(n) * 32); }

// Is the pruned path used for a call whose query cloud has nq points and candidate cloud nc?
bool spatial_applies(int nq, int nc) {
    if (g_spatial_mode == 1) return false;
    if (nq < 1 || nc < 1 || nq > kMaxCloud || nc > kMaxCloud) return false;
    if (g_spatial_mode == 2) return true;
    // measured (tools/spatial_bench.py, profiles/r2_spatial_bench.txt): 1024 queries in 4096 points, 16 scans: ball query
    // 92 -> 58 us including both sorts (33 us with the indices prebuilt), three_nn 84 -> 71 (45) us; 256 in 1024: no gain
    return nq >= 64 && nc >= 64 && (long long)nq * nc >= (1ll << 22);
}

// knn_point profits from the sort much earlier (tools/knn_bench.py: 4096 x 1024, k = 32: 0.08 vs 0.41 ms).
bool knn_applies(int nq, int nc) {
    if (g_spatial_mode == 1) return false;
    if (nq < 1 || nc < 1 || nq > kMaxCloud || nc > kMaxCloud) return false;
    if (g_spatial_mode == 2) return true;
    return nq >= 64 && nc >= 64 && (long long)nq * nc >= (1ll << 20);
}

// A cloud index in caller memory: [sorted: b x npad float4][boxes: b x npad/32 x 2 float4].
CloudDesc carve_index(int b, int n, const float *xyz, void *mem) {
    unsigned char *w = static_cast<unsigned char *>(mem);
    CloudDesc d;
    d.xyz = xyz;
    d.n = n;
    d.npad = round32(n);
    d.parts = (n + kMaxIndexed - 1) / kMaxIndexed;
    d.part_n = round32((n + d.parts - 1) / d.parts);
    d.per_thread = (d.part_n + kIdxThreads - 1) / kIdxThreads;
    d.sorted = reinterpret_cast<float4 *>(w);
    d.boxes = reinterpret_cast<float4 *>(w + (size_t)b * d.npad * 16);
    d.nsuper = super_count(n);
    d.super = d.nsuper ? reinterpret_cast<float4 *>(w + (size_t)b * d.npad * 17) : nullptr;
    return d;
}

// Second box level of a large cloud: super-box j = bounding box of buckets 32j .. 32j+31 (one warp each).  The RoI crop
// tests 64 of these instead of the 2048 bucket boxes of a 65536-point scan before it looks at buckets at all.
__global__ void __launch_bounds__(256) cloud_super_kernel(int nb, int nsuper, const float4 *__restrict__ boxes,
                                                          float4 *__restrict__ super) {
    const int lane = threadIdx.x & 31;
    const int sb = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int scene = blockIdx.y;
    if (sb >= nsuper) return;
    const int bk = sb * 32 + lane;
    float lx = INFINITY, ly = INFINITY, lz = INFINITY, hx = -INFINITY, hy = -INFINITY, hz = -INFINITY;
    if (bk < nb) {
        const float4 lo = boxes[((size_t)scene * nb + bk) * 2], hi = boxes[((size_t)scene * nb + bk) * 2 + 1];
        if (__float_as_int(lo.w) != INT_MAX) {  // the bucket holds a real point
            lx = lo.x; ly = lo.y; lz = lo.z;
            hx = hi.x; hy = hi.y; hz = hi.z;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o)); ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o));
        lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o)); hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o));
        hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
    }
    if (lane == 0) {
        super[((size_t)scene * nsuper + sb) * 2] = make_float4(lx, ly, lz, 0.f);
        super[((size_t)scene * nsuper + sb) * 2 + 1] = make_float4(hx, hy, hz, 0.f);
    }
}

// Sort one cloud (or the two clouds of a call, in one launch) into their index memory.
int launch_index(int b, const CloudDesc &a, const CloudDesc *second, cudaStream_t s) {
    HF_CUDA_OK(cudaFuncSetAttribute(cloud_index_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kIdxSmem));
    const int gparts = second && second->parts > a.parts ? second->parts : a.parts;
    cloud_index_kernel<<<dim3(b * gparts, second ? 2 : 1), kIdxThreads, kIdxSmem, s>>>(a, second ? *second : a, gparts);
    HF_LAUNCH_OK("cloud_index_kernel");
    for (const CloudDesc *d : {&a, second}) {
        if (!d || !d->nsuper) continue;
        cloud_super_kernel<<<dim3((d->nsuper + 7) / 8, b), 256, 0, s>>>(d->npad / 32, d->nsuper, d->boxes, d->super);
        HF_LAUNCH_OK("cloud_super_kernel");
    }
    return HF_OK;
}

// hf_spatial_set_pair_counter: device counter the pruned ball query / three_nn add their executed (query, candidate)
// distance evaluations to (measurement aid; per calling thread, NULL = off)
thread_local unsigned long long *g_pair_counter = nullptr;

int launch_ball_query(int b, int m, float radius, int nsample, const CloudDesc &q, const CloudDesc &c, int *idx,
                      int *pts_cnt, cudaStream_t s) {
    const size_t rowbytes = (size_t)kSqWarps * 32 * sizeof(float4) + (size_t)(kSqWarps + 1) * nsample * 33 * sizeof(int);
    const int nbq = q.npad / 32, nbc = c.npad / 32;
    if (rowbytes > 48 * 1024)
        HF_CUDA_OK(cudaFuncSetAttribute(ball_query_indexed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rowbytes));
    ball_query_indexed_kernel<<<dim3(nbq, b), kSqWarps * 32, rowbytes, s>>>(m, nsample, ball_threshold(radius), nbq, nbc,
                                                                           q.sorted, q.boxes, c.sorted, c.boxes, idx, pts_cnt,
                                                                           g_pair_counter);
    HF_LAUNCH_OK("ball_query_indexed_kernel");
    return HF_OK;
}

int launch_three_nn(int b, int n, const CloudDesc &q, const CloudDesc &c, float *dist2, int *idx, float *weight,
                    cudaStream_t s) {
    const int nbq = q.npad / 32, nbc = c.npad / 32;
    three_nn_indexed_kernel<<<dim3((nbq + kSqWarps - 1) / kSqWarps, b), kSqWarps * 32, 0, s>>>(
        n, nbq, nbc, q.sorted, q.boxes, c.sorted, c.boxes, dist2, idx, weight, g_pair_counter);
    HF_LAUNCH_OK("three_nn_indexed_kernel");
    return HF_OK;
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace
}  // namespace hf

extern "C" void hf_spatial_set_mode(int mode) { hf::g_spatial_mode = mode; }
extern "C" void hf_spatial_set_pair_counter(unsigned long long *device_counter) { hf::g_pair_counter = device_counter; }
extern "C" int hf_spatial_applies(int nq, int nc) { return hf::spatial_applies(nq, nc) ? 1 : 0; }

extern "C" size_t hf_cloud_index_bytes(int b, int n) {
    using namespace hf;
    if (b <= 0 || n <= 0 || n > kMaxCloud) return 0;
    return cloud_bytes(b, n);
}

// The one place outside carve_index that hands out pointers into an index: cropping.cu and fps_multi.cu go through it.
bool hf::cloud_index_view(int b, int n, const void *index, CloudIndexView *view) {
    if (!index || !aligned16(index) || b < 0 || hf_cloud_index_bytes(b, n) == 0) return false;
    const CloudDesc d = carve_index(b, n, nullptr, const_cast<void *>(index));
    view->sorted = d.sorted;
    view->boxes = d.boxes;
    view->super = d.super;
    view->npad = d.npad;
    view->nsuper = d.nsuper;
    return true;
}

extern "C" size_t hf_spatial_workspace_bytes(int b, int n, int m) {
    const size_t a = hf_cloud_index_bytes(b, n), c = hf_cloud_index_bytes(b, m);
    return (a && c) ? a + c : 0;
}

extern "C" int hf_cloud_index_build(int b, int n, const float *xyz, void *index, size_t index_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1 && n <= kMaxCloud, "CloudIndex: 1 <= num_points <= %d", kMaxCloud);
    HF_REQUIRE(b <= 65535, "CloudIndex: batch_size > 65535 not supported");
    HF_REQUIRE(xyz && index && aligned16(index), "CloudIndex: null or unaligned pointer");
    if (index_bytes < hf_cloud_index_bytes(b, n)) {
        set_error("CloudIndex: index memory too small (%zu < %zu bytes)", index_bytes, hf_cloud_index_bytes(b, n));
        return HF_ERR_WORKSPACE;
    }
    if (b == 0) return HF_OK;
    return launch_index(b, carve_index(b, n, xyz, index), nullptr, (cudaStream_t)stream);
}

extern "C" int hf_query_ball_point_indexed(int b, int n, int m, float radius, int nsample, const float *xyz1,
                                           const float *xyz2, int *idx, int *pts_cnt, const void *index1,
                                           const void *index2, hf_stream_t stream) {
    using namespace hf;
    if (!index1 || !index2 || b <= 0 || m <= 0 || n <= 0 || b > 65535 || !(radius > 0) || nsample <= 0 || nsample > 128 ||
        !spatial_applies(m, n) || !aligned16(index1) || !aligned16(index2))
        return hf_query_ball_point(b, n, m, radius, nsample, xyz1, xyz2, idx, pts_cnt, stream);
    HF_REQUIRE(xyz1 && xyz2 && idx && pts_cnt, "QueryBallPoint: null pointer");
    return launch_ball_query(b, m, radius, nsample, carve_index(b, m, xyz2, const_cast<void *>(index2)),
                             carve_index(b, n, xyz1, const_cast<void *>(index1)), idx, pts_cnt, (cudaStream_t)stream);
}

extern "C" int hf_three_nn_indexed(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                                   const void *index_unknown, const void *index_known, hf_stream_t stream) {
    using namespace hf;
    if (!index_unknown || !index_known || b <= 0 || n <= 0 || m <= 0 || b > 65535 || !spatial_applies(n, m) ||
        !aligned16(index_unknown) || !aligned16(index_known))
        return hf_three_nn(b, n, m, unknown, known, dist2, idx, stream);
    HF_REQUIRE(unknown && known && dist2 && idx, "ThreeNN: null pointer");
    return launch_three_nn(b, n, carve_index(b, n, unknown, const_cast<void *>(index_unknown)),
                           carve_index(b, m, known, const_cast<void *>(index_known)), dist2, idx, nullptr,
                           (cudaStream_t)stream);
}

extern "C" int hf_query_ball_point_ws(int b, int n, int m, float radius, int nsample, const float *xyz1,
                                      const float *xyz2, int *idx, int *pts_cnt, void *workspace,
                                      size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    if (!workspace || b <= 0 || m <= 0 || n <= 0 || b > 65535 || !(radius > 0) || nsample <= 0 || nsample > 128 ||
        !spatial_applies(m, n) || workspace_bytes < hf_spatial_workspace_bytes(b, n, m) || !aligned16(workspace))
        return hf_query_ball_point(b, n, m, radius, nsample, xyz1, xyz2, idx, pts_cnt, stream);
    HF_REQUIRE(xyz1 && xyz2 && idx && pts_cnt, "QueryBallPoint: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const CloudDesc c = carve_index(b, n, xyz1, workspace);
    const CloudDesc q = carve_index(b, m, xyz2, static_cast<unsigned char *>(workspace) + cloud_bytes(b, n));
    int rc = launch_index(b, c, &q, s);
    if (rc != HF_OK) return rc;
    return launch_ball_query(b, m, radius, nsample, q, c, idx, pts_cnt, s);
}

extern "C" int hf_three_nn_ws(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                              void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    if (!workspace || b <= 0 || n <= 0 || m <= 0 || b > 65535 || !spatial_applies(n, m) ||
        workspace_bytes < hf_spatial_workspace_bytes(b, n, m) || !aligned16(workspace))
        return hf_three_nn(b, n, m, unknown, known, dist2, idx, stream);
    HF_REQUIRE(unknown && known && dist2 && idx, "ThreeNN: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const CloudDesc q = carve_index(b, n, unknown, workspace);
    const CloudDesc c = carve_index(b, m, known, static_cast<unsigned char *>(workspace) + cloud_bytes(b, n));
    int rc = launch_index(b, q, &c, s);
    if (rc != HF_OK) return rc;
    return launch_three_nn(b, n, q, c, dist2, idx, nullptr, s);
}

extern "C" int hf_three_nn_weights(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx,
                                   float *weight, const void *index_unknown, const void *index_known,
                                   hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(weight, "ThreeNN: null weight pointer");
    if (!index_unknown || !index_known || b <= 0 || n <= 0 || m <= 0 || b > 65535 || !spatial_applies(n, m) ||
        !aligned16(index_unknown) || !aligned16(index_known))
        return three_nn_every_pair(b, n, m, unknown, known, dist2, idx, weight, (cudaStream_t)stream);
    HF_REQUIRE(unknown && known && dist2 && idx, "ThreeNN: null pointer");
    return launch_three_nn(b, n, carve_index(b, n, unknown, const_cast<void *>(index_unknown)),
                           carve_index(b, m, known, const_cast<void *>(index_known)), dist2, idx, weight,
                           (cudaStream_t)stream);
}

/* knn_point: k nearest points of xyz1 (b,n,3) for every query of xyz2 (b,m,3), ascending (distance, index).
 * workspace (hf_knn_workspace_bytes(b, n, m), may be NULL / 0 bytes): lets calls large enough to profit sort both
 * clouds and prune; results do not depend on it. */
extern "C" size_t hf_knn_workspace_bytes(int b, int n, int m) {
    return hf::knn_applies(m, n) ? hf_spatial_workspace_bytes(b, n, m) : 0;
}

namespace hf {
namespace {
int knn_any(int b, int n, int m, int k, const float *xyz1, const float *xyz2, float *val, int *idx, int pairs,
            void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    HF_REQUIRE(k > 0, "KnnPoint expects positive k");
    HF_REQUIRE(b >= 0 && n >= 1 && m >= 0, "KnnPoint expects (batch_size, ndataset, 3) xyz1 shape");
    HF_REQUIRE(k <= n, "KnnPoint: k must not exceed ndataset (input must have at least k columns)");
    HF_REQUIRE(b <= 65535, "KnnPoint: batch_size > 65535 not supported");
    HF_REQUIRE(k <= 128, "KnnPoint: k > 128 not supported");
    HF_REQUIRE(xyz1 && (m == 0 || (xyz2 && val && idx)), "KnnPoint: null pointer");
    if (b == 0 || m == 0) return HF_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const bool indexed = workspace && aligned16(workspace) && knn_applies(m, n) && hf_spatial_workspace_bytes(b, n, m) != 0 &&
                         workspace_bytes >= hf_spatial_workspace_bytes(b, n, m);
    CloudDesc c = {}, q = {};
    if (indexed) {
        c = carve_index(b, n, xyz1, workspace);
        q = carve_index(b, m, xyz2, static_cast<unsigned char *>(workspace) + cloud_bytes(b, n));
        int rc = launch_index(b, c, &q, s);
        if (rc != HF_OK) return rc;
    }
    if (k <= 32) return launch_knn<1>(indexed, b, n, m, k, xyz1, xyz2, &q, &c, val, idx, pairs, s);
    if (k <= 64) return launch_knn<2>(indexed, b, n, m, k, xyz1, xyz2, &q, &c, val, idx, pairs, s);
    return launch_knn<4>(indexed, b, n, m, k, xyz1, xyz2, &q, &c, val, idx, pairs, s);
}
}  // namespace
}  // namespace hf

extern "C" int hf_knn_point(int b, int n, int m, int k, const float *xyz1, const float *xyz2, float *val, int *idx,
                            void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    return hf::knn_any(b, n, m, k, xyz1, xyz2, val, idx, 0, workspace, workspace_bytes, stream);
}

extern "C" int hf_knn_indices_general(int b, int p, int n, int k, const float *queries, const float *points,
                                      float *distances, int *indices, void *workspace, size_t workspace_bytes,
                                      hf_stream_t stream) {
    return hf::knn_any(b, n, p, k, points, queries, distances, indices, 1, workspace, workspace_bytes, stream);
}
