// fps_bucket.cu — farthest point sampling with exact spatial pruning, for sm_100a.
//
// Replaces sampling/tf_sampling_g.cu:105-170 of the reference for clouds of up to 65536
// points (larger ones take the fallback in sampling.cu).  Same result bit for bit, a
// different algorithm:
//
//   The reference (and the plain on-chip kernel in sampling.cu) recomputes the distance of
//   EVERY point to the newest sample in each of the m-1 sequential iterations.  But a
//   point's running minimum td[p] only changes if the new sample c is closer than
//   sqrt(td[p]), and after the first few iterations that is true for ~1% of the cloud.
//   So the cloud is sorted along a Morton curve once (in shared memory, in the kernel's
//   prologue) and cut into BUCKETS of 32 consecutive points; each bucket keeps its bounding
//   box, the maximum of its td values and which point holds it.  Per iteration a bucket
//   is touched only if
//        LB = sqdist(clamped per-axis gap between c and the bucket's box)  <=  max td
//   Because every IEEE operation involved is monotone and LB is computed with the SAME
//   operation sequence as the point distances (common.cuh:sqdist), LB <= fl(d(p,c)) holds
//   exactly for every point p of the bucket, so a skipped bucket is one where
//   fminf(d, td[p]) == td[p] for all its points: skipping changes nothing, no slack needed.
//   On KITTI-shaped scans ~1% of the buckets are active per iteration (16384 -> 4096).
//
//   Layout: a bucket is 32 Morton-consecutive points, one per lane: bucket (w, j) = register
//   j of the 32 lanes of warp w (td lives in registers), and lane l of warp w owns the
//   metadata of bucket (w, l) (box, max td, ballot of the lanes that hold the max), so the
//   bucket test of a whole warp is one 32-lane pass.  Consecutive Morton buckets go to
//   different warps (round robin), so the few active buckets of an iteration are processed
//   by different warps in parallel.  The m-1 iterations are a pure latency chain — ~4.5
//   cycles per dependent instruction, ~30 per shared-memory or redux.sync round trip
//   (profiles/r1_fps_bucket_notes.md) — so the design minimises the number of dependent steps:
//       test (13 flops, all lanes) -> per active bucket: 3 LDS + 8 flops + redux.sync (max)
//       + ballot (who holds it), nothing else -> warp arg-max over the bucket maxima (one
//       redux.sync + ballot when unique) -> the owning lane alone resolves the point (its
//       coordinates and tie word) and publishes {key, xyz} -> __syncthreads -> arg-max over
//       the <= 32 warp candidates -> one LDS.128 gives the winner's coordinates.
//   Exact ties (duplicate points) take out-of-line slow paths.
//   xyz (SoA, sorted) and the 16-bit local index of every point: 14 B/point, 16384
//   points = 224 KB of the 227 KB a CTA can have.  Clouds of up to 65536 points are spread
//   over a cluster of 2/4/8 CTAs; their per-CTA winners are exchanged with st.async
//   (DSMEM) completing on the receivers' mbarriers.
//   The reference's tie-break (max distance, then lowest index mod 512, then lowest
//   index; tf_sampling_g.cu:146-163) is the explicit 64-bit key (distance bits, tie word).
#include "common.cuh"

namespace hf {
namespace {

__device__ __forceinline__ uint32_t fb_cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void fb_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                     : "memory");
}
__device__ __forceinline__ uint32_t fb_mapa(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void fb_st_async_v4(uint32_t raddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                               uint32_t rbar) {
    asm volatile(
        "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(
            raddr),
        "r"(a), "r"(b), "r"(c), "r"(d), "r"(rbar)
        : "memory");
}
__device__ __forceinline__ void fb_st_async_b32(uint32_t raddr, uint32_t a, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
                 "r"(a), "r"(rbar)
                 : "memory");
}


// order-preserving float <-> int (for redux.sync min/max on coordinates)
__device__ __forceinline__ int fb_f2ord(float f) {
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float fb_ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__device__ __forceinline__ uint32_t fb_spread10(uint32_t v) {  // 10 bits -> every third bit
    v &= 1023u;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

constexpr int kFbMsgBytes = 20;   // {lo, hi, x, y} + {z}
constexpr uint32_t kFbPadMorton = 0x7FFFFFFFu;


// Low word of the composite key: larger = preferred.  Encodes (k mod 512, k) ascending.
__device__ __forceinline__ uint32_t fb_tie_word(uint32_t k) { return ~(((k & 511u) << 23) | (k >> 9)); }
__device__ __forceinline__ uint32_t fb_tie_decode(uint32_t lo) {
    const uint32_t inv = ~lo;
    return ((inv & 0x7FFFFFu) << 9) | (inv >> 23);
}

// Exact ties are rare on real clouds and common on duplicate-padded ones; the slow paths
// are kept out of line so that the common case is a not-taken branch, not a predicated
// tail of dependent instructions on the per-iteration latency chain.
__device__ __noinline__ unsigned fb_tie_lanes(uint32_t hi, uint32_t lo, uint32_t hmax) {
    const uint32_t l2 = hi == hmax ? lo : 0u;
    const uint32_t lmax = __reduce_max_sync(0xffffffffu, l2);
    return __ballot_sync(0xffffffffu, hi == hmax && l2 == lmax);
}
// Arg-max of a per-lane key (hi = distance bits, lo = tie word): returns the winning lane.
// A unique maximum of `hi` costs one redux.sync and one ballot.
__device__ __forceinline__ int fb_argmax_lane(uint32_t hi, uint32_t lo) {
    const uint32_t hmax = __reduce_max_sync(0xffffffffu, hi);
    unsigned who = __ballot_sync(0xffffffffu, hi == hmax);
    if (who & (who - 1u)) who = fb_tie_lanes(hi, lo, hmax);  // warp-uniform
    return __ffs(who) - 1;
}
// Several points of one bucket hold its maximum (bit l of `who` = point at sorted position
// s0 + l): the real one (local index < cnt) with the largest tie word wins; -1 if none is real.
__device__ __noinline__ int fb_tie_point(unsigned who, const uint16_t *idx_s0, uint32_t base, int cnt) {
    int bestl = -1;
    uint32_t bk = 0u;
    while (who) {
        const int l = __ffs(who) - 1;
        who &= who - 1u;
        const uint32_t li = idx_s0[l];
        if ((int)li < cnt) {
            const uint32_t tk = fb_tie_word(base + li);
            if (bestl < 0 || tk > bk) {
                bk = tk;
                bestl = l;
            }
        }
    }
    return bestl;
}

// One bucket (32 points, one per lane) against the newest sample.  `tdj` is this lane's
// running minimum for its point of the bucket (a register of the caller's td[] array,
// selected at compile time).  Only the maximum and the ballot of the lanes that hold it
// are produced here; which point that is gets resolved later, and only for the one bucket
// per warp that becomes the warp's candidate.
template <int W>
__device__ __forceinline__ void fb_process_bucket(int jj, float &tdj, const float *sx, const float *sy,
                                                  const float *sz, int warp, int lane, float cx, float cy,
                                                  float cz, uint32_t &bmax, uint32_t &bwho) {
    const int s = (jj * W + warp) * 32 + lane;
    const float d = sqdist(sx[s] - cx, sy[s] - cy, sz[s] - cz);  // tf_sampling_g.cu:142
    const float tn = fminf(d, tdj);                               // :143
    tdj = tn;
    // distances are >= +0, so their bit patterns order like unsigned ints; padding (-1) -> 0
    const uint32_t v = __float_as_uint(fmaxf(tn, 0.f));
    const uint32_t vm = __reduce_max_sync(0xffffffffu, v);
    const unsigned who = __ballot_sync(0xffffffffu, v == vm);
    if (lane == jj) {
        bmax = vm;
        bwho = who;
    }
}

#define FB_CASE(JJ)                                                                                           \
    case JJ:                                                                                                  \
        if constexpr (JJ < J) fb_process_bucket<W>(JJ, td[JJ], sx, sy, sz, warp, lane, cx, cy, cz, bmax, bwho); \
        break;

template <int CLUSTER, int W, int J>
__global__ void __launch_bounds__(W * 32, 1)
    fps_bucket_kernel(int n, int m, int per, const float *__restrict__ inp, int *__restrict__ out) {
    constexpr int T = W * 32;
    constexpr int CAP = T * J;   // points held by one CTA: W warps x J buckets x 32 points
    static_assert((CAP & (CAP - 1)) == 0 && CAP <= 65536, "CAP must be a power of two that fits 16-bit local indices");
    static_assert(J <= 32 && CLUSTER <= 8 && W <= 32, "bucket metadata lives in one lane per bucket");

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *sx = reinterpret_cast<float *>(smem_raw);
    float *sy = sx + CAP;
    float *sz = sy + CAP;
    uint16_t *sidx = reinterpret_cast<uint16_t *>(sz + CAP);
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(smem_raw);  // prologue only: 8*CAP <= 12*CAP
    __shared__ __align__(8) uint2 cand[2][32];         // candidate keys {tie16 << 16 | warp, distance bits}, one per warp
    __shared__ __align__(16) float4 slot_xyz[2][32];   // candidate coordinates, one slot per warp
    __shared__ __align__(16) uint4 cl_key[2][8];       // cluster exchange: {key lo, key hi, x, y} per CTA
    __shared__ uint32_t cl_z[2][8];
    __shared__ __align__(8) uint64_t bars[2];
    int *red = reinterpret_cast<int *>(&slot_xyz[0][0]);  // prologue only: [6][W] ints <= 1 KB

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t rank = CLUSTER > 1 ? fb_cluster_ctarank() : 0u;
    const int scene = blockIdx.x / CLUSTER;
    const float *data = inp + (size_t)scene * n * 3;
    int *idxs = out + (size_t)scene * m;
    const uint32_t base = rank * (uint32_t)per;
    const int cnt = max(0, min(per, n - (int)base));  // real points of this CTA's slice

    // ---- prologue 1: bounding box of the slice -> Morton quantisation ----
    int mn0 = INT_MAX, mn1 = INT_MAX, mn2 = INT_MAX, mx0 = INT_MIN, mx1 = INT_MIN, mx2 = INT_MIN;
    for (int i = t; i < cnt; i += T) {
        const float *p = data + (size_t)(base + i) * 3;
        const int o0 = fb_f2ord(p[0]), o1 = fb_f2ord(p[1]), o2 = fb_f2ord(p[2]);
        mn0 = min(mn0, o0); mx0 = max(mx0, o0);
        mn1 = min(mn1, o1); mx1 = max(mx1, o1);
        mn2 = min(mn2, o2); mx2 = max(mx2, o2);
    }
    mn0 = __reduce_min_sync(0xffffffffu, mn0); mx0 = __reduce_max_sync(0xffffffffu, mx0);
    mn1 = __reduce_min_sync(0xffffffffu, mn1); mx1 = __reduce_max_sync(0xffffffffu, mx1);
    mn2 = __reduce_min_sync(0xffffffffu, mn2); mx2 = __reduce_max_sync(0xffffffffu, mx2);
    if (lane == 0) {
        red[0 * W + warp] = mn0; red[1 * W + warp] = mn1; red[2 * W + warp] = mn2;
        red[3 * W + warp] = mx0; red[4 * W + warp] = mx1; red[5 * W + warp] = mx2;
    }
    __syncthreads();
    float lo0, lo1, lo2, inv;
    {
        int l0 = INT_MAX, l1 = INT_MAX, l2 = INT_MAX, h0 = INT_MIN, h1 = INT_MIN, h2 = INT_MIN;
        for (int w = 0; w < W; ++w) {
            l0 = min(l0, red[0 * W + w]); l1 = min(l1, red[1 * W + w]); l2 = min(l2, red[2 * W + w]);
            h0 = max(h0, red[3 * W + w]); h1 = max(h1, red[4 * W + w]); h2 = max(h2, red[5 * W + w]);
        }
        lo0 = fb_ord2f(l0); lo1 = fb_ord2f(l1); lo2 = fb_ord2f(l2);
        const float ext = fmaxf(fmaxf(fb_ord2f(h0) - lo0, fb_ord2f(h1) - lo1), fb_ord2f(h2) - lo2);
        inv = (ext > 0.f && ext < 1e30f) ? 1023.f / ext : 0.f;  // cubic cells: distance pruning wants compact buckets
    }
    __syncthreads();  // red[] (aliases the slots) is dead from here

    // ---- prologue 2: (Morton code, local index) keys, bitonic sort in shared memory ----
    for (int i = t; i < CAP; i += T) {
        uint32_t mort = kFbPadMorton;
        if (i < cnt) {
            const float *p = data + (size_t)(base + i) * 3;
            const uint32_t qx = min(__float2uint_rz((p[0] - lo0) * inv), 1023u);
            const uint32_t qy = min(__float2uint_rz((p[1] - lo1) * inv), 1023u);
            const uint32_t qz = min(__float2uint_rz((p[2] - lo2) * inv), 1023u);
            mort = fb_spread10(qx) | (fb_spread10(qy) << 1) | (fb_spread10(qz) << 2);
        }
        keys[i] = ((unsigned long long)mort << 16) | (unsigned)i;
    }
    __syncthreads();
    for (unsigned k = 2; k <= (unsigned)CAP; k <<= 1) {
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            for (unsigned p = t; p < (unsigned)CAP / 2; p += T) {
                const unsigned i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
                const unsigned l = i | j;
                const bool up = (i & k) == 0;
                const unsigned long long a = keys[i], b = keys[l];
                if ((a > b) == up) {
                    keys[i] = b;
                    keys[l] = a;
                }
            }
            __syncthreads();
        }
    }

    // ---- prologue 3: sorted position (j, warp, lane) -> bucket j of this warp.  Pull my J keys
    // into registers, then overwrite the sort scratch with the sorted SoA copy of the slice.
    unsigned short li[J];
#pragma unroll
    for (int j = 0; j < J; ++j) li[j] = (unsigned short)(keys[(j * W + warp) * 32 + lane] & 0xFFFFull);
    __syncthreads();

    float td[J];
    // empty bucket: far away, so it is pruned as soon as its max is known to be 0
    float blox = 1e18f, bloy = 1e18f, bloz = 1e18f, bhix = 1e18f, bhiy = 1e18f, bhiz = 1e18f;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int s = (j * W + warp) * 32 + lane;
        const bool valid = (int)li[j] < cnt;
        float x = 0.f, y = 0.f, z = 0.f;
        if (valid) {
            const float *p = data + (size_t)(base + li[j]) * 3;
            x = p[0];
            y = p[1];
            z = p[2];
        }
        sx[s] = x;
        sy[s] = y;
        sz[s] = z;
        sidx[s] = li[j];
        td[j] = valid ? 1e38f : -1.f;  // tf_sampling_g.cu:118; padding never wins
        const int ox = fb_f2ord(x), oy = fb_f2ord(y), oz = fb_f2ord(z);
        const int lx = __reduce_min_sync(0xffffffffu, valid ? ox : INT_MAX);
        const int ly = __reduce_min_sync(0xffffffffu, valid ? oy : INT_MAX);
        const int lz = __reduce_min_sync(0xffffffffu, valid ? oz : INT_MAX);
        const int hx = __reduce_max_sync(0xffffffffu, valid ? ox : INT_MIN);
        const int hy = __reduce_max_sync(0xffffffffu, valid ? oy : INT_MIN);
        const int hz = __reduce_max_sync(0xffffffffu, valid ? oz : INT_MIN);
        if (lane == j && lx <= hx) {
            blox = fb_ord2f(lx); bloy = fb_ord2f(ly); bloz = fb_ord2f(lz);
            bhix = fb_ord2f(hx); bhiy = fb_ord2f(hy); bhiz = fb_ord2f(hz);
        }
    }
    // metadata of bucket (warp, lane): max td as ordered bits (+inf forces every bucket through
    // iteration 0) and the ballot of the lanes (= points) that hold it
    uint32_t bmax = lane < J ? 0x7F800000u : 0u, bwho = 0u;

    if (t == 0 && CLUSTER > 1) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&bars[0], CLUSTER * kFbMsgBytes);
        mbar_arrive_expect_tx(&bars[1], CLUSTER * kFbMsgBytes);
    }
    if (CLUSTER > 1) fb_cluster_sync();
    else __syncthreads();

    // start at index 0 (tf_sampling_g.cu:114-116)
    float cx = data[0], cy = data[1], cz = data[2];
    if (rank == 0 && t == 0) idxs[0] = 0;

    // cluster path: warp 0, lane r pushes this CTA's winner into slot `rank` of CTA r
    const uint32_t peer = lane < CLUSTER ? (uint32_t)lane : 0u;
    const uint32_t dst_key0 = fb_mapa(smem_u32(&cl_key[0][rank]), peer);
    const uint32_t dst_z0 = fb_mapa(smem_u32(&cl_z[0][rank]), peer);
    const uint32_t dst_bar0 = fb_mapa(smem_u32(&bars[0]), peer);

    int src = 0;  // lane (= bucket) holding this warp's candidate; recomputed only when the warp touched a bucket
    uint32_t ckey = 0u;                    // src lane only: the candidate's tie word ...
    float4 cxyz = make_float4(0.f, 0.f, 0.f, 0.f);  // ... and coordinates
    const int iters = m - 1;
    for (int it = 0; it < iters; ++it) {
        const int buf = it & 1;
        // ---- 1. bucket test: lane l tests bucket (warp, l) ----
        const float gx = fmaxf(fmaxf(blox - cx, cx - bhix), 0.f);
        const float gy = fmaxf(fmaxf(bloy - cy, cy - bhiy), 0.f);
        const float gz = fmaxf(fmaxf(bloz - cz, cz - bhiz), 0.f);
        unsigned mask = __ballot_sync(0xffffffffu, lane < J && !(sqdist(gx, gy, gz) > __uint_as_float(bmax)));
        // ---- 2./3. active buckets (compile-time register index through a jump table), then the
        // warp's candidate: the bucket with the largest maximum, resolved to a point by its owner lane
        if (mask) {
            do {
                const int j = __ffs(mask) - 1;
                mask &= mask - 1u;
                switch (j) {
                    FB_CASE(0) FB_CASE(1) FB_CASE(2) FB_CASE(3) FB_CASE(4) FB_CASE(5) FB_CASE(6) FB_CASE(7)
                    FB_CASE(8) FB_CASE(9) FB_CASE(10) FB_CASE(11) FB_CASE(12) FB_CASE(13) FB_CASE(14) FB_CASE(15)
                    FB_CASE(16) FB_CASE(17) FB_CASE(18) FB_CASE(19) FB_CASE(20) FB_CASE(21) FB_CASE(22) FB_CASE(23)
                    FB_CASE(24) FB_CASE(25) FB_CASE(26) FB_CASE(27) FB_CASE(28) FB_CASE(29) FB_CASE(30) FB_CASE(31)
                    default: break;
                }
            } while (mask);
            const uint32_t wm = __reduce_max_sync(0xffffffffu, bmax);
            unsigned wb = __ballot_sync(0xffffffffu, bmax == wm);
            const bool tied = (wb & (wb - 1u)) != 0u;  // several buckets share the maximum: every one of them resolves
            if (!tied) src = __ffs(wb) - 1;
            if (lane < J && (tied ? bmax == wm : lane == src)) {
                const int s0 = (lane * W + warp) * 32;
                int pl = __ffs(bwho) - 1;
                if (bwho & (bwho - 1u)) pl = fb_tie_point(bwho, sidx + s0, base, cnt);
                ckey = 0u;
                if (pl >= 0 && (int)sidx[s0 + pl] < cnt) ckey = fb_tie_word(base + sidx[s0 + pl]);
                if (pl < 0) pl = 0;
                cxyz = make_float4(sx[s0 + pl], sy[s0 + pl], sz[s0 + pl], 0.f);
            }
            if (tied) src = __ffs(fb_tie_lanes(bmax, ckey, wm)) - 1;
        }
        if (lane == src) {
            slot_xyz[buf][warp] = cxyz;
            cand[buf][warp] = make_uint2(ckey, bmax);
        }
        __syncthreads();
        // ---- 4. winner: arg-max over the W candidates, one per lane ----
        const uint2 cv = lane < W ? cand[buf][lane] : make_uint2(0u, 0u);
        const int wwin = fb_argmax_lane(cv.y, cv.x);
        const float4 wxyz = slot_xyz[buf][wwin];
        if (CLUSTER == 1) {
            cx = wxyz.x;
            cy = wxyz.y;
            cz = wxyz.z;
            if (warp == 0) {
                const uint32_t glo = __shfl_sync(0xffffffffu, cv.x, wwin);
                if (lane == 0) idxs[it + 1] = (int)fb_tie_decode(glo);
            }
        } else {
            const uint32_t klo = __shfl_sync(0xffffffffu, cv.x, wwin), khi = __shfl_sync(0xffffffffu, cv.y, wwin);
            if (warp == 0 && lane < CLUSTER) {
                const uint32_t dbar = dst_bar0 + buf * 8u;
                fb_st_async_v4(dst_key0 + buf * (uint32_t)sizeof(cl_key[0]), klo, khi, __float_as_uint(wxyz.x),
                               __float_as_uint(wxyz.y), dbar);
                fb_st_async_b32(dst_z0 + buf * (uint32_t)sizeof(cl_z[0]), __float_as_uint(wxyz.z), dbar);
            }
            mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
            if (t == 0 && it + 2 < iters) mbar_arrive_expect_tx(&bars[buf], CLUSTER * kFbMsgBytes);
            uint4 e = make_uint4(0u, 0u, 0u, 0u);
            uint32_t ez = 0u;
            if (lane < CLUSTER) {
                e = cl_key[buf][lane];
                ez = cl_z[buf][lane];
            }
            const int wsrc = fb_argmax_lane(e.y, e.x);
            cx = __uint_as_float(__shfl_sync(0xffffffffu, e.z, wsrc));
            cy = __uint_as_float(__shfl_sync(0xffffffffu, e.w, wsrc));
            cz = __uint_as_float(__shfl_sync(0xffffffffu, ez, wsrc));
            if (rank == 0 && warp == 0) {
                const uint32_t glo = __shfl_sync(0xffffffffu, e.x, wsrc);
                if (lane == 0) idxs[it + 1] = (int)fb_tie_decode(glo);
            }
        }
    }
    if (CLUSTER > 1) fb_cluster_sync();  // no CTA may exit while peers can still push into it
}
#undef FB_CASE

template <int CLUSTER, int W, int J>
int launch_fps_bucket(int b, int n, int m, const float *inp, int *out, cudaStream_t stream) {
    constexpr int CAP = W * 32 * J;
    auto kern = fps_bucket_kernel<CLUSTER, W, J>;
    const size_t smem = (size_t)14 * CAP;
    if (smem > 40 * 1024) HF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int per = (n + CLUSTER - 1) / CLUSTER;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)b * CLUSTER);
    cfg.blockDim = dim3(W * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HF_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, n, m, per, inp, out));
    HF_LAUNCH_OK("fps_bucket_kernel");
    return HF_OK;
}

}  // namespace

// Launch the bucket-pruned FPS.  w_req/j_req = 0 -> heuristic shape (warps per CTA, buckets
// per warp).  Returns HF_ERR_UNSUPPORTED when no instantiated shape covers the request.
int fps_bucket_dispatch(int b, int n, int m, const float *inp, int *out, cudaStream_t s, int w_req, int j_req) {
#define FB_TRY(C, W, J)                                                                            \
    if (cluster == C && w == W && j == J) return launch_fps_bucket<C, W, J>(b, n, m, inp, out, s);
    // as few CTAs per scene as hold the cloud (16384 points each, or the requested shape's capacity)
    const int cap_req = (w_req > 0 && j_req > 0) ? w_req * 32 * j_req : 16384;
    int cluster = 1;
    while (cluster < 8 && (n + cluster - 1) / cluster > cap_req) cluster <<= 1;
    const int per = (n + cluster - 1) / cluster;
    if (per > 16384) return HF_ERR_UNSUPPORTED;  // 16-bit local indices, 14 B/point of shared memory
    int cap = 256;
    while (cap < per) cap <<= 1;
    int w = w_req, j = j_req;
    if (w == 0 || j == 0) {
        switch (cap) {  // heuristic shapes (tools/fps_sweep.py)
            case 16384: w = 16; j = 32; break;
            case 8192: w = 16; j = 16; break;
            case 4096: w = 16; j = 8; break;
            case 2048: w = 16; j = 4; break;
            case 1024: w = 8; j = 4; break;
            case 512: w = 8; j = 2; break;
            default: w = 4; j = 2; break;
        }
    }
    if (w * 32 * j < per) return HF_ERR_UNSUPPORTED;
    FB_TRY(1, 16, 32) FB_TRY(1, 32, 16)
    FB_TRY(1, 8, 32) FB_TRY(1, 16, 16) FB_TRY(1, 32, 8)
    FB_TRY(1, 4, 32) FB_TRY(1, 8, 16) FB_TRY(1, 16, 8) FB_TRY(1, 32, 4)
    FB_TRY(1, 4, 16) FB_TRY(1, 8, 8) FB_TRY(1, 16, 4) FB_TRY(1, 32, 2)
    FB_TRY(1, 4, 8) FB_TRY(1, 8, 4) FB_TRY(1, 16, 2)
    FB_TRY(1, 4, 4) FB_TRY(1, 8, 2) FB_TRY(1, 16, 1)
    FB_TRY(1, 2, 4) FB_TRY(1, 4, 2) FB_TRY(1, 8, 1)
    FB_TRY(2, 16, 32) FB_TRY(4, 16, 32) FB_TRY(8, 16, 32) FB_TRY(2, 32, 16) FB_TRY(4, 32, 16) FB_TRY(8, 32, 16)
    FB_TRY(2, 16, 16) FB_TRY(4, 16, 16) FB_TRY(8, 16, 16)
#undef FB_TRY
    return HF_ERR_UNSUPPORTED;
}

}  // namespace hf
