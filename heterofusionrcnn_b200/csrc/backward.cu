// backward.cu — the four scatter-add gradients of the operator path, deterministic, for sm_100a.
//
// Replaces (see include/hfops.h):
//   scatteraddpointLauncher        sampling/tf_sampling_g.cu:183-192,209   (GatherPointGrad)
//   groupPointGradLauncher         grouping/tf_grouping_g.cu:61-78,137     (GroupPointGrad)
//   three_interpolate_grad_gpu     interpolate/tf_interpolate_g.cu:133-166 (ThreeInterpolateGrad)
//   pccropandsamplegradfts_gpu     cropping/tf_cropping_g.cu:134-150,161   (PcCropAndSampleGradFts)
// and the cudaMemset of the output each TF wrapper issues first (tf_sampling.cpp:174,
// tf_interpolate.cpp:178, ...).
//
// All four are "out[dst(e)] += w(e) * src_row(e)" over a flat list of entries e.  The reference
// does it with one float atomicAdd per element: the summation order, and with it the last
// bits of the result, changes from run to run, and every atomic is a separate L2 transaction.
// Here the scatter is turned into a gather:
//   1. count entries per destination row (integer atomics: exact in any order), exclusive scan;
//   2. fill the per-destination entry lists (CSR);
//   3. one warp per destination row: put its list into ascending entry order, then add the source
//      rows in that order with 128-bit loads along the channels, and write the row once.
// Result = the sum in ascending entry order (= the reference's loops run serially), bit for bit
// and run to run; every output row is written exactly once, so no memset pass; HBM traffic is one
// read of the incoming gradient + one write of the result (+ 12 B per entry of index traffic).
//
// Entries whose destination is outside [0, n) are skipped (the reference would write out of
// bounds).  The caller provides the workspace (hf_scatter_workspace_bytes).
#include <stdint.h>

#include "common.cuh"

namespace hf {
namespace {

constexpr int kBwThreads = 256;
constexpr int kBwWarps = kBwThreads / 32;
constexpr int kScanBlock = 1024;   // destinations per scan block
constexpr int kListCap = 1024;     // entries of one destination sorted in shared memory (per warp)

struct ScatterIndex {
    int S;                // scenes
    int n;                // destination rows per scene
    int K;                // entries per scene
    const int *idx;       // (S, K) destination row of entry e, relative to base
    const int *seg_base;  // optional (S, K / seg_len): dst = idx[e] + seg_base[e / seg_len] * seg_stride
    int seg_len, seg_stride;
    // workspace
    int *cnt;   // [T] entries per destination (T = S * n); zero again after the fill
    int *off;   // [T] exclusive scan inside the destination's scan block
    int *blk;   // [nb + 1] exclusive scan of the block totals; blk[nb] = number of listed entries
    int *list;  // [S * K] entries grouped by destination
};

__device__ __forceinline__ int dest_of(const ScatterIndex &ix, int scene, int e) {
    int d = __ldg(&ix.idx[(size_t)scene * ix.K + e]);
    if (ix.seg_base) {
        const int sb = __ldg(&ix.seg_base[(size_t)scene * (ix.K / ix.seg_len) + e / ix.seg_len]);
        if (d < 0 || d >= ix.seg_stride || sb < 0 || sb >= ix.n / ix.seg_stride) return -1;  // outside the scan / batch
        d += sb * ix.seg_stride;
    }
    return (d >= 0 && d < ix.n) ? d : -1;
}

__global__ void __launch_bounds__(kBwThreads) bw_count_kernel(ScatterIndex ix) {
    const int scene = blockIdx.y;
    for (int e = blockIdx.x * kBwThreads + threadIdx.x; e < ix.K; e += gridDim.x * kBwThreads) {
        const int d = dest_of(ix, scene, e);
        if (d >= 0) atomicAdd(&ix.cnt[(size_t)scene * ix.n + d], 1);
    }
}

// Exclusive scan of cnt inside blocks of kScanBlock destinations; block totals to blk.
__global__ void __launch_bounds__(kScanBlock) bw_scan_local_kernel(ScatterIndex ix, int T) {
    __shared__ int wsum[32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int g = blockIdx.x * kScanBlock + t;
    const int v = g < T ? ix.cnt[g] : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    int ws = wsum[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, ws, o);
        if (lane >= o) ws += up;
    }
    const int before = warp > 0 ? __shfl_sync(0xffffffffu, ws, warp - 1) : 0;
    if (g < T) ix.off[g] = before + inc - v;
    if (t == kScanBlock - 1) ix.blk[blockIdx.x] = before + inc;
}

// Exclusive scan of the nb block totals in place (one CTA, any nb); blk[nb] = grand total.
__global__ void __launch_bounds__(1024) bw_scan_blocks_kernel(int *blk, int nb) {
    __shared__ int wsum[32];
    __shared__ int carry_s;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        const int i = base + t;
        const int v = i < nb ? blk[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        int ws = wsum[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, ws, o);
            if (lane >= o) ws += up;
        }
        const int before = (warp > 0 ? __shfl_sync(0xffffffffu, ws, warp - 1) : 0) + carry_s;
        if (i < nb) blk[i] = before + inc - v;
        __syncthreads();
        if (t == 1023) carry_s = before + inc;
        __syncthreads();
    }
    if (t == 0) blk[nb] = carry_s;
}

__device__ __forceinline__ int list_start(const ScatterIndex &ix, int g, int T) {
    return g < T ? ix.blk[g / kScanBlock] + ix.off[g] : ix.blk[(T + kScanBlock - 1) / kScanBlock];
}

// Entry lists, in any order inside a destination (the reduce kernel sorts them); cnt counts back to 0.
__global__ void __launch_bounds__(kBwThreads) bw_fill_kernel(ScatterIndex ix) {
    const int scene = blockIdx.y;
    const int T = ix.S * ix.n;
    for (int e = blockIdx.x * kBwThreads + threadIdx.x; e < ix.K; e += gridDim.x * kBwThreads) {
        const int d = dest_of(ix, scene, e);
        if (d < 0) continue;
        const int g = scene * ix.n + d;
        const int pos = atomicSub(&ix.cnt[g], 1) - 1;
        ix.list[list_start(ix, g, T) + pos] = e;
    }
}

struct ScatterData {
    const float *src;  // source rows: element (scene, row, ch) at scene*src_scene + row*src_row + ch*src_ch
    long long src_scene;
    int src_row, src_ch;
    int fan;         // entries per source row: row(e) = e / fan
    const float *w;  // optional (S, K) factor per entry
    float *out;      // destination rows, same addressing
    long long out_scene;
    int out_row, out_ch;
    int c;
};

// In-place ascending bitonic sort of P (power of two) ints in shared memory by one warp.
__device__ __forceinline__ void warp_bitonic(int *a, int P, int lane) {
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < P; i += 32) {
                const int p = i ^ j;
                if (p > i) {
                    const int x = a[i], y = a[p];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) {
                        a[i] = y;
                        a[p] = x;
                    }
                }
            }
            __syncwarp();
        }
    }
}

// The next window of a destination's entry list: the entries in (lo, hi], hi chosen so that at most
// kListCap of them exist, sorted ascending into sl.  Returns their number; *hi_out = hi.
__device__ int build_window(const int *__restrict__ lst, int L, int lo, int remaining, int K, int *sl, int lane,
                            int *hi_out) {
    int hi = INT_MAX;
    if (remaining > kListCap) {  // rare: > 1024 entries on one row.  Largest hi with count((lo, hi]) <= cap
        int a = lo + 1, b = K - 1;
        while (a < b) {
            const int mid = a + (b - a + 1) / 2;
            int cnt = 0;
            for (int i = lane; i < L; i += 32) {
                const int e = lst[i];
                cnt += (e > lo && e <= mid) ? 1 : 0;
            }
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt <= kListCap) a = mid;
            else b = mid - 1;
        }
        hi = a;
    }
    int nw = 0;
    if (L <= 32) {  // common case: rank by shuffles, no sort network
        const int e = lane < L ? lst[lane] : INT_MAX;
        int rank = 0;
        for (int j = 0; j < L; ++j) rank += __shfl_sync(0xffffffffu, e, j) < e ? 1 : 0;
        if (lane < L) sl[rank] = e;
        nw = L;
        __syncwarp();
    } else {
        for (int i0 = 0; i0 < L; i0 += 32) {
            const int i = i0 + lane;
            const int e = i < L ? lst[i] : INT_MAX;
            const bool take = i < L && e > lo && e <= hi;
            const unsigned bal = __ballot_sync(0xffffffffu, take);
            if (take) sl[nw + __popc(bal & ((1u << lane) - 1u))] = e;
            nw += __popc(bal);
        }
        int P = 32;
        while (P < nw) P <<= 1;
        for (int i = nw + lane; i < P; i += 32) sl[i] = INT_MAX;
        __syncwarp();
        warp_bitonic(sl, P, lane);
    }
    *hi_out = hi;
    return nw;
}

template <typename V>
struct Vec;
template <>
struct Vec<float> {
    static __device__ __forceinline__ float zero() { return 0.f; }
    static __device__ __forceinline__ float load(const float *p) { return __ldg(p); }
    static __device__ __forceinline__ void store(float *p, float v) { *p = v; }
    static __device__ __forceinline__ float mul(float a, float w) { return __fmul_rn(a, w); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
};
template <>
struct Vec<float4> {
    static __device__ __forceinline__ float4 zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
    static __device__ __forceinline__ float4 load(const float *p) { return ldg_nc_f4(reinterpret_cast<const float4 *>(p)); }
    static __device__ __forceinline__ void store(float *p, const float4 &v) { stg_cs_f4(reinterpret_cast<float4 *>(p), v); }
    static __device__ __forceinline__ float4 mul(const float4 &a, float w) {
        return make_float4(__fmul_rn(a.x, w), __fmul_rn(a.y, w), __fmul_rn(a.z, w), __fmul_rn(a.w, w));
    }
    static __device__ __forceinline__ float4 add(const float4 &a, const float4 &b) {
        return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
    }
};

// One warp per destination row.  V = float4: channel-last rows (src_ch == out_ch == 1, c % 4 == 0,
// 16-byte aligned); V = float: any strides (xyz rows, channel-first tensors).
// A lane owns ACC columns of the row; D entries' loads are issued together before their D ordered
// adds (the adds of a column are a serial chain by contract, the loads are not): a row with
// hundreds of entries (a three_nn hub in a sparse region: 513 at 16384 -> 4096) is latency bound.
template <typename V, bool WEIGHTED, int ACC, int D>
__global__ void __launch_bounds__(kBwThreads) bw_reduce_kernel(ScatterIndex ix, ScatterData sd) {
    constexpr int W = sizeof(V) / 4;  // channels per lane access
    __shared__ int sl_all[kBwWarps][kListCap];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int *sl = sl_all[warp];
    const int T = ix.S * ix.n;
    const int cv = sd.c / W;  // columns of V per row
    for (int g = blockIdx.x * kBwWarps + warp; g < T; g += gridDim.x * kBwWarps) {
        const int scene = g / ix.n, d = g - scene * ix.n;
        const int start = list_start(ix, g, T), L = list_start(ix, g + 1, T) - start;
        const int *lst = ix.list + start;
        const float *src = sd.src + (size_t)scene * sd.src_scene;
        const float *wts = WEIGHTED ? sd.w + (size_t)scene * ix.K : nullptr;
        float *orow = sd.out + (size_t)scene * sd.out_scene + (size_t)d * sd.out_row;
        const bool single = L <= kListCap;
        int nw = 0, hi = INT_MAX;
        if (single && L > 0) nw = build_window(lst, L, -1, L, ix.K, sl, lane, &hi);
        for (int cb = 0; cb < cv; cb += 32 * ACC) {
            V acc[ACC];
#pragma unroll
            for (int a = 0; a < ACC; ++a) acc[a] = Vec<V>::zero();
            int lo = -1, remaining = L;
            while (remaining > 0) {
                if (!single) {
                    __syncwarp();
                    nw = build_window(lst, L, lo, remaining, ix.K, sl, lane, &hi);
                }
                for (int r0 = 0; r0 < nw; r0 += D) {
                    V v[D][ACC];
                    float wv[D];
#pragma unroll
                    for (int u = 0; u < D; ++u) {
                        if (r0 + u < nw) {  // warp uniform
                            const int e = sl[r0 + u];
                            const float *row = src + (size_t)(e / sd.fan) * sd.src_row;
                            wv[u] = WEIGHTED ? __ldg(&wts[e]) : 1.f;
#pragma unroll
                            for (int a = 0; a < ACC; ++a) {
                                const int col = cb + a * 32 + lane;
                                if (col < cv) v[u][a] = Vec<V>::load(row + (size_t)col * W * sd.src_ch);
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < D; ++u) {
                        if (r0 + u < nw) {
#pragma unroll
                            for (int a = 0; a < ACC; ++a) {
                                const int col = cb + a * 32 + lane;
                                if (col < cv) acc[a] = Vec<V>::add(acc[a], WEIGHTED ? Vec<V>::mul(v[u][a], wv[u]) : v[u][a]);
                            }
                        }
                    }
                }
                remaining -= nw;
                lo = hi;
            }
#pragma unroll
            for (int a = 0; a < ACC; ++a) {
                const int col = cb + a * 32 + lane;
                if (col < cv) Vec<V>::store(orow + (size_t)col * W * sd.out_ch, acc[a]);
            }
        }
        __syncwarp();  // sl is rebuilt for the next destination
    }
}

template <typename V, bool WEIGHTED>
void launch_reduce(const ScatterIndex &ix, const ScatterData &sd, int blocks, cudaStream_t s) {
    const int cv = sd.c / (int)(sizeof(V) / 4);
    if (cv <= 32) bw_reduce_kernel<V, WEIGHTED, 1, 16><<<blocks, kBwThreads, 0, s>>>(ix, sd);
    else if (cv <= 64) bw_reduce_kernel<V, WEIGHTED, 2, 8><<<blocks, kBwThreads, 0, s>>>(ix, sd);
    else bw_reduce_kernel<V, WEIGHTED, 4, 4><<<blocks, kBwThreads, 0, s>>>(ix, sd);
}

inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

struct WsLayout {
    size_t cnt, off, blk, list, total;
    int T, nb;
};
WsLayout ws_layout(int S, int n, int K) {
    WsLayout l;
    l.T = S * n;
    l.nb = (l.T + kScanBlock - 1) / kScanBlock;
    l.cnt = 0;
    l.off = l.cnt + align16((size_t)l.T * 4);
    l.blk = l.off + align16((size_t)l.T * 4);
    l.list = l.blk + align16((size_t)(l.nb + 1) * 4);
    l.total = l.list + align16((size_t)S * K * 4);
    return l;
}

// Build the CSR of `ix` in the workspace and run the reduction described by `sd`.
int scatter_rows(ScatterIndex ix, ScatterData sd, void *workspace, size_t workspace_bytes, cudaStream_t s,
                 const char *what) {
    HF_REQUIRE((long long)ix.S * ix.n < (1ll << 31) && (long long)ix.S * ix.K < (1ll << 31),
               "%s: more than 2^31 destination rows or entries", what);
    const WsLayout l = ws_layout(ix.S, ix.n, ix.K);
    HF_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "%s: null or unaligned workspace", what);
    if (workspace_bytes < l.total) {
        set_error("%s: workspace too small (%zu < %zu bytes)", what, workspace_bytes, l.total);
        return HF_ERR_WORKSPACE;
    }
    unsigned char *w = static_cast<unsigned char *>(workspace);
    ix.cnt = reinterpret_cast<int *>(w + l.cnt);
    ix.off = reinterpret_cast<int *>(w + l.off);
    ix.blk = reinterpret_cast<int *>(w + l.blk);
    ix.list = reinterpret_cast<int *>(w + l.list);
    const int sms = num_sms();
    HF_CUDA_OK(cudaMemsetAsync(ix.cnt, 0, (size_t)l.T * 4, s));
    if (ix.K > 0) {
        int bx = (ix.K + kBwThreads - 1) / kBwThreads;
        const int cap = (sms * 8 + ix.S - 1) / ix.S;
        if (bx > cap) bx = cap;
        bw_count_kernel<<<dim3(bx, ix.S), kBwThreads, 0, s>>>(ix);
        HF_LAUNCH_OK("bw_count_kernel");
        bw_scan_local_kernel<<<l.nb, kScanBlock, 0, s>>>(ix, l.T);
        HF_LAUNCH_OK("bw_scan_local_kernel");
        bw_scan_blocks_kernel<<<1, 1024, 0, s>>>(ix.blk, l.nb);
        HF_LAUNCH_OK("bw_scan_blocks_kernel");
        bw_fill_kernel<<<dim3(bx, ix.S), kBwThreads, 0, s>>>(ix);
        HF_LAUNCH_OK("bw_fill_kernel");
    } else {
        HF_CUDA_OK(cudaMemsetAsync(ix.off, 0, l.list - l.off, s));
    }
    int blocks = (l.T + kBwWarps - 1) / kBwWarps;
    if (blocks > sms * 8) blocks = sms * 8;
    const bool vec = sd.c % 4 == 0 && sd.src_ch == 1 && sd.out_ch == 1 && sd.src_row % 4 == 0 && sd.out_row % 4 == 0 &&
                     sd.src_scene % 4 == 0 && sd.out_scene % 4 == 0 && (reinterpret_cast<uintptr_t>(sd.src) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(sd.out) & 15) == 0;
    if (vec) {
        if (sd.w) launch_reduce<float4, true>(ix, sd, blocks, s);
        else launch_reduce<float4, false>(ix, sd, blocks, s);
    } else {
        if (sd.w) launch_reduce<float, true>(ix, sd, blocks, s);
        else launch_reduce<float, false>(ix, sd, blocks, s);
    }
    HF_LAUNCH_OK("bw_reduce_kernel");
    return HF_OK;
}

}  // namespace
}  // namespace hf

extern "C" size_t hf_scatter_workspace_bytes(int scenes, int dst_rows, int entries) {
    if (scenes <= 0 || dst_rows <= 0 || entries < 0) return 16;
    if ((long long)scenes * dst_rows >= (1ll << 31) || (long long)scenes * entries >= (1ll << 31)) return 0;
    return hf::ws_layout(scenes, dst_rows, entries).total;
}

extern "C" int hf_gather_point_grad(int b, int n, int m, const float *out_g, const int *idx, float *inp_g,
                                    void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1 && m >= 0, "GatherPointGradGpuOp expects (batch_size,num_points,3) inp");
    HF_REQUIRE(inp_g && (m == 0 || (out_g && idx)), "GatherPointGrad: null pointer");
    if (b == 0) return HF_OK;
    ScatterIndex ix = {b, n, m, idx, nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr};
    ScatterData sd = {out_g, (long long)m * 3, 3, 1, 1, nullptr, inp_g, (long long)n * 3, 3, 1, 3};
    return scatter_rows(ix, sd, workspace, workspace_bytes, (cudaStream_t)stream, "GatherPointGrad");
}

extern "C" int hf_group_point_grad(int b, int n, int c, int m, int nsample, const float *grad_out, const int *idx,
                                   float *grad_points, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1 && c >= 1 && m >= 0 && nsample >= 0,
               "GroupPointGrad expects (batch_size, num_points, channel) points shape");
    HF_REQUIRE((long long)m * nsample < (1ll << 31), "GroupPointGrad: npoint*nsample must fit int32");
    HF_REQUIRE(grad_points && ((long long)m * nsample == 0 || (grad_out && idx)), "GroupPointGrad: null pointer");
    if (b == 0) return HF_OK;
    const int K = m * nsample;
    ScatterIndex ix = {b, n, K, idx, nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr};
    ScatterData sd = {grad_out, (long long)K * c, c, 1, 1, nullptr, grad_points, (long long)n * c, c, 1, c};
    return scatter_rows(ix, sd, workspace, workspace_bytes, (cudaStream_t)stream, "GroupPointGrad");
}

extern "C" int hf_three_interpolate_grad(int b, int c, int n, int m, const float *grad_out, const int *idx,
                                         const float *weight, float *grad_points, void *workspace,
                                         size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && c >= 1 && n >= 0 && m >= 1, "ThreeInterpolateGrad expects (b,c,m) points shape");
    HF_REQUIRE((long long)n * 3 < (1ll << 31), "ThreeInterpolateGrad: n*3 must fit int32");
    HF_REQUIRE(grad_points && (n == 0 || (grad_out && idx && weight)), "ThreeInterpolateGrad: null pointer");
    if (b == 0) return HF_OK;
    ScatterIndex ix = {b, m, n * 3, idx, nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr};
    ScatterData sd = {grad_out, (long long)c * n, 1, n, 3, weight, grad_points, (long long)c * m, 1, m, c};  // (b,c,n) -> (b,c,m)
    return scatter_rows(ix, sd, workspace, workspace_bytes, (cudaStream_t)stream, "ThreeInterpolateGrad");
}

extern "C" int hf_three_interpolate_grad_nlc(int b, int n, int c, int m, const float *grad_out, const int *idx,
                                             const float *weight, float *grad_points, void *workspace,
                                             size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && c >= 1 && n >= 0 && m >= 1, "ThreeInterpolateGrad expects (b,m,c) points shape");
    HF_REQUIRE((long long)n * 3 < (1ll << 31), "ThreeInterpolateGrad: n*3 must fit int32");
    HF_REQUIRE(grad_points && (n == 0 || (grad_out && idx && weight)), "ThreeInterpolateGrad: null pointer");
    if (b == 0) return HF_OK;
    ScatterIndex ix = {b, m, n * 3, idx, nullptr, 1, 0, nullptr, nullptr, nullptr, nullptr};
    ScatterData sd = {grad_out, (long long)n * c, c, 1, 3, weight, grad_points, (long long)m * c, c, 1, c};  // (b,n,c) -> (b,m,c)
    return scatter_rows(ix, sd, workspace, workspace_bytes, (cudaStream_t)stream, "ThreeInterpolateGrad");
}

extern "C" int hf_pc_crop_and_sample_grad_fts(const int *box_ind, const int *crop_ind, const float *grad_crop_fts,
                                              int num_boxes, int batch, int npts, int resize, int channel,
                                              float *grad_fts, void *workspace, size_t workspace_bytes,
                                              hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num_boxes >= 0 && batch >= 1 && npts >= 1 && resize >= 1 && channel >= 1,
               "PcCropAndSampleGradFts expects fts and grad_crop_fts has same shape(2)");
    HF_REQUIRE((long long)num_boxes * resize < (1ll << 31) && (long long)batch * npts < (1ll << 31),
               "PcCropAndSampleGradFts: num_boxes*resize and batch*npts must fit int32");
    HF_REQUIRE(grad_fts && (num_boxes == 0 || (box_ind && crop_ind && grad_crop_fts)), "PcCropAndSampleGradFts: null pointer");
    // one flat "scene": destination row = box_ind[box] * npts + crop_ind[box, slot]
    ScatterIndex ix = {1, batch * npts, num_boxes * resize, crop_ind, box_ind, resize, npts, nullptr, nullptr, nullptr, nullptr};
    ScatterData sd = {grad_crop_fts, 0, channel, 1, 1, nullptr, grad_fts, 0, channel, 1, channel};
    return scatter_rows(ix, sd, workspace, workspace_bytes, (cudaStream_t)stream, "PcCropAndSampleGradFts");
}
