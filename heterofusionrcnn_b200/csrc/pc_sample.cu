// pc_sample.cu — the fixed-size point sampling of the KITTI loader on the device.
//
// Replaces hf/datasets/kitti/kitti_dataset.py:341-371 of the reference: from the n points hf_lidar_fov_filter kept,
// choose pc_sample_pts = P rows — every point at depth >= 40 m plus a random subset of the nearer ones (n > P), or
// every point plus random repeats (n < P) — shuffle them, and shift the intensity to [-0.5, 0.5].
//
// The reference draws with numpy's global RNG, whose stream cannot be reproduced on a GPU and which needs the
// near / far counts before it can draw.  So the randomness is an INPUT, drawn by the host without knowing anything
// about the scan, and everything else runs here, so the scan never goes back to the host between
// hf_lidar_fov_filter and farthest_point_sample:
//   keys    (S, n_cap) uint32   np.random.choice(a, k, replace=False)  := the k candidates with the smallest
//                               (keys[i], i), returned in ascending index order
//   extra   (S, P) uint32       np.random.choice(a, k, replace=True)   := a[extra[e] % len(a)], e = 0..k-1
//   slot_of (S, P) int32        np.random.shuffle(choice)              := new[slot_of[j]] = old[j]  (a permutation)
// Each is a valid implementation of the numpy call it stands for (uniform keys / a uniform permutation give the
// same distribution); tests/golden/make_golden_pc_sample.py executes the reference's own lines with np.random
// replaced by exactly these definitions; this kernel reproduces that output bit for bit (tests/test_gpu_pc_sample.py).
//
// One CTA per scan.  Selecting the k smallest keys is a 4-pass radix select (256-bin shared-memory histograms);
// "ascending index order" makes every position a prefix count, so there is no sort: three bitmaps by point index
// (candidate, key < T, key == T) built with ballots, their popcount prefixes by one block scan, and one output pass.
#include "common.cuh"

namespace hf {
namespace {

constexpr int kPsThreads = 1024;
constexpr int kPsMaxPoints = 131072;  // 3 bitmaps + 3 prefix arrays of n/32 words = 96 KB of shared memory

// exclusive prefix sums of popc(bits[w]) over `words` words -> pref[w]; returns the total (to every thread)
__device__ unsigned popc_prefix(const unsigned *bits, unsigned *pref, int words, unsigned *warp_tot) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (words + kPsThreads - 1) / kPsThreads;  // consecutive words per thread
    const int w0 = t * per;
    unsigned mine = 0;
    for (int k = 0; k < per; ++k)
        if (w0 + k < words) mine += __popc(bits[w0 + k]);
    unsigned inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    __syncthreads();  // warp_tot may still be read by the previous call
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    unsigned base = 0, total = 0;
    for (int k = 0; k < kPsThreads / 32; ++k) {
        const unsigned v = warp_tot[k];
        if (k < warp) base += v;
        total += v;
    }
    unsigned run = base + inc - mine;
    for (int k = 0; k < per; ++k)
        if (w0 + k < words) {
            pref[w0 + k] = run;
            run += __popc(bits[w0 + k]);
        }
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(kPsThreads)
    pc_sample_kernel(int n_cap, int P, const float4 *__restrict__ pts, const int *__restrict__ counts,
                     const unsigned *__restrict__ keys, const int *__restrict__ slot_of, const unsigned *__restrict__ extra,
                     float split, float shift, float4 *__restrict__ out, int *__restrict__ choice, int *__restrict__ status) {
    extern __shared__ unsigned sm[];
    __shared__ unsigned hist[256];
    __shared__ unsigned warp_tot[kPsThreads / 32];
    __shared__ unsigned s_prefix, s_kk;
    const int t = threadIdx.x, lane = t & 31;
    const int scene = blockIdx.x;
    const int words_cap = (n_cap + 31) >> 5;
    unsigned *wcand = sm, *wlt = sm + words_cap, *weq = sm + 2 * words_cap;
    unsigned *pfar = sm + 3 * words_cap, *plt = sm + 4 * words_cap, *peq = sm + 5 * words_cap;

    int n = counts[scene];
    n = n < 0 ? 0 : (n > n_cap ? n_cap : n);
    const int words = (n + 31) >> 5;
    pts += (size_t)scene * n_cap;
    keys += (size_t)scene * n_cap;
    slot_of += (size_t)scene * P;
    out += (size_t)scene * P;
    if (choice) choice += (size_t)scene * P;
    if (extra) extra += (size_t)scene * P;

    auto fail = [&]() {  // the reference raises here (np.random.choice with a negative size / from an empty array)
        for (int j = t; j < P; j += kPsThreads) {
            out[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (choice) choice[j] = -1;
        }
        if (t == 0) status[scene] = 0;
    };
    bool bad_slot = false;  // this thread met a slot_of entry outside [0, P)
    auto emit = [&](int j, int i) {  // pre-shuffle position j holds point i
        const int slot = slot_of[j];
        if (slot < 0 || slot >= P) {  // not a permutation: nothing is written through it, the scan is failed below
            bad_slot = true;
            return;
        }
        float4 p = pts[i];
        p.w = __fsub_rn(p.w, shift);
        out[slot] = p;
        if (choice) choice[slot] = i;
    };
    if (P <= 0) {
        if (t == 0) status[scene] = 1;
        return;
    }
    if (n == 0) {
        fail();
        return;
    }
    const bool caseA = P < n;  // more points than wanted: all far points + a draw from the near ones
    // ---- candidates of the draw without replacement: near points (case A) / every point (case B) ----
    for (int base = 0; base < words * 32; base += kPsThreads) {
        const int i = base + t;
        const bool valid = i < n;
        bool c = valid;
        if (caseA && valid) c = pts[i].z < split;  // pts_near_flag = pts_depth < 40.0 (NaN depth counts as far)
        const unsigned b = __ballot_sync(0xffffffffu, c);
        if (lane == 0 && (i >> 5) < words) wcand[i >> 5] = b;  // the last round of the loop runs past `words`
    }
    __syncthreads();
    // pfar <- prefix of the far bitmap (valid & ~cand); borrow wlt as its scratch
    for (int w = t; w < words; w += kPsThreads) {
        const unsigned validmask = (w == words - 1 && (n & 31)) ? ((1u << (n & 31)) - 1u) : 0xffffffffu;
        wlt[w] = validmask & ~wcand[w];
    }
    __syncthreads();
    const int n_far = (int)popc_prefix(wlt, pfar, words, warp_tot);
    const int n_cand = n - n_far;
    int k;  // size of the draw without replacement
    bool with_replacement = false;
    if (caseA) {
        k = P - n_far;
        if (k < 0) {
            fail();
            return;
        }
    } else {
        k = P - n;
        if (P > 2 * n) {  // replace=True (kitti_dataset.py:357-361)
            with_replacement = true;
            k = 0;
        }
    }
    // ---- radix select: T = the k-th smallest candidate key, r = how many of the keys equal to T are taken ----
    unsigned T = 0, r = 0;
    if (k > 0) {
        if (t == 0) {
            s_prefix = 0u;
            s_kk = (unsigned)k;
        }
        for (int pass = 0; pass < 4; ++pass) {
            const int sh = 24 - 8 * pass;
            for (int bn = t; bn < 256; bn += kPsThreads) hist[bn] = 0u;
            __syncthreads();
            const unsigned prefix = s_prefix;
            for (int i = t; i < n; i += kPsThreads) {
                if (!((wcand[i >> 5] >> (i & 31)) & 1u)) continue;
                const unsigned key = keys[i];
                if (pass == 0 || (key >> (sh + 8)) == (prefix >> (sh + 8))) atomicAdd(&hist[(key >> sh) & 255u], 1u);
            }
            __syncthreads();
            if (t == 0) {
                unsigned cum = 0, kk = s_kk;
                for (int bn = 0; bn < 256; ++bn) {
                    if (cum + hist[bn] >= kk) {
                        s_prefix = prefix | ((unsigned)bn << sh);
                        s_kk = kk - cum;
                        break;
                    }
                    cum += hist[bn];
                }
            }
            __syncthreads();
        }
        T = s_prefix;
        r = s_kk;
    }
    // ---- bitmaps key < T / key == T over the candidates, and their prefixes ----
    for (int base = 0; base < words * 32; base += kPsThreads) {
        const int i = base + t;
        bool lt = false, eq = false;
        if (k > 0 && i < n && ((wcand[i >> 5] >> (i & 31)) & 1u)) {
            const unsigned key = keys[i];
            lt = key < T;
            eq = key == T;
        }
        const unsigned bl = __ballot_sync(0xffffffffu, lt), be = __ballot_sync(0xffffffffu, eq);
        if (lane == 0 && (i >> 5) < words) {
            wlt[i >> 5] = bl;
            weq[i >> 5] = be;
        }
    }
    __syncthreads();
    popc_prefix(wlt, plt, words, warp_tot);
    popc_prefix(weq, peq, words, warp_tot);
    // ---- output: position before the shuffle -> slot_of -> row ----
    for (int i = t; i < n; i += kPsThreads) {
        const int w = i >> 5;
        const unsigned bit = 1u << (i & 31), below = bit - 1u;
        const unsigned eqr = peq[w] + __popc(weq[w] & below);
        const bool sel = (wlt[w] & bit) || ((weq[w] & bit) && eqr < r);
        const int selrank = (int)(plt[w] + __popc(wlt[w] & below) + (eqr < r ? eqr : r));
        if (caseA) {
            if (sel) {
                emit(selrank, i);  // near_idxs_choice, ascending index
            } else if (!(wcand[w] & bit)) {
                const unsigned validmask = (w == words - 1 && (n & 31)) ? ((1u << (n & 31)) - 1u) : 0xffffffffu;
                emit(k + (int)(pfar[w] + __popc(validmask & ~wcand[w] & below)), i);  // far_idxs_choice
            }
        } else {
            emit(i, i);                       // choice = np.arange(n)
            if (sel) emit(n + selrank, i);    // extra_choice, replace=False
        }
    }
    if (with_replacement) {
        if (!extra) {
            __syncthreads();
            fail();
            return;
        }
        for (int e = t; e < P - n; e += kPsThreads) emit(n + e, (int)(extra[e] % (unsigned)n));
    }
    (void)n_cand;
    // An out-of-range slot_of entry would leave rows of `out` unwritten: report the scan like the reference's own
    // failures (status 0, rows zero).  The barrier also orders every emit() before fail()'s rewrite of the rows.
    // (Duplicate in-range slots are the caller's precondition and are not detected.)
    if (__syncthreads_or(bad_slot ? 1 : 0)) {
        fail();
        return;
    }
    if (t == 0) status[scene] = 1;
}

}  // namespace
}  // namespace hf

extern "C" int hf_pc_sample_points(int scans, int n_cap, int num_out, const float *pts, const int *counts,
                                   const unsigned *keys, const int *slot_of, const unsigned *extra, float depth_split,
                                   float intensity_shift, float *out, int *choice, int *status, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(scans >= 0 && n_cap >= 1 && num_out >= 0, "pc_sample_points expects (S,n,4) points and a sample size >= 0");
    HF_REQUIRE(n_cap <= kPsMaxPoints, "pc_sample_points: at most %d points per scan", kPsMaxPoints);
    if (scans == 0) return HF_OK;
    HF_REQUIRE(pts && counts && keys && status && (num_out == 0 || (slot_of && out)), "pc_sample_points: null pointer");
    HF_REQUIRE((reinterpret_cast<uintptr_t>(pts) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
               "pc_sample_points: pts and out must be 16-byte aligned");
    const size_t smem = (size_t)6 * ((n_cap + 31) / 32) * sizeof(unsigned);
    if (smem > 40 * 1024)
        HF_CUDA_OK(cudaFuncSetAttribute(pc_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pc_sample_kernel<<<scans, kPsThreads, smem, (cudaStream_t)stream>>>(
        n_cap, num_out, reinterpret_cast<const float4 *>(pts), counts, keys, slot_of, extra, depth_split, intensity_shift,
        reinterpret_cast<float4 *>(out), choice, status);
    HF_LAUNCH_OK("pc_sample_kernel");
    return HF_OK;
}
