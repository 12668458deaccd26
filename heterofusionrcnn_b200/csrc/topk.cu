// topk.cu — select_top_k (the reference's SelectionSort op) for sm_100a.
//
// Replaces selectionSortLauncher / selection_sort_gpu, grouping/tf_grouping_g.cu:83-123,129 (see
// include/hfops.h).  The reference gives every row of the (b,m,n) distance matrix to ONE thread, which
// runs k steps of selection sort over it (n*k serial compares per row, 16 blocks for the whole batch).
// The full outputs — including the tail, which keeps whatever permutation the k swaps leave behind —
// are part of the op's contract (tf_grouping.py:29-39), so the algorithm is kept and parallelised:
// one CTA per row, the row (values and indices) lives in shared memory, each of the k steps is a
// block-wide arg-min over positions s..n-1 ordered by (value, position) — exactly what the
// reference's ascending scan with a strict '<' selects — followed by the same swap.
#include <stdint.h>

#include "common.cuh"

namespace hf {
namespace {

constexpr int kTkThreads = 256;

__global__ void __launch_bounds__(kTkThreads)
    select_top_k_kernel(int n, int k, const float *__restrict__ dist, int *__restrict__ outi, float *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char tk_smem[];
    float *v = reinterpret_cast<float *>(tk_smem);  // [n]
    int *ix = reinterpret_cast<int *>(v + n);        // [n]
    __shared__ float wv[kTkThreads / 32];
    __shared__ int wp[kTkThreads / 32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const size_t row = (size_t)blockIdx.x * n;
    for (int i = t; i < n; i += kTkThreads) {
        v[i] = __ldg(&dist[row + i]);
        ix[i] = i;
    }
    __syncthreads();
    const int steps = k < n ? k : n;
    for (int s = 0; s < steps; ++s) {
        // first position in [s, n) holding the minimum: (value, position) lexicographic, NaN never wins
        float bv = INFINITY;
        int bp = INT_MAX;
        for (int i = s + t; i < n; i += kTkThreads) {
            const float x = v[i];
            if (x < bv || (x == bv && i < bp)) {
                bv = x;
                bp = i;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int op = __shfl_xor_sync(0xffffffffu, bp, o);
            if (ov < bv || (ov == bv && op < bp)) {
                bv = ov;
                bp = op;
            }
        }
        if (lane == 0) {
            wv[warp] = bv;
            wp[warp] = bp;
        }
        __syncthreads();
        if (t == 0) {
            for (int w = 1; w < kTkThreads / 32; ++w)
                if (wv[w] < bv || (wv[w] == bv && wp[w] < bp)) {
                    bv = wv[w];
                    bp = wp[w];
                }
            // the reference starts from min = s and only moves on a strictly smaller value: a NaN at s stays
            const float cur = v[s];
            if (bp != INT_MAX && bp != s && !(cur != cur)) {
                v[bp] = cur;
                v[s] = bv;
                const int ti = ix[bp];
                ix[bp] = ix[s];
                ix[s] = ti;
            }
        }
        __syncthreads();
    }
    for (int i = t; i < n; i += kTkThreads) {
        out[row + i] = v[i];
        outi[row + i] = ix[i];
    }
}

}  // namespace
}  // namespace hf

extern "C" int hf_select_top_k(int b, int n, int m, int k, const float *dist, int *outi, float *out, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(k > 0, "SelectionSort expects positive k");
    HF_REQUIRE(b >= 0 && m >= 0 && n >= 1, "SelectionSort expects (b,m,n) dist shape.");
    HF_REQUIRE(dist && outi && out, "SelectionSort: null pointer");
    const long long rows = (long long)b * m;
    HF_REQUIRE(rows < (1ll << 31), "SelectionSort: b*m must fit int32");
    if (rows == 0) return HF_OK;
    const size_t smem = (size_t)n * 8;
    if (smem > 200 * 1024) {
        set_error("SelectionSort: rows of more than %d elements are not supported", 200 * 1024 / 8);
        return HF_ERR_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
        HF_CUDA_OK(cudaFuncSetAttribute(select_top_k_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    select_top_k_kernel<<<(unsigned)rows, kTkThreads, smem, (cudaStream_t)stream>>>(n, k, dist, outi, out);
    HF_LAUNCH_OK("select_top_k_kernel");
    return HF_OK;
}
