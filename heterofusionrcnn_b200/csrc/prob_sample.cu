// prob_sample.cu — weighted sampling by inverse CDF (the reference's ProbSample op) for sm_100a.
//
// Replaces probsampleLauncher / cumsumLauncher, sampling/tf_sampling_g.cu:7-104,194-201 (see include/hfops.h).
// The op's output is the position a binary search finds in a float32 prefix sum of the weights, so the prefix
// sum's rounding — i.e. its association order — is part of the contract: 8192-element chunks, quads
// (a, a+b, c+(a+b), (c+d)+(a+b)), a work-efficient up-sweep / down-sweep scan of the quad totals, the running sum
// carried between chunks with a compensated two-float add.  The tree is reproduced node for node (float add is
// commutative, so only the pairing matters); what differs is the machine mapping: one 1024-thread CTA per row
// with the quads in registers (the reference runs 32 CTAs of 512 threads over b rows through two shared-memory
// arrays), and the search is fused behind the scan of its row in the same kernel, so a call is ONE launch and
// the prefix sum is read back from L2 / shared memory, not from a second kernel.
#include <stdint.h>

#include "common.cuh"

namespace hf {
namespace {

constexpr int kPsThreads = 1024;
constexpr int kPsChunk = 8192;             // elements per chunk
constexpr int kPsQuads = kPsChunk / 4;     // 2048 quad totals per chunk, two per thread

__global__ void __launch_bounds__(kPsThreads)
    prob_sample_kernel(int n, int m, const float *__restrict__ inp_p, const float *__restrict__ inp_r,
                       float *__restrict__ temp, int *__restrict__ out) {
    __shared__ float tot[kPsQuads];
    __shared__ float carry[2];  // running sum and its compensation term
    const int t = threadIdx.x;
    const int row = blockIdx.x;
    const float *src = inp_p + (size_t)row * n;
    float *cs = temp + (size_t)row * n;
    if (t == 0) carry[0] = carry[1] = 0.f;
    __syncthreads();
    for (int j = 0; j < n; j += kPsChunk) {
        const int ni = min(n - j, kPsChunk), n2 = ((ni + 3) & ~3) >> 2;
        float q[2][4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int quad = t + h * kPsThreads, k = quad * 4;
            if (quad < n2) {
                if (k + 3 < ni) {
                    const float a = src[j + k], b = src[j + k + 1], c = src[j + k + 2], d = src[j + k + 3];
                    const float ab = __fadd_rn(b, a), cd = __fadd_rn(d, c);
                    q[h][0] = a;
                    q[h][1] = ab;
                    q[h][2] = __fadd_rn(c, ab);
                    q[h][3] = __fadd_rn(cd, ab);
                    tot[quad] = q[h][3];
                } else {  // the ragged last quad of the row: a plain running sum
                    float v = 0.f;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (k + e < ni) v = __fadd_rn(v, src[j + k + e]);
                        q[h][e] = v;
                    }
                    tot[quad] = v;
                }
            }
        }
        // up-sweep, then down-sweep: the reference's node pairs and bounds (tf_sampling_g.cu:46-65)
        int u = 0;
        for (; (2 << u) <= n2; ++u) {
            __syncthreads();
            for (int k = t; k < (n2 >> (u + 1)); k += kPsThreads) {
                const int i1 = (((k << 1) + 2) << u) - 1, i2 = (((k << 1) + 1) << u) - 1;
                tot[i1] = __fadd_rn(tot[i1], tot[i2]);
            }
        }
        for (--u; u >= 0; --u) {
            __syncthreads();
            for (int k = t; k < ((n2 - (1 << u)) >> (u + 1)); k += kPsThreads) {
                const int i1 = (((k << 1) + 3) << u) - 1, i2 = (((k << 1) + 2) << u) - 1;
                tot[i1] = __fadd_rn(tot[i1], tot[i2]);
            }
        }
        __syncthreads();
        const float running = carry[0];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int quad = t + h * kPsThreads, k = quad * 4;
            if (quad < n2) {
                const float before = quad ? tot[quad - 1] : 0.f;
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (k + e < ni) cs[j + k + e] = __fadd_rn(quad ? __fadd_rn(q[h][e], before) : q[h][e], running);
            }
        }
        __syncthreads();
        if (t == 0) {
            const float tt = __fadd_rn(tot[n2 - 1], carry[1]);
            const float r2 = __fadd_rn(running, tt);
            carry[1] = __fsub_rn(tt, __fsub_rn(r2, running));
            carry[0] = r2;
        }
        __syncthreads();
    }
    if (m == 0) return;
    __threadfence_block();  // the row's prefix sum, written above by this CTA, is searched below
    __syncthreads();
    int base = 1;
    while (base < n) base <<= 1;
    const float total = cs[n - 1];
    for (int s = t; s < m; s += kPsThreads) {
        const float q = __fmul_rn(inp_r[(size_t)row * m + s], total);
        int r = n - 1;
        for (int k = base; k >= 1; k >>= 1)
            if (r >= k && cs[r - k] >= q) r -= k;
        out[(size_t)row * m + s] = r;
    }
}

}  // namespace
}  // namespace hf

extern "C" int hf_prob_sample(int b, int n, int m, const float *inp_p, const float *inp_r, float *temp, int *out,
                              hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(b >= 0 && n >= 1, "ProbSample expects (batch_size,num_choices) inp shape");
    HF_REQUIRE(m >= 0, "ProbSample expects (batch_size,num_points) inpr shape");
    HF_REQUIRE(inp_p && temp && (m == 0 || (inp_r && out)), "ProbSample: null pointer");
    if (b == 0) return HF_OK;
    prob_sample_kernel<<<b, kPsThreads, 0, (cudaStream_t)stream>>>(n, m, inp_p, inp_r, temp, out);
    HF_LAUNCH_OK("prob_sample_kernel");
    return HF_OK;
}
