// api.cu — library-wide state of libhfops: version, per-thread last error, launch counter.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace hf {
namespace {
thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};
}  // namespace

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace hf

#ifndef HF_BUILD_ID
#define HF_BUILD_ID "unknown"
#endif
extern "C" const char *hf_version(void) { return "hfops-b200 0.2.0 (sm_100a) build " HF_BUILD_ID; }
extern "C" const char *hf_last_error(void) { return hf::g_err; }
extern "C" unsigned long long hf_launch_count(void) { return hf::g_launches.load(std::memory_order_relaxed); }

// ---- measurement aid: the FP32 (non-tensor) peak of the device, for roofline denominators ----
// Every thread runs 8 independent FMA chains (16 with packed = 1: fma.rn.f32x2, two FMAs per instruction),
// `iters` times; nothing but FFMA issues in the loop.  The caller times the launch with CUDA events and divides
// hf_fp32_probe_flops() by the duration.  The sums are written so the loop cannot be removed.
namespace hf {
namespace {
constexpr int kProbeThreads = 1024;
constexpr int kProbeChains = 8;

template <bool PACKED>
__global__ void __launch_bounds__(kProbeThreads) fp32_probe_kernel(int iters, float seed, float *out) {
    const float a = seed, b = 1.f - seed * seed;
    if (PACKED) {
        unsigned long long acc[kProbeChains];
#pragma unroll
        for (int k = 0; k < kProbeChains; ++k) {
            const float v = (float)(threadIdx.x + k);
            asm("mov.b64 %0, {%1, %2};" : "=l"(acc[k]) : "f"(v), "f"(v + 0.5f));
        }
        unsigned long long aa, bb;
        asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
        asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < kProbeChains; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[k]) : "l"(aa), "l"(bb));
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kProbeChains; ++k) {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[k]));
            s += lo + hi;
        }
        out[(size_t)blockIdx.x * kProbeThreads + threadIdx.x] = s;
    } else {
        float acc[kProbeChains];
#pragma unroll
        for (int k = 0; k < kProbeChains; ++k) acc[k] = (float)(threadIdx.x + k);
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < kProbeChains; ++k) acc[k] = __fmaf_rn(acc[k], a, b);
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kProbeChains; ++k) s += acc[k];
        out[(size_t)blockIdx.x * kProbeThreads + threadIdx.x] = s;
    }
}
}  // namespace
}  // namespace hf

extern "C" size_t hf_fp32_probe_out_floats(int blocks) { return (size_t)(blocks > 0 ? blocks : 0) * hf::kProbeThreads; }
extern "C" double hf_fp32_probe_flops(int blocks, int iters, int packed) {
    return 2.0 * (packed ? 2.0 : 1.0) * hf::kProbeChains * (double)hf::kProbeThreads * (double)blocks * (double)iters;
}
extern "C" int hf_fp32_probe(int blocks, int iters, int packed, float *out, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(blocks > 0 && iters > 0 && out, "fp32 probe: blocks, iters > 0 and an output buffer are required");
    if (packed) fp32_probe_kernel<true><<<blocks, kProbeThreads, 0, (cudaStream_t)stream>>>(iters, 0.999f, out);
    else fp32_probe_kernel<false><<<blocks, kProbeThreads, 0, (cudaStream_t)stream>>>(iters, 0.999f, out);
    HF_LAUNCH_OK("fp32_probe_kernel");
    return HF_OK;
}
