// common.cuh — shared host/device helpers for libhfops (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hfops.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libhfops is written for sm_100a (B200) only"
#endif

namespace hf {

// ---- host side: status / error string / launch counter -------------------
void set_error(const char *fmt, ...);
void count_launch(unsigned n = 1);

#define HF_REQUIRE(cond, ...)              \
    do {                                   \
        if (!(cond)) {                     \
            hf::set_error(__VA_ARGS__);    \
            return HF_ERR_INVALID_ARG;     \
        }                                  \
    } while (0)

#define HF_CUDA_OK(expr)                                                              \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            hf::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                          __FILE__, __LINE__);                                        \
            return HF_ERR_CUDA;                                                       \
        }                                                                             \
    } while (0)

// Check the launch that was just issued.
#define HF_LAUNCH_OK(name)                                                            \
    do {                                                                              \
        cudaError_t _e = cudaGetLastError();                                          \
        if (_e != cudaSuccess) {                                                      \
            hf::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));   \
            return HF_ERR_CUDA;                                                       \
        }                                                                             \
        hf::count_launch();                                                           \
    } while (0)

// SM count of the CURRENT device, cached per device ordinal (a process may drive several device types).
inline int num_sms() {
    constexpr int kMaxDev = 64;
    static int cache[kMaxDev] = {0};  // 0 = unknown; racing first calls write the same value
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev >= 0 && dev < kMaxDev && cache[dev] > 0) return cache[dev];
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    if (dev >= 0 && dev < kMaxDev) cache[dev] = sms;
    return sms;
}

// Read-only view of a cloud index (hf_cloud_index_build) for the kernels outside spatial.cu that consume one.  The
// layout has ONE owner, spatial.cu (carve_index / super_count / kMaxIndexed): consumers ask for the view instead of
// re-deriving offsets.  Returns false (and leaves *view alone) when `index` is null or not 16-byte aligned, or when a
// cloud of (b, n) cannot be indexed at all; sorted/boxes/super are float4 records as documented in spatial.cu.
struct CloudIndexView {
    const float4 *sorted;  // (b, npad) {x, y, z, original index as int bits}; padding = {NaN, NaN, NaN, -1}
    const float4 *boxes;   // (b, npad / 32, 2) bucket boxes {lo.xyz, smallest original index}, {hi.xyz, -}
    const float4 *super;   // (b, nsuper, 2) boxes of 32 consecutive buckets; nullptr when nsuper == 0
    int npad, nsuper;
};
bool cloud_index_view(int b, int n, const void *index, CloudIndexView *view);

#ifdef __CUDACC__
// ---- device side ----------------------------------------------------------

// The squared distance every index-producing op shares.  Order fixed to what nvcc
// 12.9 emits for the reference source on sm_100a (FMUL on dy, then FFMA dx, FFMA dz),
// written with explicit intrinsics so no compiler choice can change the last bit.
__device__ __forceinline__ float sqdist(float dx, float dy, float dz) {
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// Inverse-distance weights of the three nearest neighbours (hf/core/feature_extractors/
// pointnet_util.py:304-307): d = max(d, 1e-10); w_k = (1/d_k) / sum_k (1/d_k).
__device__ __forceinline__ void fp_weights3(float d1, float d2, float d3, float *w) {
    const float i1 = __fdiv_rn(1.f, fmaxf(d1, 1e-10f)), i2 = __fdiv_rn(1.f, fmaxf(d2, 1e-10f)),
                i3 = __fdiv_rn(1.f, fmaxf(d3, 1e-10f));
    const float norm = __fadd_rn(__fadd_rn(i1, i2), i3);
    w[0] = __fdiv_rn(i1, norm);
    w[1] = __fdiv_rn(i2, norm);
    w[2] = __fdiv_rn(i3, norm);
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier (shared::cta) ----
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// ---- TMA 1-D bulk copy global -> shared (SASS: UBLKCP), completes on an mbarrier ----
// dst and src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- TMA 1-D bulk copy shared -> global (SASS: UBLKCP.G.S), tracked by the thread's bulk async-group ----
__device__ __forceinline__ void tma_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// Wait until at most N of this thread's committed bulk groups still READ their shared-memory source.
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// Order generic-proxy writes to shared memory (st.shared) before a following bulk store reads them.
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- streaming 128-bit global access ----
__device__ __forceinline__ float4 ldg_nc_f4(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_cs_f4(float4 *p, const float4 &v) {
    asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
#endif  // __CUDACC__

}  // namespace hf
