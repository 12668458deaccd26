// fps_lab.cu — GPU micro-benchmark of FPS exchange designs (not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/fps_lab tools/fps_lab.cu && /tmp/fps_lab
// Variants (all must produce identical indices):
//   A  library kernel (owner warp pushes, one __syncthreads + st.async/mbarrier)
//   B  every warp pushes its own candidate to every CTA (no __syncthreads in the loop)
//   C  single CTA, every warp writes its candidate to shared memory, one __syncthreads
//   D  single warp per scene (no barrier at all), n <= 1024
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../heterofusionrcnn_b200/csrc/api.cu"
#include "../heterofusionrcnn_b200/csrc/sampling.cu"

namespace hf {
namespace {

// ---- B: every warp pushes ------------------------------------------------------------
template <int CLUSTER, int THREADS, int PTS>
__global__ void __launch_bounds__(THREADS, 1)
    fps_warp_push_kernel(int n, int m, const float *__restrict__ inp, int *__restrict__ out) {
    constexpr int CAP = THREADS * PTS;
    constexpr int NW = THREADS / 32;
    constexpr int NSLOT = CLUSTER * NW;  // <= 64 handled (2 per lane)
    static_assert(NSLOT <= 64, "");
    __shared__ __align__(16) uint32_t slots[2][64][8];
    __shared__ __align__(8) uint64_t bars[2];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t rank = CLUSTER > 1 ? cluster_ctarank() : 0u;
    const int scene = blockIdx.x / CLUSTER;
    const float *data = inp + (size_t)scene * n * 3;
    int *idxs = out + (size_t)scene * m;
    const uint32_t base = rank * CAP;

    float x[PTS], y[PTS], z[PTS], td[PTS];
    uint32_t tw[PTS];
#pragma unroll
    for (int p = 0; p < PTS; ++p) {
        uint32_t k = base + p * THREADS + t;
        tw[p] = tie_word(k);
        if (k < (uint32_t)n) {
            x[p] = data[(size_t)k * 3 + 0];
            y[p] = data[(size_t)k * 3 + 1];
            z[p] = data[(size_t)k * 3 + 2];
            td[p] = 1e38f;
        } else {
            x[p] = y[p] = z[p] = 0.f;
            td[p] = -1.f;
            tw[p] = 0u;
        }
    }
    for (int i = t; i < 2 * 64 * 8; i += THREADS) (&slots[0][0][0])[i] = 0u;
    if (t == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&bars[0], NSLOT * kMsgBytes);
        mbar_arrive_expect_tx(&bars[1], NSLOT * kMsgBytes);
    }
    if (CLUSTER > 1) cluster_sync_all();
    else __syncthreads();

    float cx = data[0], cy = data[1], cz = data[2];
    if (rank == 0 && t == 0) idxs[0] = 0;
    const uint32_t myslot = rank * NW + warp;
    const int iters = m - 1;
    for (int it = 0; it < iters; ++it) {
        const int buf = it & 1;
        float mx = -1.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            float d = sqdist(x[p] - cx, y[p] - cy, z[p] - cz);
            td[p] = fminf(d, td[p]);
            mx = fmaxf(mx, td[p]);
        }
        const uint32_t hi = __float_as_uint(fmaxf(mx, 0.f));
        const uint32_t wm = __reduce_max_sync(0xffffffffu, hi);
        // branch-free recovery of this lane's best point
        uint32_t lo = 0u;
        float bx = 0.f, by = 0.f, bz = 0.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            bool better = td[p] == mx && tw[p] > lo;
            lo = better ? tw[p] : lo;
            bx = better ? x[p] : bx;
            by = better ? y[p] : by;
            bz = better ? z[p] : bz;
        }
        lo = hi == wm ? lo : 0u;
        const uint32_t wl = __reduce_max_sync(0xffffffffu, lo);
        // the winning lane (unique when wl != 0) hands its coordinates to lanes 0..CLUSTER-1
        const unsigned who = __ballot_sync(0xffffffffu, lo == wl);
        const int src = __ffs(who) - 1;
        const float wx = __shfl_sync(0xffffffffu, bx, src);
        const float wy = __shfl_sync(0xffffffffu, by, src);
        const float wz = __shfl_sync(0xffffffffu, bz, src);
        if (lane < CLUSTER) {
            uint32_t dst = mapa_u32(smem_u32(&slots[buf][myslot][0]), (uint32_t)lane);
            uint32_t dbar = mapa_u32(smem_u32(&bars[buf]), (uint32_t)lane);
            st_async_v4(dst, wl, wl != 0u ? wm : 0u, __float_as_uint(wx), __float_as_uint(wy), dbar);
            st_async_b32(dst + 16, __float_as_uint(wz), dbar);
        }
        mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
        if (t == 0 && it + 2 < iters) mbar_arrive_expect_tx(&bars[buf], NSLOT * kMsgBytes);
        // 64 slots, two per lane (unused ones stay zero)
        const uint2 k0 = *reinterpret_cast<const uint2 *>(&slots[buf][lane][0]);
        const uint2 k1 = *reinterpret_cast<const uint2 *>(&slots[buf][lane + 32][0]);
        const bool second = k1.y > k0.y || (k1.y == k0.y && k1.x > k0.x);
        const uint32_t khi = second ? k1.y : k0.y, klo = second ? k1.x : k0.x;
        const uint32_t gh = __reduce_max_sync(0xffffffffu, khi);
        const uint32_t gl = __reduce_max_sync(0xffffffffu, khi == gh ? klo : 0u);
        const uint32_t win = tie_decode(gl);
        const uint32_t ws = (win / CAP) * NW + ((win % THREADS) >> 5);
        cx = __uint_as_float(slots[buf][ws][2]);
        cy = __uint_as_float(slots[buf][ws][3]);
        cz = __uint_as_float(slots[buf][ws][4]);
        if (rank == 0 && t == 0) idxs[it + 1] = (int)win;
    }
    if (CLUSTER > 1) cluster_sync_all();
}

// ---- C: single CTA, shared-memory slots + one __syncthreads ---------------------------
template <int THREADS, int PTS>
__global__ void __launch_bounds__(THREADS, 1)
    fps_cta_kernel(int n, int m, const float *__restrict__ inp, int *__restrict__ out) {
    constexpr int NW = THREADS / 32;
    __shared__ __align__(16) uint32_t slots[2][32][8];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int scene = blockIdx.x;
    const float *data = inp + (size_t)scene * n * 3;
    int *idxs = out + (size_t)scene * m;
    float x[PTS], y[PTS], z[PTS], td[PTS];
    uint32_t tw[PTS];
#pragma unroll
    for (int p = 0; p < PTS; ++p) {
        uint32_t k = p * THREADS + t;
        tw[p] = tie_word(k);
        if (k < (uint32_t)n) {
            x[p] = data[(size_t)k * 3 + 0];
            y[p] = data[(size_t)k * 3 + 1];
            z[p] = data[(size_t)k * 3 + 2];
            td[p] = 1e38f;
        } else {
            x[p] = y[p] = z[p] = 0.f;
            td[p] = -1.f;
            tw[p] = 0u;
        }
    }
    for (int i = t; i < 2 * 32 * 8; i += THREADS) (&slots[0][0][0])[i] = 0u;
    __syncthreads();
    float cx = data[0], cy = data[1], cz = data[2];
    if (t == 0) idxs[0] = 0;
    const int iters = m - 1;
    for (int it = 0; it < iters; ++it) {
        const int buf = it & 1;
        float mx = -1.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            float d = sqdist(x[p] - cx, y[p] - cy, z[p] - cz);
            td[p] = fminf(d, td[p]);
            mx = fmaxf(mx, td[p]);
        }
        const uint32_t hi = __float_as_uint(fmaxf(mx, 0.f));
        const uint32_t wm = __reduce_max_sync(0xffffffffu, hi);
        uint32_t lo = 0u;
        float bx = 0.f, by = 0.f, bz = 0.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            bool better = td[p] == mx && tw[p] > lo;
            lo = better ? tw[p] : lo;
            bx = better ? x[p] : bx;
            by = better ? y[p] : by;
            bz = better ? z[p] : bz;
        }
        lo = hi == wm ? lo : 0u;
        const uint32_t wl = __reduce_max_sync(0xffffffffu, lo);
        if (lo == wl && (wl != 0u || lane == 0)) {
            *reinterpret_cast<uint4 *>(&slots[buf][warp][0]) =
                make_uint4(wl, wl != 0u ? wm : 0u, __float_as_uint(bx), __float_as_uint(by));
            slots[buf][warp][4] = __float_as_uint(bz);
        }
        __syncthreads();
        const uint2 k0 = *reinterpret_cast<const uint2 *>(&slots[buf][lane][0]);
        const uint32_t gh = __reduce_max_sync(0xffffffffu, k0.y);
        const uint32_t gl = __reduce_max_sync(0xffffffffu, k0.y == gh ? k0.x : 0u);
        const uint32_t win = tie_decode(gl);
        const uint32_t ws = (win % THREADS) >> 5;
        cx = __uint_as_float(slots[buf][ws][2]);
        cy = __uint_as_float(slots[buf][ws][3]);
        cz = __uint_as_float(slots[buf][ws][4]);
        if (t == 0) idxs[it + 1] = (int)win;
    }
    (void)NW;
}

// ---- D: one warp per scene -----------------------------------------------------------
template <int PTS>
__global__ void __launch_bounds__(32)
    fps_warp_kernel(int n, int m, const float *__restrict__ inp, int *__restrict__ out) {
    const int lane = threadIdx.x;
    const int scene = blockIdx.x;
    const float *data = inp + (size_t)scene * n * 3;
    int *idxs = out + (size_t)scene * m;
    float x[PTS], y[PTS], z[PTS], td[PTS];
    uint32_t tw[PTS];
#pragma unroll
    for (int p = 0; p < PTS; ++p) {
        uint32_t k = p * 32 + lane;
        tw[p] = tie_word(k);
        if (k < (uint32_t)n) {
            x[p] = data[(size_t)k * 3 + 0];
            y[p] = data[(size_t)k * 3 + 1];
            z[p] = data[(size_t)k * 3 + 2];
            td[p] = 1e38f;
        } else {
            x[p] = y[p] = z[p] = 0.f;
            td[p] = -1.f;
            tw[p] = 0u;
        }
    }
    float cx = data[0], cy = data[1], cz = data[2];
    if (lane == 0) idxs[0] = 0;
    for (int it = 0; it < m - 1; ++it) {
        float mx = -1.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            float d = sqdist(x[p] - cx, y[p] - cy, z[p] - cz);
            td[p] = fminf(d, td[p]);
            mx = fmaxf(mx, td[p]);
        }
        const uint32_t hi = __float_as_uint(fmaxf(mx, 0.f));
        const uint32_t wm = __reduce_max_sync(0xffffffffu, hi);
        uint32_t lo = 0u;
        float bx = 0.f, by = 0.f, bz = 0.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            bool better = td[p] == mx && tw[p] > lo;
            lo = better ? tw[p] : lo;
            bx = better ? x[p] : bx;
            by = better ? y[p] : by;
            bz = better ? z[p] : bz;
        }
        lo = hi == wm ? lo : 0u;
        const uint32_t wl = __reduce_max_sync(0xffffffffu, lo);
        const int src = __ffs(__ballot_sync(0xffffffffu, lo == wl)) - 1;
        cx = __shfl_sync(0xffffffffu, bx, src);
        cy = __shfl_sync(0xffffffffu, by, src);
        cz = __shfl_sync(0xffffffffu, bz, src);
        if (lane == 0) idxs[it + 1] = (int)tie_decode(wl);
    }
}

template <typename K>
void launch_cluster(K kern, int cluster, int threads, int b, int n, int m, const float *inp, int *out) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)b * cluster);
    cfg.blockDim = dim3(threads);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, n, m, inp, out);
}

}  // namespace
}  // namespace hf

using namespace hf;

template <typename F>
float timeit(F f, int reps = 5) {
    f();
    cudaDeviceSynchronize();
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main() {
    const int B = 16;
    struct Case { int n, m; } cases[] = {{16384, 4096}, {4096, 1024}, {1024, 256}, {256, 64}};
    for (auto c : cases) {
        int n = c.n, m = c.m;
        std::vector<float> h((size_t)B * n * 3);
        srand(7 + n);
        for (size_t i = 0; i < h.size(); ++i) h[i] = (float)rand() / RAND_MAX * 40.f;
        // duplicates to exercise ties
        for (int i = 0; i < n / 8; ++i) for (int k = 0; k < 3; ++k) h[(size_t)(n - 1 - i) * 3 + k] = h[(size_t)i * 3 + k];
        float *d;
        int *o0, *o1;
        cudaMalloc(&d, h.size() * 4);
        cudaMalloc(&o0, (size_t)B * m * 4);
        cudaMalloc(&o1, (size_t)B * m * 4);
        cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
        std::vector<int> r0((size_t)B * m), r1((size_t)B * m);
        auto check = [&](const char *name, float ms) {
            cudaMemcpy(r1.data(), o1, r1.size() * 4, cudaMemcpyDeviceToHost);
            cudaError_t e = cudaGetLastError();
            printf("n=%6d m=%5d %-28s %8.4f ms %7.1f ns/iter  same=%d %s\n", n, m, name, ms, ms * 1e6 / (m - 1),
                   (int)(r0 == r1), e == cudaSuccess ? "" : cudaGetErrorString(e));
            cudaMemset(o1, 0xff, r1.size() * 4);
        };
        float ms = timeit([&] { hf_farthest_point_sample(B, n, m, d, nullptr, o0, nullptr); });
        cudaMemcpy(r0.data(), o0, r0.size() * 4, cudaMemcpyDeviceToHost);
        printf("n=%6d m=%5d %-28s %8.4f ms %7.1f ns/iter\n", n, m, "A library", ms, ms * 1e6 / (m - 1));
        if (n == 16384) {
            check("B push c8 t256 p8", timeit([&] { launch_cluster(fps_warp_push_kernel<8, 256, 8>, 8, 256, B, n, m, d, o1); }));
            check("B push c4 t512 p8", timeit([&] { launch_cluster(fps_warp_push_kernel<4, 512, 8>, 4, 512, B, n, m, d, o1); }));
            check("B push c8 t128 p16", timeit([&] { launch_cluster(fps_warp_push_kernel<8, 128, 16>, 8, 128, B, n, m, d, o1); }));
            check("C cta t1024 p16", timeit([&] { fps_cta_kernel<1024, 16><<<B, 1024>>>(n, m, d, o1); }));
        }
        if (n == 4096) {
            check("B push c8 t256 p2", timeit([&] { launch_cluster(fps_warp_push_kernel<8, 256, 2>, 8, 256, B, n, m, d, o1); }));
            check("B push c8 t128 p4", timeit([&] { launch_cluster(fps_warp_push_kernel<8, 128, 4>, 8, 128, B, n, m, d, o1); }));
            check("B push c4 t256 p4", timeit([&] { launch_cluster(fps_warp_push_kernel<4, 256, 4>, 4, 256, B, n, m, d, o1); }));
            check("B push c1 t512 p8", timeit([&] { launch_cluster(fps_warp_push_kernel<1, 512, 8>, 1, 512, B, n, m, d, o1); }));
            check("C cta t512 p8", timeit([&] { fps_cta_kernel<512, 8><<<B, 512>>>(n, m, d, o1); }));
            check("C cta t256 p16", timeit([&] { fps_cta_kernel<256, 16><<<B, 256>>>(n, m, d, o1); }));
            check("C cta t1024 p4", timeit([&] { fps_cta_kernel<1024, 4><<<B, 1024>>>(n, m, d, o1); }));
        }
        if (n == 1024) {
            check("B push c1 t256 p4", timeit([&] { launch_cluster(fps_warp_push_kernel<1, 256, 4>, 1, 256, B, n, m, d, o1); }));
            check("C cta t256 p4", timeit([&] { fps_cta_kernel<256, 4><<<B, 256>>>(n, m, d, o1); }));
            check("C cta t128 p8", timeit([&] { fps_cta_kernel<128, 8><<<B, 128>>>(n, m, d, o1); }));
            check("C cta t64 p16", timeit([&] { fps_cta_kernel<64, 16><<<B, 64>>>(n, m, d, o1); }));
            check("D warp p32", timeit([&] { fps_warp_kernel<32><<<B, 32>>>(n, m, d, o1); }));
        }
        if (n == 256) {
            check("B push c1 t256 p1", timeit([&] { launch_cluster(fps_warp_push_kernel<1, 256, 1>, 1, 256, B, n, m, d, o1); }));
            check("C cta t256 p1", timeit([&] { fps_cta_kernel<256, 1><<<B, 256>>>(n, m, d, o1); }));
            check("C cta t64 p4", timeit([&] { fps_cta_kernel<64, 4><<<B, 64>>>(n, m, d, o1); }));
            check("D warp p8", timeit([&] { fps_warp_kernel<8><<<B, 32>>>(n, m, d, o1); }));
        }
        cudaFree(d);
        cudaFree(o0);
        cudaFree(o1);
    }
    return 0;
}
