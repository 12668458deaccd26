// gpc_probe.cu — how many thread-block clusters of each size does this GPU co-schedule, and how are its
// SMs grouped into GPCs?  (A cluster must sit inside one GPC; FPS runs one cluster per scan.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o tools/build/gpc_probe tools/gpc_probe.cu
//   tools/build/gpc_probe
// Prints cudaOccupancyMaxActiveClusters for cluster sizes 1..16 at the FPS kernel's footprint (128 / 256
// threads, 24 KB / 48 KB of shared memory, 1 CTA per SM forced by __launch_bounds__(T,1) + big smem), then the
// GPC partition derived from which SMs ever share a cluster of 2.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

__global__ void where_kernel(int *smid_of_cta) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0) smid_of_cta[blockIdx.x] = (int)smid;
    // stay resident long enough that the whole grid is co-scheduled wave by wave
    long long t0 = clock64();
    while (clock64() - t0 < 200000) {
    }
}

static int find(std::vector<int> &p, int a) { return p[a] == a ? a : p[a] = find(p, p[a]); }

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    printf("# %s, %d SMs\n", prop.name, sms);
    cudaFuncSetAttribute(where_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(where_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int smems[] = {24 * 1024, 120 * 1024, 200 * 1024};  // 200 KB: one CTA per SM for certain
    for (int threads : {128, 256}) {
        for (int smem : smems) {
            printf("threads=%d smem=%dKB  max co-resident clusters by size:", threads, smem / 1024);
            for (int c = 1; c <= 16; ++c) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(c * 64);
                cfg.blockDim = dim3(threads);
                cfg.dynamicSmemBytes = smem;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = c;
                at[0].val.clusterDim.y = 1;
                at[0].val.clusterDim.z = 1;
                cfg.attrs = at;
                cfg.numAttrs = 1;
                int n = -1;
                cudaError_t e = cudaOccupancyMaxActiveClusters(&n, where_kernel, &cfg);
                if (e != cudaSuccess) {
                    cudaGetLastError();
                    printf(" %d:err", c);
                } else {
                    printf(" %d:%d(%dSM)", c, n, n * c);
                }
            }
            printf("\n");
        }
    }
    // GPC partition: clusters of 2 with 200 KB smem (1 CTA / SM); union the SMs of each cluster over many launches
    std::vector<int> parent(sms);
    for (int i = 0; i < sms; ++i) parent[i] = i;
    int *d;
    const int grid = 2 * 74;
    cudaMalloc(&d, grid * sizeof(int));
    std::vector<int> h(grid);
    for (int c : {2, 3, 4, 5, 6, 7, 8}) {
        for (int rep = 0; rep < 6; ++rep) {
            const int g = (grid / c) * c;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(g);
            cfg.blockDim = dim3(128);
            cfg.dynamicSmemBytes = 200 * 1024;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = c;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            if (cudaLaunchKernelEx(&cfg, where_kernel, d) != cudaSuccess) {
                printf("launch c=%d failed: %s\n", c, cudaGetErrorString(cudaGetLastError()));
                continue;
            }
            cudaDeviceSynchronize();
            cudaMemcpy(h.data(), d, g * sizeof(int), cudaMemcpyDeviceToHost);
            for (int k = 0; k < g; k += c)
                for (int j = 1; j < c; ++j) parent[find(parent, h[k + j])] = find(parent, h[k]);
        }
    }
    std::vector<std::vector<int>> groups(sms);
    for (int i = 0; i < sms; ++i) groups[find(parent, i)].push_back(i);
    std::vector<int> sizes;
    printf("GPC partition (SMs that ever shared a cluster):\n");
    for (auto &g : groups)
        if (!g.empty()) {
            sizes.push_back((int)g.size());
            printf("  %2d SMs:", (int)g.size());
            for (int s : g) printf(" %d", s);
            printf("\n");
        }
    std::sort(sizes.begin(), sizes.end());
    printf("GPC sizes:");
    for (int s : sizes) printf(" %d", s);
    printf("\n");
    return 0;
}
