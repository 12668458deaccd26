// Input side of the path: a raw KITTI velodyne scan -> rectified camera frame ->
// in-front-of-camera and image field-of-view filter, order preserving, as hf/core/obj_utils.py:221-277
// (get_lidar_point_cloud) with hf/core/calib_utils.py:370-407 (lidar_to_cam_frame) and :280-296 (project_to_image).
//
// The reference does this in float64 numpy (the float32 scan is promoted by the appended column of ones), so the
// arithmetic here is double too; the kept points are written as float32 [x, y, z, intensity], the type the network
// input is cast to (hf/datasets/kitti/kitti_dataset.py:341-370).
//
// Two launches, no atomics, deterministic order: (1) every CTA tests its tile of 1024 points, stores one keep bit
// per point and the tile's count, (2) every CTA sums the counts of the tiles before it (at most a few hundred
// values), redoes the frame change of its kept points only (no projection, no divisions) and writes them at their
// rank.  Traffic: 16 B read per point, 16 B read again + 16 B written per kept point.
#include "common.cuh"

namespace hf {
namespace {

constexpr int kTile = 1024;

struct Calib {
    double rect[12];  // first three rows of r0_rect(4x4) @ tr_velodyne_to_cam(4x4)
    double p2[12];    // camera matrix (3,4)
};

__device__ __forceinline__ double row4(const double *m, double x, double y, double z) {
    return fma(m[2], z, fma(m[1], y, m[0] * x)) + m[3];
}

// Camera-frame point and the filter of obj_utils.py:252-265.
__device__ __forceinline__ bool fov_keep(const Calib &c, float4 p, double im_w, double im_h, double &X, double &Y,
                                         double &Z) {
    X = row4(c.rect + 0, p.x, p.y, p.z);
    Y = row4(c.rect + 4, p.x, p.y, p.z);
    Z = row4(c.rect + 8, p.x, p.y, p.z);
    if (!(Z > 0.0)) return false;
    const double a = row4(c.p2 + 0, X, Y, Z);
    const double b = row4(c.p2 + 4, X, Y, Z);
    const double d = row4(c.p2 + 8, X, Y, Z);
    const double u = a / d, v = b / d;
    return u > 0.0 && u < im_w && v > 0.0 && v < im_h;
}

__device__ __forceinline__ int block_count(unsigned ballot, int *warp_sums, int &before_in_block) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_sums[warp] = __popc(ballot);
    __syncthreads();
    int mine = lane < (kTile / 32) ? warp_sums[lane] : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    const int warp_excl = __shfl_sync(0xffffffffu, incl - mine, warp);
    before_in_block = warp_excl + __popc(ballot & ((1u << lane) - 1u));
    return total;
}

template <bool WRITE>
__global__ void __launch_bounds__(kTile) fov_filter_kernel(int n, int tiles, const float4 *__restrict__ pts,
                                                           const Calib *__restrict__ calib, double im_w, double im_h,
                                                           int *__restrict__ tile_counts, unsigned *__restrict__ keep_bits,
                                                           float4 *__restrict__ out, int *__restrict__ count) {
    __shared__ int warp_sums[kTile / 32];
    __shared__ int tile_base;
    __shared__ Calib c;
    const int scan = blockIdx.y, tile = blockIdx.x;
    const int i = tile * kTile + threadIdx.x;
    const int words = tiles * (kTile / 32);
    unsigned *bits = keep_bits + (size_t)scan * words + (i >> 5);
    // every global load of the thread is issued before the first barrier: one memory latency per CTA, not four
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n) p = __ldg(pts + (size_t)scan * n + i);
    const unsigned stored = WRITE ? *bits : 0u;
    if (threadIdx.x < 24) reinterpret_cast<double *>(&c)[threadIdx.x] = reinterpret_cast<const double *>(calib + scan)[threadIdx.x];
    if (WRITE && threadIdx.x >= 32 && threadIdx.x < 64) {  // warp 1: rank of the tile's first kept point
        int s = 0;
        for (int t = threadIdx.x - 32; t < tile; t += 32) s += tile_counts[(size_t)scan * tiles + t];
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 32) tile_base = s;
    }
    __syncthreads();
    double X = 0, Y = 0, Z = 0;
    int before;
    if (!WRITE) {
        const bool keep = i < n && fov_keep(c, p, im_w, im_h, X, Y, Z);
        const unsigned ballot = __ballot_sync(0xffffffffu, keep);
        if ((threadIdx.x & 31) == 0) *bits = ballot;
        const int total = block_count(ballot, warp_sums, before);
        if (threadIdx.x == 0) tile_counts[(size_t)scan * tiles + tile] = total;
        return;
    }
    const int total = block_count(stored, warp_sums, before);
    if ((stored >> (threadIdx.x & 31)) & 1u) {
        X = row4(c.rect + 0, p.x, p.y, p.z);
        Y = row4(c.rect + 4, p.x, p.y, p.z);
        Z = row4(c.rect + 8, p.x, p.y, p.z);
        out[(size_t)scan * n + tile_base + before] = make_float4((float)X, (float)Y, (float)Z, p.w);
    }
    if (tile == tiles - 1 && threadIdx.x == 0) count[scan] = tile_base + total;
}

}  // namespace
}  // namespace hf

extern "C" size_t hf_lidar_fov_filter_workspace_bytes(int scans, int n) {
    if (scans <= 0 || n <= 0) return 16;
    const size_t tiles = ((size_t)n + hf::kTile - 1) / hf::kTile;
    return (((size_t)scans * tiles * sizeof(int) + 15) & ~(size_t)15) + (size_t)scans * tiles * (hf::kTile / 32) * sizeof(unsigned);
}

extern "C" int hf_lidar_fov_filter(int scans, int n, const float *points, const double *calib, int im_w, int im_h,
                                   float *out, int *count, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(scans >= 0 && n >= 0, "lidar_fov_filter expects (scans, n, 4) points");
    HF_REQUIRE(im_w > 0 && im_h > 0, "lidar_fov_filter expects a positive image size [w, h]");
    if (scans == 0) return HF_OK;
    HF_REQUIRE(count && calib, "lidar_fov_filter: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        HF_CUDA_OK(cudaMemsetAsync(count, 0, (size_t)scans * sizeof(int), s));
        return HF_OK;
    }
    HF_REQUIRE(points && out, "lidar_fov_filter: null pointer");
    HF_REQUIRE((reinterpret_cast<uintptr_t>(points) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(calib) & 7) == 0,
               "lidar_fov_filter: points/out must be 16-byte aligned, calib 8-byte aligned");
    HF_REQUIRE(scans <= 65535, "lidar_fov_filter: at most 65535 scans per call");
    HF_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "lidar_fov_filter: null or unaligned workspace");
    if (workspace_bytes < hf_lidar_fov_filter_workspace_bytes(scans, n)) {
        set_error("lidar_fov_filter: workspace too small");
        return HF_ERR_WORKSPACE;
    }
    const int tiles = (n + kTile - 1) / kTile;
    const dim3 grid(tiles, scans);
    int *tile_counts = static_cast<int *>(workspace);
    unsigned *keep_bits = reinterpret_cast<unsigned *>(static_cast<unsigned char *>(workspace) +
                                                       (((size_t)scans * tiles * sizeof(int) + 15) & ~(size_t)15));
    const Calib *c = reinterpret_cast<const Calib *>(calib);
    const float4 *p = reinterpret_cast<const float4 *>(points);
    fov_filter_kernel<false><<<grid, kTile, 0, s>>>(n, tiles, p, c, (double)im_w, (double)im_h, tile_counts, keep_bits, nullptr, nullptr);
    HF_LAUNCH_OK("fov_filter_kernel<count>");
    fov_filter_kernel<true><<<grid, kTile, 0, s>>>(n, tiles, p, c, (double)im_w, (double)im_h, tile_counts,
                                                  keep_bits, reinterpret_cast<float4 *>(out), count);
    HF_LAUNCH_OK("fov_filter_kernel<write>");
    return HF_OK;
}
