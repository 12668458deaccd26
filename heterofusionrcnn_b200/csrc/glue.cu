// glue.cu — the callers' glue around BEV IoU / NMS (hf/core/compute_iou.py), on the device, for sm_100a.
//
//   boxes3d_to_bev_tf   compute_iou.py:7-20    [x,y,z,l,w,h,ry] -> [x-l/2, z-w/2, x+l/2, z+w/2, ry]
//   box3d_iou_tf        compute_iou.py:23-64   BEV overlap x height overlap -> 3-D IoU (+ the 2-D IoU)
//   oriented_nms_tf     compute_iou.py:67-80   BEV boxes, tf.nn.top_k over ALL scores (a full descending
//                                               sort, equal scores keep the lower index first), oriented_nms,
//                                               keep list mapped back to the caller's box order
// The reference runs these as a dozen small TF ops (tile / reshape / gather) around the two custom ops, and
// per scene inside a tf.map_fn (rpn_model.py:683).  Here they are batched over scenes and stay on the device.
#include <float.h>
#include <stdint.h>

#include <mutex>

#include "common.cuh"

namespace hf {
namespace {

__global__ void boxes3d_to_bev_kernel(long long total, const float *__restrict__ b3, const int *__restrict__ order,
                                      int per_scene, float *__restrict__ bev) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        long long src = i;
        if (order) src = (i / per_scene) * per_scene + order[i];  // row i of the sorted list
        const float *p = b3 + src * 7;
        const float hl = p[3] / 2, hw = p[4] / 2;
        float *o = bev + i * 5;
        o[0] = p[0] - hl;
        o[1] = p[2] - hw;
        o[2] = p[0] + hl;
        o[3] = p[2] + hw;
        o[4] = p[6];
    }
}

// box_8c_encoder.py:101-184 (tf_box_3d_to_box_8co) after the proposal expansion of rcnn_model.py:462-476:
// corners (N,3,8), rows x/y/z, columns P1..P8; x' = cos*x + sin*z, z' = -sin*x + cos*z about the box origin.
__global__ void boxes3d_to_corners_kernel(long long num, const float *__restrict__ b3, float expand,
                                          float *__restrict__ corners) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < num * 8; i += (long long)gridDim.x * blockDim.x) {
        const long long bx = i >> 3;
        const int k = (int)(i & 7);
        const float *p = b3 + bx * 7;
        const float e2 = __fmul_rn(2.f, expand);
        const float l = __fadd_rn(p[3], e2), w = __fadd_rn(p[4], e2), h = __fadd_rn(p[5], e2);
        const float cy = __fadd_rn(p[1], expand);
        const float hl = l / 2, hw = w / 2;
        const float xc = (k & 2) ? -hl : hl;                      // P1 P2 +, P3 P4 -
        const float zc = ((k & 3) == 1 || (k & 3) == 2) ? -hw : hw;  // P1 +, P2 -, P3 -, P4 +
        const float yc = (k & 4) ? -h : 0.f;
        const float c = cosf(p[6]), sn = sinf(p[6]);
        float *o = corners + bx * 24;
        o[k] = __fadd_rn(__fadd_rn(__fmul_rn(c, xc), __fmul_rn(sn, zc)), p[0]);
        o[8 + k] = __fadd_rn(yc, cy);
        o[16 + k] = __fadd_rn(__fadd_rn(__fmul_rn(-sn, xc), __fmul_rn(c, zc)), p[2]);
    }
}

// iou3d holds the BEV overlap on entry (compute_iou.py:38-62).
__global__ void box3d_iou_kernel(int na, int nb, const float *__restrict__ a, const float *__restrict__ b,
                                 float *__restrict__ iou3d) {
    const long long total = (long long)na * nb;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const float *pa = a + (i / nb) * 7, *pb = b + (i % nb) * 7;
        const float amin = __fsub_rn(pa[1], pa[5]), amax = pa[1], bmin = __fsub_rn(pb[1], pb[5]), bmax = pb[1];
        const float oh = fminf(fmaxf(__fsub_rn(fminf(amax, bmax), fmaxf(amin, bmin)), 0.f), FLT_MAX);
        const float o3 = __fmul_rn(iou3d[i], oh);
        const float va = __fmul_rn(__fmul_rn(pa[3], pa[4]), pa[5]), vb = __fmul_rn(__fmul_rn(pb[3], pb[4]), pb[5]);
        const float den = fminf(fmaxf(__fsub_rn(__fadd_rn(va, vb), o3), 1e-7f), FLT_MAX);
        iou3d[i] = __fdiv_rn(o3, den);
    }
}

// Descending stable order of one scene's scores: 64-bit keys {~ordered(score), index}, bitonic sort in shared memory.
constexpr int kSortThreads = 1024;
__global__ void __launch_bounds__(kSortThreads) score_order_kernel(int n, int P, const float *__restrict__ scores,
                                                                   int *__restrict__ order) {
    extern __shared__ __align__(16) unsigned long long keys[];
    const int t = threadIdx.x;
    const float *sc = scores + (size_t)blockIdx.x * n;
    for (int i = t; i < P; i += kSortThreads) {
        unsigned long long k = ~0ull;
        if (i < n) {
            float f = sc[i];
            if (f == 0.f) f = 0.f;  // -0 and +0 are one score
            unsigned u = __float_as_uint(f);
            u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending order of the float as unsigned
            if (f != f) u = 0u;                              // NaN scores sort last
            k = ((unsigned long long)(~u) << 32) | (unsigned)i;
        }
        keys[i] = k;
    }
    __syncthreads();
    for (int k2 = 2; k2 <= P; k2 <<= 1) {
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            for (int i = t; i < P; i += kSortThreads) {
                const int p = i ^ j;
                if (p > i) {
                    const unsigned long long x = keys[i], y = keys[p];
                    if ((x > y) == ((i & k2) == 0)) {
                        keys[i] = y;
                        keys[p] = x;
                    }
                }
            }
            __syncthreads();
        }
    }
    int *o = order + (size_t)blockIdx.x * n;
    for (int i = t; i < n; i += kSortThreads) o[i] = (int)(unsigned)keys[i];
}

__global__ void map_keep_kernel(long long total, int per_scene, const int *__restrict__ order,
                                const int *__restrict__ keep_sorted, int *__restrict__ keep) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long base = (i / per_scene) * per_scene;
        keep[i] = order[base + keep_sorted[i]];
    }
}

// sb_nms_fn's tail (hf/core/models/model_util.py:124-142): the first nms_size entries of the keep list, mapped back to
// the caller's box order; entries from `limit` on are -1, limit = min(nms_size, N) when the duplicate padding of the
// keep list is kept (fixed_num_proposal_nms) and min(nms_size, kept) when tf.unique removes it (the kept boxes are
// distinct and the padding repeats keep[0], so unique leaves exactly the kept ones in order).
__global__ void sb_nms_select_kernel(int scenes, int per_scene, int keep_stride, int nms_size, int fixed,
                                     const int *__restrict__ order, const int *__restrict__ keep_sorted,
                                     const int *__restrict__ num_kept, int *__restrict__ out, int *__restrict__ out_num) {
    const long long total = (long long)scenes * nms_size;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int scene = (int)(i / nms_size), j = (int)(i - (long long)scene * nms_size);
        const int limit = min(nms_size, fixed ? per_scene : num_kept[scene]);
        const long long base = (long long)scene * per_scene;
        out[i] = j < limit ? order[base + keep_sorted[(long long)scene * keep_stride + j]] : -1;
        if (j == 0) out_num[scene] = limit;
    }
}

inline size_t a16(size_t v) { return (v + 15) & ~(size_t)15; }
inline int grid_for(long long total) {
    long long g = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace hf

extern "C" int hf_boxes3d_to_bev(long long num, const float *boxes3d, float *boxes_bev, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num >= 0, "boxes3d_to_bev expects (N,7) boxes3d");
    HF_REQUIRE(num == 0 || (boxes3d && boxes_bev), "boxes3d_to_bev: null pointer");
    if (num == 0) return HF_OK;
    boxes3d_to_bev_kernel<<<grid_for(num), 256, 0, (cudaStream_t)stream>>>(num, boxes3d, nullptr, 1, boxes_bev);
    HF_LAUNCH_OK("boxes3d_to_bev_kernel");
    return HF_OK;
}

extern "C" int hf_boxes3d_to_corners(long long num, const float *boxes3d, float expand_length, float *corners,
                                     hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num >= 0, "boxes3d_to_corners expects (N,7) boxes3d");
    HF_REQUIRE(num == 0 || (boxes3d && corners), "boxes3d_to_corners: null pointer");
    if (num == 0) return HF_OK;
    boxes3d_to_corners_kernel<<<grid_for(num * 8), 256, 0, (cudaStream_t)stream>>>(num, boxes3d, expand_length, corners);
    HF_LAUNCH_OK("boxes3d_to_corners_kernel");
    return HF_OK;
}

extern "C" size_t hf_box3d_iou_workspace_bytes(int num_a, int num_b) {
    if (num_a < 0 || num_b < 0) return 0;
    return hf::a16((size_t)num_a * 20) + hf::a16((size_t)num_b * 20) + 16;
}

extern "C" int hf_box3d_iou(int num_a, const float *boxes_a, int num_b, const float *boxes_b, float *iou_3d, float *iou_2d,
                            void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num_a >= 0 && num_b >= 0, "box3d_iou expects (N,7) and (M,7) boxes");
    if (num_a == 0 || num_b == 0) return HF_OK;
    HF_REQUIRE(boxes_a && boxes_b && iou_3d && iou_2d, "box3d_iou: null pointer");
    HF_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "box3d_iou: null or unaligned workspace");
    if (workspace_bytes < hf_box3d_iou_workspace_bytes(num_a, num_b)) {
        set_error("box3d_iou: workspace too small");
        return HF_ERR_WORKSPACE;
    }
    float *bev_a = static_cast<float *>(workspace);
    float *bev_b = reinterpret_cast<float *>(static_cast<unsigned char *>(workspace) + a16((size_t)num_a * 20));
    int rc = hf_boxes3d_to_bev(num_a, boxes_a, bev_a, stream);
    if (rc == HF_OK) rc = hf_boxes3d_to_bev(num_b, boxes_b, bev_b, stream);
    if (rc == HF_OK) rc = hf_compute_bev_iou(num_a, bev_a, num_b, bev_b, iou_3d, iou_2d, stream);  // iou_3d <- overlap
    if (rc != HF_OK) return rc;
    box3d_iou_kernel<<<grid_for((long long)num_a * num_b), 256, 0, (cudaStream_t)stream>>>(num_a, num_b, boxes_a, boxes_b, iou_3d);
    HF_LAUNCH_OK("box3d_iou_kernel");
    return HF_OK;
}

extern "C" size_t hf_oriented_nms_scored_workspace_bytes(int scenes, int boxes_num) {
    if (scenes <= 0 || boxes_num <= 0) return 16;
    const size_t sn = (size_t)scenes * boxes_num;
    return hf::a16(sn * 4) + hf::a16(sn * 20) + hf::a16(sn * 4) + hf_oriented_nms_workspace_bytes(scenes, boxes_num);
}

namespace hf {
namespace {
// keep != null: the whole keep list (S,N) mapped back to the caller's order (oriented_nms_tf);
// sel != null: sb_nms_fn's (S,nms_size) selection with -1 padding and its count.
int nms_scored(const float *boxes3d, const float *scores, int scenes, int boxes_num, float thresh, int *keep, int *num_kept,
               int nms_size, int fixed, int *sel, int *sel_num, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    HF_REQUIRE(thresh >= 0, "Need nms_threshold >= 0, got %f", (double)thresh);
    HF_REQUIRE(scenes >= 0 && boxes_num >= 0, "oriented_nms expects (S,N,7) boxes and (S,N) scores");
    if (scenes == 0 || boxes_num == 0) return HF_OK;
    HF_REQUIRE(boxes3d && scores && (keep || (sel && sel_num)), "oriented_nms: null pointer");
    HF_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "oriented_nms: null or unaligned workspace");
    if (workspace_bytes < hf_oriented_nms_scored_workspace_bytes(scenes, boxes_num)) {
        set_error("oriented_nms: workspace too small");
        return HF_ERR_WORKSPACE;
    }
    int P = 32;
    while (P < boxes_num) P <<= 1;
    if ((size_t)P * 8 > 200 * 1024) {
        set_error("oriented_nms: more than %d boxes per scene are not supported by the score sort", 200 * 1024 / 8);
        return HF_ERR_UNSUPPORTED;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t sn = (size_t)scenes * boxes_num;
    unsigned char *w = static_cast<unsigned char *>(workspace);
    int *order = reinterpret_cast<int *>(w);
    float *bev = reinterpret_cast<float *>(w + a16(sn * 4));
    int *keep_sorted = reinterpret_cast<int *>(w + a16(sn * 4) + a16(sn * 20));
    unsigned char *nms_ws = w + a16(sn * 4) + a16(sn * 20) + a16(sn * 4);
    if (!num_kept && sel) {  // the selection needs the count: it lives behind the NMS workspace (hf_sb_nms_workspace_bytes)
        num_kept = reinterpret_cast<int *>(nms_ws + a16(hf_oriented_nms_workspace_bytes(scenes, boxes_num)));
    }
    const size_t smem = (size_t)P * 8;
    if (smem > 48 * 1024) HF_CUDA_OK(cudaFuncSetAttribute(score_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    score_order_kernel<<<scenes, kSortThreads, smem, s>>>(boxes_num, P, scores, order);
    HF_LAUNCH_OK("score_order_kernel");
    boxes3d_to_bev_kernel<<<grid_for((long long)sn), 256, 0, s>>>((long long)sn, boxes3d, order, boxes_num, bev);
    HF_LAUNCH_OK("boxes3d_to_bev_kernel");
    // the selection only looks at the first nms_size entries of the keep list: the sweep stops there
    const int kstride = (sel && !keep && nms_size < boxes_num) ? nms_size : boxes_num;
    int rc = hf_oriented_nms_topk(bev, scenes, boxes_num, thresh, kstride, keep_sorted, num_kept, nms_ws,
                                  hf_oriented_nms_workspace_bytes(scenes, boxes_num), stream);
    if (rc != HF_OK) return rc;
    if (keep) {
        map_keep_kernel<<<grid_for((long long)sn), 256, 0, s>>>((long long)sn, boxes_num, order, keep_sorted, keep);
        HF_LAUNCH_OK("map_keep_kernel");
    }
    if (sel) {
        sb_nms_select_kernel<<<grid_for((long long)scenes * nms_size), 256, 0, s>>>(scenes, boxes_num, kstride, nms_size, fixed,
                                                                                   order, keep_sorted, num_kept, sel, sel_num);
        HF_LAUNCH_OK("sb_nms_select_kernel");
    }
    return HF_OK;
}
}  // namespace
}  // namespace hf

extern "C" int hf_oriented_nms_scored(const float *boxes3d, const float *scores, int scenes, int boxes_num, float thresh,
                                      int *keep, int *num_kept, void *workspace, size_t workspace_bytes,
                                      hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(scenes == 0 || boxes_num == 0 || keep, "oriented_nms: null pointer");
    if (workspace_bytes < hf_oriented_nms_scored_workspace_bytes(scenes, boxes_num)) {
        set_error("oriented_nms: workspace too small");
        return HF_ERR_WORKSPACE;
    }
    return nms_scored(boxes3d, scores, scenes, boxes_num, thresh, keep, num_kept, 0, 0, nullptr, nullptr,
                      workspace, workspace_bytes, stream);
}

extern "C" size_t hf_sb_nms_workspace_bytes(int scenes, int boxes_num) {
    return hf_oriented_nms_scored_workspace_bytes(scenes, boxes_num) + hf::a16((size_t)(scenes > 0 ? scenes : 0) * 4) + 16;
}

extern "C" int hf_sb_nms(const float *boxes3d, const float *scores, int scenes, int boxes_num, float thresh, int nms_size,
                         int fixed_num_proposal_nms, int *nms_indices, int *num_selected, void *workspace,
                         size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(nms_size >= 0, "sb_nms_fn: nms_size must be >= 0");
    HF_REQUIRE(scenes >= 0 && boxes_num >= 0, "oriented_nms expects (S,N,7) boxes and (S,N) scores");
    if (scenes == 0) return HF_OK;
    HF_REQUIRE(num_selected && (nms_size == 0 || nms_indices), "sb_nms_fn: null pointer");
    if (workspace_bytes < hf_sb_nms_workspace_bytes(scenes, boxes_num)) {
        set_error("sb_nms_fn: workspace too small");
        return HF_ERR_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (boxes_num == 0 || nms_size == 0) {  // nothing to select: all padding, zero selected
        if (nms_size) HF_CUDA_OK(cudaMemsetAsync(nms_indices, 0xff, (size_t)scenes * nms_size * 4, s));
        HF_CUDA_OK(cudaMemsetAsync(num_selected, 0, (size_t)scenes * 4, s));
        return HF_OK;
    }
    return nms_scored(boxes3d, scores, scenes, boxes_num, thresh, nullptr, nullptr, nms_size, fixed_num_proposal_nms ? 1 : 0,
                      nms_indices, num_selected, workspace, workspace_bytes, stream);
}


// ---- hf_box3d_iou_static: box3d_iou for hosts that may not allocate ---------------------------------------------------
// Same four launches as hf_box3d_iou; the one difference is where the two BEV box lists are staged.  hf_box3d_iou
// takes that memory from the caller.  A caller that cannot obtain device memory at call time (an executor that bans
// cudaMalloc on its worker threads because it stalls the whole device, a control loop with a frozen memory budget)
// gets it from the array below, which is part of this library's device image: it occupies 320 KiB on every GPU the
// library is loaded on, from load until the process exits.
namespace hf {
namespace {

constexpr size_t kStaticSideBytes = (size_t)HF_BOX3D_IOU_STATIC_MAX_BOXES * 5 * sizeof(float);  // one BEV list
static_assert(kStaticSideBytes % 256 == 0, "the second list must start on the array's alignment");
// [0, kStaticSideBytes): BEV boxes of boxes_a; [kStaticSideBytes, 2 * kStaticSideBytes): BEV boxes of boxes_b.
static __device__ __align__(256) unsigned char g_box3d_iou_static_scratch[2 * kStaticSideBytes];

// In host code `g_box3d_iou_static_scratch` is only a handle for the runtime, never an address a kernel may be given.
// The device address differs per GPU; it is looked up once per device ordinal, with that device current, and kept.
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
unsigned char *g_static_device_ptr[kStaticMaxDevices];  // nullptr = not looked up on that device yet

int static_scratch_of_current_device(unsigned char **out) {
    int dev = -1;
    HF_CUDA_OK(cudaGetDevice(&dev));
    const bool cached_slot = dev >= 0 && dev < kStaticMaxDevices;
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (cached_slot && g_static_device_ptr[dev]) {
        *out = g_static_device_ptr[dev];
        return HF_OK;
    }
    unsigned char *ptr = nullptr;
    HF_CUDA_OK(cudaGetSymbolAddress((void **)&ptr, g_box3d_iou_static_scratch));
    if (!ptr) {
        set_error("box3d_iou_static: the runtime returned a null address for the library's staging array on device %d", dev);
        return HF_ERR_CUDA;
    }
    if (cached_slot) g_static_device_ptr[dev] = ptr;  // ordinals past the table are looked up on every call
    *out = ptr;
    return HF_OK;
}

}  // namespace
}  // namespace hf

extern "C" int hf_box3d_iou_static(int num_a, const float *boxes_a, int num_b, const float *boxes_b, float *iou_3d,
                                   float *iou_2d, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num_a >= 0 && num_b >= 0, "box3d_iou expects (N,7) and (M,7) boxes");
    if (num_a > HF_BOX3D_IOU_STATIC_MAX_BOXES || num_b > HF_BOX3D_IOU_STATIC_MAX_BOXES) {
        set_error("box3d_iou_static: %d x %d boxes, but the library's built-in staging array holds at most %d boxes per side; "
                  "use hf_box3d_iou with a caller workspace for larger sets",
                  num_a, num_b, HF_BOX3D_IOU_STATIC_MAX_BOXES);
        return HF_ERR_UNSUPPORTED;
    }
    if (num_a == 0 || num_b == 0) return HF_OK;
    HF_REQUIRE(boxes_a && boxes_b && iou_3d && iou_2d, "box3d_iou: null pointer");
    unsigned char *scratch = nullptr;
    int rc = static_scratch_of_current_device(&scratch);
    if (rc != HF_OK) return rc;
    // Whatever an earlier call left in the array is dead: the two conversions below write the num_a and num_b entries
    // that the IoU kernel then reads, in stream order.  Calls that overlap on the GPU would race on these bytes.
    float *bev_a = reinterpret_cast<float *>(scratch);
    float *bev_b = reinterpret_cast<float *>(scratch + kStaticSideBytes);
    rc = hf_boxes3d_to_bev(num_a, boxes_a, bev_a, stream);
    if (rc == HF_OK) rc = hf_boxes3d_to_bev(num_b, boxes_b, bev_b, stream);
    if (rc == HF_OK) rc = hf_compute_bev_iou(num_a, bev_a, num_b, bev_b, iou_3d, iou_2d, stream);  // iou_3d <- overlap
    if (rc != HF_OK) return rc;
    box3d_iou_kernel<<<grid_for((long long)num_a * num_b), 256, 0, (cudaStream_t)stream>>>(num_a, num_b, boxes_a, boxes_b, iou_3d);
    HF_LAUNCH_OK("box3d_iou_kernel");
    return HF_OK;
}
