// bev_iou.cu — rotated BEV overlap / IoU matrix and oriented NMS for sm_100a.
//
// Replaces bev_iou/bev_iou_g.cu:227-322 and the host half of OrientedNMSOp::Compute
// (bev_iou/bev_iou.cpp:74-116) of the reference (see include/hfops.h).
//
// Structure (not the reference's one-thread-per-pair / full N^2 / host sweep):
//   * per 64x64 tile, everything that depends on ONE box only (centre, sin/cos of +-angle,
//     the four rotated corners, a padded bounding radius) is computed once per box into
//     shared memory — the reference recomputes 2 cos + 2 sin per pair and 2 more per
//     corner test (:125-126, :50);
//   * pairs whose padded bounding circles are disjoint have overlap exactly 0 (no edge
//     crossing, no corner within the 1e-5 margin) and are settled by that test alone;
//   * the surviving pairs are compacted across the CTA (warp ballot + one shared counter)
//     so the expensive polygon clip runs with full warps instead of 1-2 live lanes;
//   * NMS: only the tiles on/above the diagonal are evaluated (the sweep never reads the
//     others, bev_iou.cpp:101-103); suppression bits are assembled with shared-memory
//     atomicOr; the greedy sweep runs on the device, one CTA per scene, 64 boxes per step
//     (a register-resident serial resolve of the diagonal word + a parallel OR of the kept
//     rows), so there is no 10 MB mask D2H copy, no host loop and no H2D copy.
// The clip itself follows the reference's arithmetic step by step (same candidate order,
// same strict/non-strict comparisons, same stable angular ordering, same fan), because
// the NMS keep list is a bit-exact output and depends on every comparison.
#include <math.h>

#include "common.cuh"

namespace hf {
namespace {

constexpr int kTile = 64;
constexpr int kBevThreads = 256;
constexpr float kEps = 1e-8f;     // bev_iou_g.cu:7
constexpr float kMargin = 1e-5f;  // bev_iou_g.cu:46

// Per-box precomputed data (floats): 0-4 box, 5-6 centre, 7-8 cos/sin(angle),
// 9-10 cos/sin(-angle), 11-18 rotated corners (x0,y0..x3,y3), 19 padded radius.
constexpr int kPre = 20;

__device__ __forceinline__ void precompute_box(const float *box, float *pre) {
    float x1 = box[0], y1 = box[1], x2 = box[2], y2 = box[3], ang = box[4];
    float cx = (x1 + x2) / 2, cy = (y1 + y2) / 2;
    float ca = cosf(ang), sa = sinf(ang);
    float cn = cosf(-ang), sn = sinf(-ang);
    pre[0] = x1; pre[1] = y1; pre[2] = x2; pre[3] = y2; pre[4] = ang;
    pre[5] = cx; pre[6] = cy; pre[7] = ca; pre[8] = sa; pre[9] = cn; pre[10] = sn;
    const float px[4] = {x1, x2, x2, x1};
    const float py[4] = {y1, y1, y2, y2};
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // rotate_around_center, :92-96
        pre[11 + 2 * k] = (px[k] - cx) * ca + (py[k] - cy) * sa + cx;
        pre[12 + 2 * k] = -(px[k] - cx) * sa + (py[k] - cy) * ca + cy;
    }
    float hx = x2 - x1, hy = y2 - y1;
    // half diagonal, padded far beyond the 1e-5 inclusion margin and fp32 corner error
    pre[19] = 0.5f * sqrtf(hx * hx + hy * hy) * 1.0001f + 1e-3f + 1e-5f * (fabsf(cx) + fabsf(cy));
}

// true if the two boxes cannot touch: overlap is exactly 0.
__device__ __forceinline__ bool circles_disjoint(const float *A, const float *B) {
    float dx = A[5] - B[5], dy = A[6] - B[6], r = A[19] + B[19];
    return dx * dx + dy * dy > r * r;
}

__device__ __forceinline__ float cross3(float p1x, float p1y, float p2x, float p2y, float p0x, float p0y) {
    return (p1x - p0x) * (p2y - p0y) - (p2x - p0x) * (p1y - p0y);  // :32-34
}

// Segment P (p0->p1) against segment Q (q0->q1); bev_iou_g.cu:61-90.
__device__ __forceinline__ bool seg_intersection(float p1x, float p1y, float p0x, float p0y, float q1x, float q1y,
                                                 float q0x, float q0y, float &ax, float &ay) {
    bool rc = fminf(p0x, p1x) <= fmaxf(q0x, q1x) && fminf(q0x, q1x) <= fmaxf(p0x, p1x) &&
              fminf(p0y, p1y) <= fmaxf(q0y, q1y) && fminf(q0y, q1y) <= fmaxf(p0y, p1y);
    if (!rc) return false;
    float s1 = cross3(q0x, q0y, p1x, p1y, p0x, p0y);
    float s2 = cross3(p1x, p1y, q1x, q1y, p0x, p0y);
    float s3 = cross3(p0x, p0y, q1x, q1y, q0x, q0y);
    float s4 = cross3(q1x, q1y, p1x, p1y, q0x, q0y);
    if (!(s1 * s2 > 0 && s3 * s4 > 0)) return false;
    float s5 = cross3(q1x, q1y, p1x, p1y, p0x, p0y);
    if (fabsf(s5 - s1) > kEps) {
        ax = (s5 * q0x - s1 * q1x) / (s5 - s1);
        ay = (s5 * q0y - s1 * q1y) / (s5 - s1);
    } else {
        float a0 = p0y - p1y, b0 = p1x - p0x, c0 = p0x * p1y - p1x * p0y;
        float a1 = q0y - q1y, b1 = q1x - q0x, c1 = q0x * q1y - q1x * q0y;
        float D = a0 * b1 - a1 * b0;
        ax = (b0 * c1 - b1 * c0) / D;
        ay = (a1 * c0 - a0 * c1) / D;
    }
    return true;
}

// check_in_box2d (:44-59) with the box's centre and cos/sin(-angle) precomputed.
__device__ __forceinline__ bool in_box(const float *B, float px, float py) {
    float cx = B[5], cy = B[6], c = B[9], s = B[10];
    float rx = (px - cx) * c + (py - cy) * s + cx;
    float ry = -(px - cx) * s + (py - cy) * c + cy;
    return rx > B[0] - kMargin && rx < B[2] + kMargin && ry > B[1] - kMargin && ry < B[3] + kMargin;
}

// box_overlap (:102-206) on precomputed boxes.
__device__ float overlap_pre(const float *A, const float *B) {
    float ptx[24], pty[24], ang[24];
    int cnt = 0;
    float sx = 0.f, sy = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int i1 = (i + 1) & 3;
        float a0x = A[11 + 2 * i], a0y = A[12 + 2 * i], a1x = A[11 + 2 * i1], a1y = A[12 + 2 * i1];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int j1 = (j + 1) & 3;
            float ix, iy;
            if (seg_intersection(a1x, a1y, a0x, a0y, B[11 + 2 * j1], B[12 + 2 * j1], B[11 + 2 * j], B[12 + 2 * j], ix, iy)) {
                sx = sx + ix;
                sy = sy + iy;
                ptx[cnt] = ix;
                pty[cnt] = iy;
                ++cnt;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float bx = B[11 + 2 * k], by = B[12 + 2 * k];
        if (in_box(A, bx, by)) {
            sx = sx + bx; sy = sy + by;
            ptx[cnt] = bx; pty[cnt] = by; ++cnt;
        }
        float ax = A[11 + 2 * k], ay = A[12 + 2 * k];
        if (in_box(B, ax, ay)) {
            sx = sx + ax; sy = sy + ay;
            ptx[cnt] = ax; pty[cnt] = ay; ++cnt;
        }
    }
    if (cnt == 0) return 0.f;  // reference: 0/0 centre, empty loops, fabs(0)/2
    float cx = sx / cnt, cy = sy / cnt;
    for (int k = 0; k < cnt; ++k) ang[k] = atan2f(pty[k] - cy, ptx[k] - cx);
    // bubble sort, swap on strict '>' (:98-100, :177-185)
    for (int j = 0; j < cnt - 1; ++j)
        for (int i = 0; i < cnt - j - 1; ++i)
            if (ang[i] > ang[i + 1]) {
                float t;
                t = ang[i]; ang[i] = ang[i + 1]; ang[i + 1] = t;
                t = ptx[i]; ptx[i] = ptx[i + 1]; ptx[i + 1] = t;
                t = pty[i]; pty[i] = pty[i + 1]; pty[i + 1] = t;
            }
    float area = 0.f;
    for (int k = 0; k < cnt - 1; ++k) {
        float ax = ptx[k] - ptx[0], ay = pty[k] - pty[0];
        float bx = ptx[k + 1] - ptx[0], by = pty[k + 1] - pty[0];
        area += ax * by - ay * bx;
    }
    return fabsf(area) / 2.0f;
}

__device__ __forceinline__ float iou_from_overlap(const float *A, const float *B, float s) {
    float sa = (A[2] - A[0]) * (A[3] - A[1]);
    float sb = (B[2] - B[0]) * (B[3] - B[1]);
    return s / fmaxf(sa + sb - s, kEps);  // :208-215
}

// Cheap-test all pairs of a 64x64 tile, compact the survivors into `list`.
// pair id e = i*64 + j.  diag: only j > i.  Returns the survivor count (after a barrier).
template <typename OnReject>
__device__ __forceinline__ int tile_filter(const float *rowpre, const float *colpre, int row_size, int col_size,
                                           bool diag, unsigned short *list, int *counter, OnReject on_reject) {
    const int t = threadIdx.x, lane = t & 31;
    if (t == 0) *counter = 0;
    __syncthreads();
#pragma unroll 4
    for (int e = t; e < kTile * kTile; e += kBevThreads) {
        int i = e >> 6, j = e & 63;
        bool valid = i < row_size && j < col_size && (!diag || j > i);
        bool live = false;
        if (valid) {
            live = !circles_disjoint(rowpre + i * kPre, colpre + j * kPre);
            if (!live) on_reject(i, j);
        }
        unsigned bal = __ballot_sync(0xffffffffu, live);
        if (bal) {
            int base = 0;
            if (lane == 0) base = atomicAdd(counter, __popc(bal));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (live) list[base + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)e;
        }
    }
    __syncthreads();
    return *counter;
}

// ---- NMS mask ---------------------------------------------------------------------
// grid.x enumerates tiles: upper_only -> the cb*(cb+1)/2 tiles with col >= row, else all
// cb*cb; grid.y = scene.
__global__ void __launch_bounds__(kBevThreads)
    nms_mask_kernel(int boxes_num, float thresh, const float *__restrict__ boxes,
                    unsigned long long *__restrict__ mask, int upper_only) {
    __shared__ float rowpre[kTile * kPre];
    __shared__ float colpre[kTile * kPre];
    __shared__ unsigned short list[kTile * kTile];
    __shared__ unsigned long long smask[kTile];
    __shared__ int counter;

    const int cb = (boxes_num + kTile - 1) / kTile;
    int row_start, col_start;
    if (upper_only) {
        // invert L = r*cb - r*(r-1)/2 + (c - r)
        int L = blockIdx.x;
        float fb = 2.f * cb + 1.f;
        int r = (int)((fb - sqrtf(fb * fb - 8.f * (float)L)) * 0.5f);
        if (r < 0) r = 0;
        if (r > cb - 1) r = cb - 1;
        while (r > 0 && (long long)r * cb - (long long)r * (r - 1) / 2 > L) --r;
        while ((long long)(r + 1) * cb - (long long)(r + 1) * r / 2 <= L) ++r;
        row_start = r;
        col_start = r + (L - (r * cb - r * (r - 1) / 2));
    } else {
        row_start = blockIdx.x / cb;
        col_start = blockIdx.x - row_start * cb;
    }
    const float *sb = boxes + (size_t)blockIdx.y * boxes_num * 5;
    unsigned long long *sm = mask + (size_t)blockIdx.y * boxes_num * cb;
    const int row_size = min(boxes_num - row_start * kTile, kTile);
    const int col_size = min(boxes_num - col_start * kTile, kTile);
    const int t = threadIdx.x;

    if (t < kTile) {
        if (t < col_size) precompute_box(sb + (size_t)(col_start * kTile + t) * 5, colpre + t * kPre);
        smask[t] = 0ull;
    } else if (t < 2 * kTile) {
        int i = t - kTile;
        if (i < row_size) precompute_box(sb + (size_t)(row_start * kTile + i) * 5, rowpre + i * kPre);
    }
    __syncthreads();
    const bool diag = row_start == col_start;
    int ns = tile_filter(rowpre, colpre, row_size, col_size, diag, list, &counter, [](int, int) {});
    for (int q = t; q < ns; q += kBevThreads) {
        int e = list[q], i = e >> 6, j = e & 63;
        const float *A = rowpre + i * kPre, *B = colpre + j * kPre;
        float s = overlap_pre(A, B);
        if (iou_from_overlap(A, B, s) > thresh) atomicOr(&smask[i], 1ull << j);
    }
    __syncthreads();
    if (t < row_size) sm[(size_t)(row_start * kTile + t) * cb + col_start] = smask[t];
}

// ---- greedy sweep -----------------------------------------------------------------
// One CTA per scene.  bev_iou.cpp:87-112.
constexpr int kSweepThreads = 256;
__global__ void __launch_bounds__(kSweepThreads)
    nms_sweep_kernel(int boxes_num, const unsigned long long *__restrict__ mask, int *__restrict__ keep,
                     int *__restrict__ num_kept) {
    extern __shared__ unsigned long long remv[];  // [cb]
    __shared__ unsigned long long sdiag[kTile];
    __shared__ unsigned long long skeep;
    const int cb = (boxes_num + kTile - 1) / kTile;
    const int scene = blockIdx.x, t = threadIdx.x;
    const unsigned long long *sm = mask + (size_t)scene * boxes_num * cb;
    int *kp = keep + (size_t)scene * boxes_num;
    for (int c = t; c < cb; c += kSweepThreads) remv[c] = 0ull;
    int kept = 0;
    __syncthreads();
    for (int nb = 0; nb < cb; ++nb) {
        const int size = min(boxes_num - nb * kTile, kTile);
        if (t < kTile) sdiag[t] = t < size ? sm[(size_t)(nb * kTile + t) * cb + nb] : 0ull;
        __syncthreads();
        if (t == 0) {
            unsigned long long d[kTile];
#pragma unroll
            for (int i = 0; i < kTile; ++i) d[i] = sdiag[i];
            unsigned long long word = remv[nb];
            if (size < kTile) word |= ~0ull << size;  // boxes past the end are not candidates
            unsigned long long kb = 0ull;
#pragma unroll
            for (int i = 0; i < kTile; ++i) {
                bool alive = !((word >> i) & 1ull);
                kb |= alive ? (1ull << i) : 0ull;
                word |= alive ? d[i] : 0ull;
            }
            skeep = kb;
        }
        __syncthreads();
        const unsigned long long kb = skeep;
        // kept indices, ascending
        if (t < kTile && ((kb >> t) & 1ull)) kp[kept + __popcll(kb & ((1ull << t) - 1ull))] = nb * kTile + t;
        kept += __popcll(kb);
        // OR the kept rows into the removal words of the later column blocks
        for (int c = nb + 1 + t; c < cb; c += kSweepThreads) {
            unsigned long long acc = 0ull, bits = kb;
            while (bits) {
                int i = __ffsll((long long)bits) - 1;
                bits &= bits - 1ull;
                acc |= sm[(size_t)(nb * kTile + i) * cb + c];
            }
            remv[c] |= acc;
        }
        __syncthreads();
    }
    // pad with keep[0] (:110-112)
    __syncthreads();
    const int first = kp[0];
    for (int p = kept + t; p < boxes_num; p += kSweepThreads) kp[p] = first;
    if (num_kept && t == 0) num_kept[scene] = kept;
}

// ---- IoU matrix -------------------------------------------------------------------
__global__ void __launch_bounds__(kBevThreads)
    bev_iou_matrix_kernel(int num_a, const float *__restrict__ boxes_a, int num_b, const float *__restrict__ boxes_b,
                          float *__restrict__ ans_overlap, float *__restrict__ ans_iou) {
    __shared__ float rowpre[kTile * kPre];
    __shared__ float colpre[kTile * kPre];
    __shared__ unsigned short list[kTile * kTile];
    __shared__ int counter;
    const int row_start = blockIdx.y, col_start = blockIdx.x;
    const int row_size = min(num_a - row_start * kTile, kTile);
    const int col_size = min(num_b - col_start * kTile, kTile);
    const int t = threadIdx.x;
    if (t < kTile) {
        if (t < col_size) precompute_box(boxes_b + (size_t)(col_start * kTile + t) * 5, colpre + t * kPre);
    } else if (t < 2 * kTile) {
        int i = t - kTile;
        if (i < row_size) precompute_box(boxes_a + (size_t)(row_start * kTile + i) * 5, rowpre + i * kPre);
    }
    __syncthreads();
    auto at = [&](int i, int j) { return (size_t)(row_start * kTile + i) * num_b + (size_t)col_start * kTile + j; };
    int ns = tile_filter(rowpre, colpre, row_size, col_size, false, list, &counter, [&](int i, int j) {
        size_t o = at(i, j);
        ans_overlap[o] = 0.f;
        ans_iou[o] = 0.f;
    });
    for (int q = t; q < ns; q += kBevThreads) {
        int e = list[q], i = e >> 6, j = e & 63;
        const float *A = rowpre + i * kPre, *B = colpre + j * kPre;
        float s = overlap_pre(A, B);
        size_t o = at(i, j);
        ans_overlap[o] = s;
        ans_iou[o] = iou_from_overlap(A, B, s);
    }
}

}  // namespace
}  // namespace hf

extern "C" int hf_compute_bev_iou(int num_a, const float *boxes_a, int num_b, const float *boxes_b,
                                  float *ans_overlap, float *ans_iou, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(num_a > 0, "ComputeIOU3D expects (N, 5) proposals shape");
    HF_REQUIRE(num_b > 0, "ComputeIOU3D expects (M, 5) gt_bboxes shape");
    HF_REQUIRE(boxes_a && boxes_b && ans_overlap && ans_iou, "ComputeBevIOU: null pointer");
    dim3 grid((num_b + kTile - 1) / kTile, (num_a + kTile - 1) / kTile);
    HF_REQUIRE(grid.y <= 65535, "ComputeBevIOU: more than 65535*64 proposals not supported");
    bev_iou_matrix_kernel<<<grid, kBevThreads, 0, (cudaStream_t)stream>>>(num_a, boxes_a, num_b, boxes_b, ans_overlap, ans_iou);
    HF_LAUNCH_OK("bev_iou_matrix_kernel");
    return HF_OK;
}

// zero_lower: also clear the never-evaluated tiles below the diagonal (public mask API);
// the fused NMS path skips that, its sweep only reads column blocks >= the row's own.
static int launch_mask(const float *boxes, unsigned long long *mask, int scenes, int boxes_num, float thresh,
                       int upper_only, bool zero_lower, cudaStream_t s) {
    using namespace hf;
    const long long cb = (boxes_num + kTile - 1) / kTile;
    if (upper_only && zero_lower) HF_CUDA_OK(cudaMemsetAsync(mask, 0, (size_t)scenes * boxes_num * cb * sizeof(unsigned long long), s));
    long long tiles = upper_only ? cb * (cb + 1) / 2 : cb * cb;
    HF_REQUIRE(tiles < (1ll << 31) && scenes <= 65535, "OrientedNMS: too many boxes / scenes");
    nms_mask_kernel<<<dim3((unsigned)tiles, scenes), kBevThreads, 0, s>>>(boxes_num, thresh, boxes, mask, upper_only);
    HF_LAUNCH_OK("nms_mask_kernel");
    return HF_OK;
}

extern "C" int hf_oriented_nms_mask(const float *boxes, unsigned long long *mask, int boxes_num,
                                    float nms_overlap_thresh, int upper_only, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(nms_overlap_thresh >= 0, "Need nms_threshold >= 0");
    HF_REQUIRE(boxes_num > 0, "OrientendNMS expects (N, 5) boxes shape");
    HF_REQUIRE(boxes && mask, "OrientedNMS: null pointer");
    return launch_mask(boxes, mask, 1, boxes_num, nms_overlap_thresh, upper_only, true, (cudaStream_t)stream);
}

extern "C" size_t hf_oriented_nms_workspace_bytes(int scenes, int boxes_num) {
    if (scenes <= 0 || boxes_num <= 0) return 0;
    size_t cb = ((size_t)boxes_num + hf::kTile - 1) / hf::kTile;
    return (size_t)scenes * boxes_num * cb * sizeof(unsigned long long);
}

extern "C" int hf_oriented_nms(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int *keep,
                               int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(nms_overlap_thresh >= 0, "Need nms_threshold >= 0");
    HF_REQUIRE(boxes_num > 0 && scenes >= 0, "OrientendNMS expects (N, 5) boxes shape");
    HF_REQUIRE(boxes && keep, "OrientedNMS: null pointer");
    if (scenes == 0) return HF_OK;
    size_t need = hf_oriented_nms_workspace_bytes(scenes, boxes_num);
    if (!workspace || workspace_bytes < need) {
        set_error("OrientedNMS: workspace of %zu bytes required, got %zu", need, workspace ? workspace_bytes : (size_t)0);
        return HF_ERR_WORKSPACE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    unsigned long long *mask = static_cast<unsigned long long *>(workspace);
    int rc = launch_mask(boxes, mask, scenes, boxes_num, nms_overlap_thresh, 1, false, s);
    if (rc != HF_OK) return rc;
    const int cb = (boxes_num + kTile - 1) / kTile;
    size_t smem = (size_t)cb * sizeof(unsigned long long);
    if (smem > 200 * 1024) {
        set_error("OrientedNMS: %d boxes exceed the sweep kernel's shared memory", boxes_num);
        return HF_ERR_UNSUPPORTED;
    }
    if (smem > 40 * 1024)
        HF_CUDA_OK(cudaFuncSetAttribute(nms_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nms_sweep_kernel<<<scenes, kSweepThreads, smem, s>>>(boxes_num, mask, keep, num_kept);
    HF_LAUNCH_OK("nms_sweep_kernel");
    return HF_OK;
}
