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
//     rows), so there is no 10 MB mask D2H copy, no host loop and no H2D copy;
//   * the fused NMS entry (hf_oriented_nms) goes further — see "Fused NMS path" below.
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
__device__ __forceinline__ float overlap_pre(const float *A, const float *B) {
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

// ===================================================================================
// Fused NMS path (hf_oriented_nms): prepare -> filter -> clip -> sweep.
//
// The 64x64-tile kernel above spends its time in the polygon clip of every pair whose
// bounding circles touch (7% of the pairs of a 9000-proposal scene).  For the
// *thresholded* question "iou > thresh" far fewer pairs need the exact clip:
//   sort     one CTA per scene: Morton code of every box centre, bitonic sort in shared memory
//            -> a spatial permutation.  Score order matters only to the sweep; WHICH pairs
//            overlap does not depend on the order they are enumerated in, and in Morton order
//            a tile of 64 consecutive boxes is spatially compact;
//   prepare  one thread per box, in sorted order: the per-box record of precompute_box
//            (sin/cos, rotated corners, padded radius) + a compact 8-float "lite" record,
//            written once per scene instead of once per tile, + per 64-box tile the bounding
//            rectangle of its padded circles;
//   filter   per row tile: column tiles whose bounding rectangle is disjoint from the row
//            tile's cannot hold a touching pair and are skipped whole (~95% of the tile pairs
//            of a 9000-proposal scene).  For the others: stage 1 bounding circles (column box in
//            registers, row boxes broadcast from shared memory, warp-private ballot
//            compaction, no atomics); stage 2 a separating-axis UPPER BOUND on the overlap
//            (intersection is contained in A cut by B's extent along A's axes and vice
//            versa) turned into an upper bound on the IoU — pairs whose bound is safely
//            below the threshold have bit 0 without clipping.  Survivors are appended to a
//            global pair list (one atomicAdd per CTA flush);
//   clip     one thread per surviving pair, full warps, both records in registers; exact
//            reference arithmetic (overlap_pre) with the box of lower ORIGINAL index as the
//            reference's `a` (the clip is not bit-symmetric in its arguments); atomicOr of
//            the suppression bit at the pair's original (score-order) indices;
//   sweep    see nms_sweep_kernel2.
// The bound only ever removes pairs whose exact IoU is below the threshold by a margin
// (1e-3 relative + float slack), so the mask — and the keep list — are unchanged.
// ===================================================================================
constexpr int kRec = 24;    // full record: kPre floats + hx, hy, area, pad (96 B, float4 aligned)
constexpr int kLite = 8;    // cx, cy, radius, hx, hy, cos, sin, mb (see lite_area / cheap_passes)

constexpr int kSortThreads = 1024;
constexpr int kSortCells = 4096;  // 64 x 64 grid over the bounding rectangle of the centres, Morton order

__device__ __forceinline__ uint32_t spread6(uint32_t v) {  // 6 bits -> every second bit
    v &= 63u;
    v = (v | (v << 4)) & 0x0F0Fu;
    v = (v | (v << 2)) & 0x3333u;
    v = (v | (v << 1)) & 0x5555u;
    return v;
}

// perm[scene][p] = original index of the box at sorted position p: a counting sort of the box
// centres into a 64 x 64 grid walked in Morton order (one CTA per scene, any boxes_num).
// Positions inside one cell are handed out by atomicAdd, so the permutation may differ from
// run to run; the NMS result does not depend on it (see clip_pair).  tmp: boxes_num u64 per scene.
__global__ void __launch_bounds__(kSortThreads)
    nms_sort_kernel(int boxes_num, const float *__restrict__ boxes, int *__restrict__ perm,
                    unsigned long long *__restrict__ tmp_all) {
    __shared__ int hist[kSortCells];
    __shared__ float red[4][32];
    __shared__ int wsum[32];
    const int scene = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const float *sb = boxes + (size_t)scene * boxes_num * 5;
    unsigned long long *tmp = tmp_all + (size_t)scene * boxes_num;
    int *pp = perm + (size_t)scene * boxes_num;
    // ---- bounding rectangle of the centres (fminf/fmaxf skip NaN) ----
    float x0 = INFINITY, y0 = INFINITY, x1 = -INFINITY, y1 = -INFINITY;
    for (int i = t; i < boxes_num; i += kSortThreads) {
        const float cx = (sb[i * 5] + sb[i * 5 + 2]) * 0.5f, cy = (sb[i * 5 + 1] + sb[i * 5 + 3]) * 0.5f;
        x0 = fminf(x0, cx); y0 = fminf(y0, cy);
        x1 = fmaxf(x1, cx); y1 = fmaxf(y1, cy);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x0 = fminf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
        y0 = fminf(y0, __shfl_xor_sync(0xffffffffu, y0, o));
        x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
        y1 = fmaxf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
    }
    if (lane == 0) {
        red[0][warp] = x0; red[1][warp] = y0; red[2][warp] = x1; red[3][warp] = y1;
    }
    for (int c = t; c < kSortCells; c += kSortThreads) hist[c] = 0;
    __syncthreads();
    x0 = red[0][lane]; y0 = red[1][lane]; x1 = red[2][lane]; y1 = red[3][lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        x0 = fminf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
        y0 = fminf(y0, __shfl_xor_sync(0xffffffffu, y0, o));
        x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
        y1 = fmaxf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
    }
    const float ext = fmaxf(x1 - x0, y1 - y0);  // square cells
    const float inv = (ext > 0.f && ext < 1e30f) ? 63.999f / ext : 0.f;
    // ---- cell of every box, rank inside the cell ----
    for (int i = t; i < boxes_num; i += kSortThreads) {
        const float cx = (sb[i * 5] + sb[i * 5 + 2]) * 0.5f, cy = (sb[i * 5 + 1] + sb[i * 5 + 3]) * 0.5f;
        const uint32_t qx = min(__float2uint_rz(fmaxf((cx - x0) * inv, 0.f)), 63u);  // NaN -> 0
        const uint32_t qy = min(__float2uint_rz(fmaxf((cy - y0) * inv, 0.f)), 63u);
        const uint32_t cell = spread6(qx) | (spread6(qy) << 1);
        const int r = atomicAdd(&hist[cell], 1);
        tmp[i] = ((unsigned long long)cell << 32) | (unsigned)r;
    }
    __syncthreads();
    // ---- exclusive scan of the 4096 counts: 4 per thread, warp scan, scan of the warp sums ----
    int c0 = hist[4 * t], c1 = hist[4 * t + 1], c2 = hist[4 * t + 2], c3 = hist[4 * t + 3];
    const int mine = c0 + c1 + c2 + c3;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int wincl = wsum[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, wincl, o);
        if (lane >= o) wincl += v;
    }
    const int wbase = __shfl_sync(0xffffffffu, wincl, warp) - wsum[warp];
    const int e0 = wbase + incl - mine;
    __syncthreads();  // every thread has read its counts
    hist[4 * t] = e0;
    hist[4 * t + 1] = e0 + c0;
    hist[4 * t + 2] = e0 + c0 + c1;
    hist[4 * t + 3] = e0 + c0 + c1 + c2;
    __syncthreads();
    for (int i = t; i < boxes_num; i += kSortThreads) {
        const unsigned long long v = tmp[i];
        pp[hist[(uint32_t)(v >> 32)] + (int)(uint32_t)v] = i;
    }
}

// One 64-thread CTA per tile of 64 boxes in sorted order (perm == nullptr: identity).
// grid = (cb, scenes).  tile_box = bounding rectangle of the tile's padded circles, half_box =
// the same for each half (32 boxes) of the tile: a tile that straddles two clusters is large,
// its halves mostly are not.
__global__ void __launch_bounds__(kTile)
    nms_prepare_kernel(int boxes_num, const float *__restrict__ boxes, const int *__restrict__ perm,
                       float *__restrict__ rec, float *__restrict__ lite, float4 *__restrict__ tile_box,
                       float4 *__restrict__ half_box, float pad_scale) {
    __shared__ float4 wbox[2];
    const int cb = (boxes_num + kTile - 1) / kTile;
    const int scene = blockIdx.y, tile = blockIdx.x, t = threadIdx.x;
    const int p = tile * kTile + t;
    float bx0 = INFINITY, by0 = INFINITY, bx1 = -INFINITY, by1 = -INFINITY;
    if (p < boxes_num) {
        const size_t g = (size_t)scene * boxes_num + p;
        const int src = perm ? perm[g] : p;
        float pre[kRec];
        precompute_box(boxes + ((size_t)scene * boxes_num + src) * 5, pre);
        const float hx = (pre[2] - pre[0]) * 0.5f, hy = (pre[3] - pre[1]) * 0.5f;
        pre[20] = hx;
        pre[21] = hy;
        pre[22] = (pre[2] - pre[0]) * (pre[3] - pre[1]);
        pre[23] = 0.f;
        float4 *o = reinterpret_cast<float4 *>(rec + g * kRec);
#pragma unroll
        for (int k = 0; k < kRec / 4; ++k) o[k] = make_float4(pre[4 * k], pre[4 * k + 1], pre[4 * k + 2], pre[4 * k + 3]);
        float4 *l = reinterpret_cast<float4 *>(lite + g * kLite);
        l[0] = make_float4(pre[5], pre[6], pre[19], hx);
        l[1] = make_float4(hy, pre[7], pre[8], 5e-5f + 4e-6f * (fabsf(pre[5]) + fabsf(pre[6]) + pre[19]));
        // thresh >= 0.5: centres of a passing pair are within rscale * min(r_a, r_b) <= rscale * (r_a + r_b)/2
        bx0 = pre[5] - pre[19] * pad_scale;
        by0 = pre[6] - pre[19] * pad_scale;
        bx1 = pre[5] + pre[19] * pad_scale;
        by1 = pre[6] + pre[19] * pad_scale;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        bx0 = fminf(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
        by0 = fminf(by0, __shfl_xor_sync(0xffffffffu, by0, o));
        bx1 = fmaxf(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
        by1 = fmaxf(by1, __shfl_xor_sync(0xffffffffu, by1, o));
    }
    if ((t & 31) == 0) {
        wbox[t >> 5] = make_float4(bx0, by0, bx1, by1);
        half_box[((size_t)scene * cb + tile) * 2 + (t >> 5)] = make_float4(bx0, by0, bx1, by1);  // boxes 32h .. 32h+31 of the tile
    }
    __syncthreads();
    if (t == 0) {
        const float4 a = wbox[0], b = wbox[1];
        tile_box[(size_t)scene * cb + tile] =
            make_float4(fminf(a.x, b.x), fminf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
    }
}

// How far apart the centres of a pair with iou > tq >= 0.5 can be, as a fraction of the smaller
// of the two bounding radii.  From the reflection bound (cheap_passes (3)) and the area-ratio
// bound (1): 4 (ahx-|du|)(ahy-|dv|) >= 2|I| - |B| >= |A| (2 tq - 1)/tq, so |du| <= ahx (1 - g) and
// |dv| <= ahy (1 - g) with g = (2 tq - 1)/tq, i.e. |offset| <= (1 - g) * half-diagonal, for both
// boxes.  tq = 0.5 gives 1 (each centre inside the other box), tq = 0.85 gives 0.18.
__host__ __device__ __forceinline__ float nms_rscale(float tq) { return tq >= 0.5f ? (1.f - tq) / tq : 1.f; }

// Lite record (8 floats, 32 B): cx, cy, padded radius, hx, hy, cos, sin, mb.
// mb = half of the slack a test on a pair needs for float error of the projections
// (coordinates ~1e2 m -> ~1e-5) and the reference's 1e-5 inclusion margin: m = mb_a + mb_b.
__device__ __forceinline__ float lite_area(const float *L) { return 4.f * L[3] * L[4]; }

// Rotation-free necessary conditions for iou > thresh (stage 1).  a*/b*: the two lite records,
// (dx,dy) = B's centre - A's; tq = thresh minus margin (> 0); tight = tq >= 0.5.
// Every comparison is written so that NaN input PASSES (the exact clip then decides).
//  (1) iou <= min area / max area.
//  (2) tight: the intersection is a convex subset of each (centrally symmetric) box holding more
//      than half of its area, so it contains both centres: each centre lies inside the other box.
//  (3) reflection bound: with I = A n B and s the point reflection about B's centre, s(I) is inside
//      B too, so |I n s(I)| >= 2|I| - |B|; and I n s(I) lies in A n s(A), where s(A) is A shifted by
//      twice the centre offset: |A n s(A)| = 4 (ahx-|du|)+ (ahy-|dv|)+ in A's frame.  Hence
//      |I| <= (|B| + 4 (ahx-|du|)+ (ahy-|dv|)+) / 2, and the same with the roles swapped.
__device__ __forceinline__ bool cheap_passes(float dx, float dy, float ahx, float ahy, float aca, float asa, float aar,
                                             float bhx, float bhy, float bca, float bsa, float bar, float m, float tq,
                                             bool tight) {
    // centre offset in A's frame (local x axis = (cos,-sin), local y axis = (sin,cos); :92-96) and in B's
    const float du = fabsf(dx * aca - dy * asa), dv = fabsf(dx * asa + dy * aca);
    const float du2 = fabsf(dx * bca - dy * bsa), dv2 = fabsf(dx * bsa + dy * bca);
    bool ok = !(fminf(aar, bar) * 1.00001f + m < tq * fmaxf(aar, bar));
    if (tight) ok = ok && !(du > ahx + m || dv > ahy + m || du2 > bhx + m || dv2 > bhy + m);
    const float ra = 4.f * fmaxf(ahx - du + m, 0.f) * fmaxf(ahy - dv + m, 0.f);
    const float rb = 4.f * fmaxf(bhx - du2 + m, 0.f) * fmaxf(bhy - dv2 + m, 0.f);
    const float ubr = fminf(0.5f * (bar + ra), 0.5f * (aar + rb)) * 1.00001f + m;
    const float dr = aar + bar - ubr;
    return ok && !(dr > 0.f && ubr < tq * dr);
}

// Separating-axis upper bound (stage 2).  A, B: lite records.  true = the exact test iou > thresh
// is certainly false (overlap bound too small), so the pair needs no clip.
__device__ __forceinline__ bool sat_rejects(const float *A, const float *B, float tq) {
    const float acx = A[0], acy = A[1], ahx = A[3], ahy = A[4], aca = A[5], asa = A[6], aar = lite_area(A);
    const float bcx = B[0], bcy = B[1], bhx = B[3], bhy = B[4], bca = B[5], bsa = B[6], bar = lite_area(B);
    if (!(ahx > 0.f && ahy > 0.f && bhx > 0.f && bhy > 0.f)) return false;  // degenerate input: always clip
    const float dx = bcx - acx, dy = bcy - acy;
    const float m = A[7] + B[7];
    const float cd = fabsf(aca * bca + asa * bsa), sd = fabsf(asa * bca - aca * bsa);
    // in A's frame: the intersection is contained in A cut by B's extent along A's axes
    float du = dx * aca - dy * asa, dv = dx * asa + dy * aca;
    float eu = bhx * cd + bhy * sd + m, ev = bhx * sd + bhy * cd + m;
    float ou = fminf(ahx, du + eu) - fmaxf(-ahx, du - eu);
    float ov = fminf(ahy, dv + ev) - fmaxf(-ahy, dv - ev);
    if (ou <= 0.f || ov <= 0.f) return true;  // separated along one of A's axes: overlap is exactly 0
    // in B's frame
    float du2 = -(dx * bca - dy * bsa), dv2 = -(dx * bsa + dy * bca);
    float eu2 = ahx * cd + ahy * sd + m, ev2 = ahx * sd + ahy * cd + m;
    float ou2 = fminf(bhx, du2 + eu2) - fmaxf(-bhx, du2 - eu2);
    float ov2 = fminf(bhy, dv2 + ev2) - fmaxf(-bhy, dv2 - ev2);
    if (ou2 <= 0.f || ov2 <= 0.f) return true;
    if (!(tq > 0.f)) return false;
    float ub = fminf(fminf(ou * ov, ou2 * ov2), fminf(aar, bar) * 1.00001f + m);
    float denom = aar + bar - ub;
    if (!(denom > 0.f)) return false;
    return ub < tq * denom;  // iou <= ub/denom < tq <= thresh - margin
}

// Liang-Barsky: parameter range [t0,t1] of the segment p + t*d, t in [0,1], inside the
// axis-aligned rectangle [-hx,hx] x [-hy,hy].  Empty when t0 >= t1.
__device__ __forceinline__ void clip_to_rect(float px, float py, float dx, float dy, float hx, float hy, float &t0,
                                             float &t1) {
    // a (nearly) parallel segment gets a huge finite slope: inside -> (-huge, +huge), outside -> empty
    const float ix = fabsf(dx) > 1e-12f ? __frcp_rn(dx) : 1e12f;
    const float iy = fabsf(dy) > 1e-12f ? __frcp_rn(dy) : 1e12f;
    const float ax = (-hx - px) * ix, bx = (hx - px) * ix;
    const float ay = (-hy - py) * iy, by = (hy - py) * iy;
    t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), 0.f);
    t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), 1.f);
}

// Area of the intersection of two rotated rectangles by Green's theorem: the boundary of A n B
// is made of the parts of A's edges inside B and the parts of B's edges inside A; each part is
// one segment (convexity) found by clipping against an axis-aligned rectangle in the other
// box's own frame, and contributes cross(P, Q)/2, all evaluated in A's frame.  Coincident edges
// are counted twice, i.e. the result can only OVER-estimate — which is the safe side for its
// one use, an upper bound that spares the exact (reference-arithmetic) clip.  A, B: lite records.
__device__ __forceinline__ float green_overlap(const float *A, const float *B) {
    const float ahx = A[3], ahy = A[4], aca = A[5], asa = A[6];
    const float bhx = B[3], bhy = B[4], bca = B[5], bsa = B[6];
    const float ox = B[0] - A[0], oy = B[1] - A[1];  // world offset of B's centre from A's
    // relative rotation: B-local -> world is [[c,s],[-s,c]], world -> A-local is its transpose for A
    const float rc = bca * aca + bsa * asa, rs = bsa * aca - bca * asa;
    // B-local (u,v) -> A-local: ( u*rc + v*rs + tx, -u*rs + v*rc + ty ),  t = world->A-local of the offset
    const float tx = ox * aca - oy * asa, ty = ox * asa + oy * aca;
    // A-local (u,v) -> B-local: inverse rotation, and B's centre maps to the origin
    const float sx = -(tx * rc - ty * rs), sy = -(tx * rs + ty * rc);
    const float su[4] = {-1.f, 1.f, 1.f, -1.f}, sv[4] = {-1.f, -1.f, 1.f, 1.f};  // counter-clockwise corners
    float acc = 0.f;
    float bax[4], bay[4], abx[4], aby[4];  // B's corners in A's frame, A's corners in B's frame
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float u = su[k] * bhx, v = sv[k] * bhy;
        bax[k] = u * rc + v * rs + tx;
        bay[k] = -u * rs + v * rc + ty;
        const float p = su[k] * ahx, q = sv[k] * ahy;
        abx[k] = p * rc - q * rs + sx;
        aby[k] = p * rs + q * rc + sy;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int k1 = (k + 1) & 3;
        float t0, t1;
        // edge k of B, clipped by A (in A's frame), integrated in A's frame
        float dx = bax[k1] - bax[k], dy = bay[k1] - bay[k];
        clip_to_rect(bax[k], bay[k], dx, dy, ahx, ahy, t0, t1);
        if (t1 > t0) {
            const float px = bax[k] + t0 * dx, py = bay[k] + t0 * dy, qx = bax[k] + t1 * dx, qy = bay[k] + t1 * dy;
            acc += px * qy - py * qx;
        }
        // edge k of A, clipped by B (parameters from B's frame), integrated in A's frame
        dx = abx[k1] - abx[k];
        dy = aby[k1] - aby[k];
        clip_to_rect(abx[k], aby[k], dx, dy, bhx, bhy, t0, t1);
        if (t1 > t0) {
            const float ex = (su[k1] - su[k]) * ahx, ey = (sv[k1] - sv[k]) * ahy;
            const float px = su[k] * ahx + t0 * ex, py = sv[k] * ahy + t0 * ey, qx = su[k] * ahx + t1 * ex, qy = sv[k] * ahy + t1 * ey;
            acc += px * qy - py * qx;
        }
    }
    return 0.5f * acc;
}

// true = the exact test iou > thresh is certainly false: even an over-estimate of the true
// intersection area (+1e-4 relative, +1e-4 m^2 for float error) stays below the threshold.
// The reference's own polygon (edge crossings + contained corners) lies inside the true
// intersection up to its 1e-5 margins, so its area cannot exceed the bound either.
__device__ __forceinline__ bool area_rejects(const float *A, const float *B, float tq) {
    if (!(tq > 0.f) || !(A[3] > 0.f && A[4] > 0.f && B[3] > 0.f && B[4] > 0.f)) return false;
    const float est = green_overlap(A, B);
    if (!(est == est)) return false;  // NaN: let the exact clip decide
    const float aar = lite_area(A), bar = lite_area(B);
    const float ub = fminf(est * 1.0001f + 1e-4f, fminf(aar, bar) * 1.00001f + 1e-4f);
    const float denom = aar + bar - ub;
    if (!(denom > 0.f)) return false;
    return ub < tq * denom;
}

// Exact test of one pair on full records held in registers; sets the mask bit.  row/col are
// SORTED positions; the reference's (a, b) = (lower, higher) ORIGINAL index, and the bit lives
// at mask[a][b].
__device__ __forceinline__ void clip_pair(const float *__restrict__ rec, const int *__restrict__ perm,
                                          unsigned long long *__restrict__ mask, int cb, int row, int col, float thresh) {
    int oa = perm ? __ldg(perm + row) : row, ob = perm ? __ldg(perm + col) : col;
    if (oa > ob) {
        int tmp = oa; oa = ob; ob = tmp;
        tmp = row; row = col; col = tmp;
    }
    float a[kRec], b[kRec];
    const float4 *pa = reinterpret_cast<const float4 *>(rec + (size_t)row * kRec);
    const float4 *pb = reinterpret_cast<const float4 *>(rec + (size_t)col * kRec);
#pragma unroll
    for (int k = 0; k < 5; ++k) {  // floats 0..19 are all the clip reads
        float4 va = __ldg(pa + k), vb = __ldg(pb + k);
        a[4 * k] = va.x; a[4 * k + 1] = va.y; a[4 * k + 2] = va.z; a[4 * k + 3] = va.w;
        b[4 * k] = vb.x; b[4 * k + 1] = vb.y; b[4 * k + 2] = vb.z; b[4 * k + 3] = vb.w;
    }
    float s = overlap_pre(a, b);
    if (iou_from_overlap(a, b, s) > thresh) atomicOr(&mask[(size_t)oa * cb + (ob >> 6)], 1ull << (ob & 63));
}

__device__ __noinline__ void clip_pair_overflow(const float *rec, const int *perm, unsigned long long *mask, int cb,
                                                int row, int col, float thresh) {
    clip_pair(rec, perm, mask, cb, row, col, thresh);
}

struct NmsLists {
    unsigned long long *pairs;  // [cap] (scene << 48) | (row << 24) | col, sorted positions
    unsigned long long *count;  // pairs appended so far (may exceed cap: the excess was clipped inline)
    unsigned long long cap;
};

// grid = (cb, scenes); CTA rb handles row tile rb (sorted order) against the column tiles
// ct >= rb whose bounding rectangle touches its own.  After the tile list is built the eight
// warps run without any CTA barrier: warp w owns the block (rows 16*(w>>1) .. +16) x (columns
// 32*(w&1) .. +32) of every listed tile, keeps its 32 column records in registers (one per
// lane, straight from global memory) and reads the row records as shared-memory broadcasts.
//   stage 1  centre distance (see nms_rscale), ballot-compacted into a warp-private list;
//   stage 2  rotation-free bounds (cheap_passes), then the separating-axis bound, full warps;
//   stage 3  exact-area bound (green_overlap) on what is left, batched 32 at a time through a
//            warp-private queue; the survivors go to the global pair list for the exact clip.
constexpr int kTileListCap = 256;
constexpr int kWarpQueue = 64;  // <= 31 left over + 32 from one stage-2 batch
__global__ void __launch_bounds__(256, 3)
    nms_filter_kernel(int boxes_num, float thresh, const float *__restrict__ rec_all, const float *__restrict__ lite_all,
                      const float4 *__restrict__ tile_box_all, const float4 *__restrict__ half_box_all,
                      const int *__restrict__ perm_all, unsigned long long *__restrict__ mask_all, NmsLists lists) {
    __shared__ __align__(16) float rowl[kTile * kLite];
    __shared__ unsigned short list1[8][512];
    __shared__ unsigned int wq[8][kWarpQueue];
    __shared__ int tiles[kTileListCap];
    __shared__ float4 rsubs[4];
    __shared__ int ntiles;

    const int cb = (boxes_num + kTile - 1) / kTile;
    const int rb = blockIdx.x, scene = blockIdx.y;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const float *lite = lite_all + (size_t)scene * boxes_num * kLite;
    const float *rec = rec_all + (size_t)scene * boxes_num * kRec;
    const int *perm = perm_all ? perm_all + (size_t)scene * boxes_num : nullptr;
    const float4 *tbox = tile_box_all + (size_t)scene * cb;
    const float4 *hbox = half_box_all + (size_t)scene * cb * 2;
    unsigned long long *mask = mask_all + (size_t)scene * boxes_num * cb;
    const int row_size = min(boxes_num - rb * kTile, kTile);
    const float tq = thresh * (1.f - 1e-3f) - 1e-6f;
    // thresh >= 0.5 (with margin): a pair can only pass if each centre lies inside the other box, so
    // the centres are no farther apart than the smaller bounding radius (see cheap_passes)
    const bool tight = tq >= 0.5f;
    const float rscale = nms_rscale(tq);

    for (int w = t; w < kTile * kLite; w += 256) rowl[w] = (w >> 3) < row_size ? lite[(size_t)(rb * kTile) * kLite + w] : 0.f;
    __syncthreads();
    // bounding rectangle of the 16 rows this warp tests in stage 1 (rows 16*(warp>>1) .. +16)
    const int row0 = 16 * (warp >> 1);
    float4 rsub;
    {
        const int i = row0 + (lane & 15);
        const bool v = i < row_size;
        const float cx = rowl[i * kLite], cy = rowl[i * kLite + 1], r = rowl[i * kLite + 2] * (tight ? 0.5f * rscale : 1.f);
        float x0 = v ? cx - r : INFINITY, y0 = v ? cy - r : INFINITY, x1 = v ? cx + r : -INFINITY, y1 = v ? cy + r : -INFINITY;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            x0 = fminf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
            y0 = fminf(y0, __shfl_xor_sync(0xffffffffu, y0, o));
            x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
            y1 = fmaxf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
        }
        rsub = make_float4(x0, y0, x1, y1);
        if (lane == 0 && (warp & 1) == 0) rsubs[warp >> 1] = rsub;
    }

    // ---- warp-private queue of pairs that passed stage 2 (entry = row-in-tile | sorted col << 6) ----
    unsigned int *myq = wq[warp];
    int nq = 0;  // warp-uniform
    auto drain = [&](int keep_below) {  // stage 3 + output, 32 entries at a time, while nq > keep_below
        while (nq > keep_below) {
            const int take = min(nq, 32);
            const int q = nq - take + lane;
            bool pass = false;
            unsigned int e = 0;
            if (lane < take) {
                e = myq[q];
                const int i = (int)(e & 63u), col = (int)(e >> 6);
                float bl[kLite];
                const float4 b0 = __ldg(reinterpret_cast<const float4 *>(lite + (size_t)col * kLite));
                const float4 b1 = __ldg(reinterpret_cast<const float4 *>(lite + (size_t)col * kLite) + 1);
                bl[0] = b0.x; bl[1] = b0.y; bl[2] = b0.z; bl[3] = b0.w;
                bl[4] = b1.x; bl[5] = b1.y; bl[6] = b1.z; bl[7] = b1.w;
                pass = !area_rejects(rowl + i * kLite, bl, tq);
            }
            nq -= take;
            const unsigned bal = __ballot_sync(0xffffffffu, pass);
            if (bal) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(lists.count, (unsigned long long)__popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0) + __popc(bal & ((1u << lane) - 1u));
                if (pass) {
                    const int row = rb * kTile + (int)(e & 63u), col = (int)(e >> 6);
                    if (base < lists.cap)
                        lists.pairs[base] = ((unsigned long long)scene << 48) | ((unsigned long long)row << 24) | (unsigned)col;
                    else
                        clip_pair_overflow(rec, perm, mask, cb, row, col, thresh);
                }
            }
        }
    };

    const float4 rbox = __ldg(tbox + rb);
    for (int ct0 = rb; ct0 < cb; ct0 += kTileListCap) {
        // ---- stage 0: which column tiles can hold a passing pair at all ----
        __syncthreads();  // the previous pass's readers of tiles[] are done
        if (t == 0) ntiles = 0;
        __syncthreads();
        const int ct1 = min(cb, ct0 + kTileListCap);
        {
            const int ct = ct0 + t;
            bool hit = false;
            if (ct < ct1) {
                const float4 cbx = __ldg(tbox + ct);
                // negated so that a NaN rectangle is never culled
                if (!(cbx.x > rbox.z || rbox.x > cbx.z || cbx.y > rbox.w || rbox.y > cbx.w)) {
                    // refine: some (16 rows) x (32 columns) block must touch
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float4 c = __ldg(hbox + (size_t)ct * 2 + h);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const float4 r = rsubs[g];
                            hit = hit || !(c.x > r.z || r.x > c.z || c.y > r.w || r.y > c.w);
                        }
                    }
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (bal) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&ntiles, __popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (hit) tiles[base + __popc(bal & ((1u << lane) - 1u))] = ct;
            }
        }
        __syncthreads();
        const int nt = ntiles;
        unsigned short *mine = list1[warp];
        for (int ti = 0; ti < nt; ++ti) {  // no CTA barrier in here
            const int ct = tiles[ti];
            // this warp's block = 16 rows x 32 columns: skip it when the two rectangles are disjoint
            const float4 csub = __ldg(hbox + (size_t)ct * 2 + (warp & 1));
            if (csub.x > rsub.z || rsub.x > csub.z || csub.y > rsub.w || rsub.y > csub.w) continue;
            const int j = 32 * (warp & 1) + lane, col = ct * kTile + j;
            const bool jvalid = col < boxes_num;
            float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f);  // cx cy r hx of this lane's column box
            if (jvalid) b0 = __ldg(reinterpret_cast<const float4 *>(lite + (size_t)col * kLite));
            const bool diag = ct == rb;
            // ---- stage 1: centre distance.  tight: |offset| <= rscale * min(r_a, r_b), else touching circles
            int cnt1 = 0;
#pragma unroll 4
            for (int k = 0; k < 16; ++k) {
                const int i = row0 + k;
                const float4 a0 = *reinterpret_cast<const float4 *>(rowl + i * kLite);  // cx cy r hx (broadcast)
                const float dx = b0.x - a0.x, dy = b0.y - a0.y;
                const float rr = tight ? rscale * fminf(a0.z, b0.z) : a0.z + b0.z;
                const bool live = jvalid && i < row_size && (!diag || j > i) && !(dx * dx + dy * dy > rr * rr);
                const unsigned bal = __ballot_sync(0xffffffffu, live);
                if (live) mine[cnt1 + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)((i << 6) | j);
                cnt1 += __popc(bal);
            }
            __syncwarp();
            // ---- stage 2: rotation-free bounds, then the separating-axis bound; survivors into the queue ----
            for (int q0 = 0; q0 < cnt1; q0 += 32) {
                const int q = q0 + lane;
                bool pass = false;
                unsigned int entry = 0;
                if (q < cnt1) {
                    const int e = mine[q], i = e >> 6, jj = e & 63, c2 = ct * kTile + jj;
                    float bl[kLite];
                    const float4 v0 = __ldg(reinterpret_cast<const float4 *>(lite + (size_t)c2 * kLite));
                    const float4 v1 = __ldg(reinterpret_cast<const float4 *>(lite + (size_t)c2 * kLite) + 1);
                    bl[0] = v0.x; bl[1] = v0.y; bl[2] = v0.z; bl[3] = v0.w;
                    bl[4] = v1.x; bl[5] = v1.y; bl[6] = v1.z; bl[7] = v1.w;
                    const float *al = rowl + i * kLite;
                    const bool deg = !(al[3] > 0.f && al[4] > 0.f && bl[3] > 0.f && bl[4] > 0.f);  // degenerate box: always clip
                    pass = deg || !(tq > 0.f) ||
                           cheap_passes(bl[0] - al[0], bl[1] - al[1], al[3], al[4], al[5], al[6], lite_area(al), bl[3], bl[4],
                                        bl[5], bl[6], lite_area(bl), al[7] + bl[7], tq, tight);
                    pass = pass && !sat_rejects(al, bl, tq);
                    entry = (unsigned)i | ((unsigned)c2 << 6);
                }
                const unsigned bal = __ballot_sync(0xffffffffu, pass);
                if (pass) myq[nq + __popc(bal & ((1u << lane) - 1u))] = entry;  // nq <= 31 here, + 32 <= kWarpQueue
                nq += __popc(bal);
                __syncwarp();
                drain(31);
            }
            __syncwarp();
        }
    }
    __syncwarp();
    drain(0);
}

// One thread per listed pair (grid-stride over the list).
__global__ void __launch_bounds__(256)
    nms_clip_kernel(int boxes_num, float thresh, const float *__restrict__ rec_all, const int *__restrict__ perm_all,
                    unsigned long long *__restrict__ mask_all, NmsLists lists) {
    const int cb = (boxes_num + kTile - 1) / kTile;
    unsigned long long n = *lists.count;
    if (n > lists.cap) n = lists.cap;
    for (unsigned long long q = (unsigned long long)blockIdx.x * 256 + threadIdx.x; q < n;
         q += (unsigned long long)gridDim.x * 256) {
        unsigned long long e = lists.pairs[q];
        int scene = (int)(e >> 48), row = (int)((e >> 24) & 0xFFFFFFull), col = (int)(e & 0xFFFFFFull);
        clip_pair(rec_all + (size_t)scene * boxes_num * kRec, perm_all ? perm_all + (size_t)scene * boxes_num : nullptr,
                  mask_all + (size_t)scene * boxes_num * cb, cb, row, col, thresh);
    }
}

// Greedy sweep, one 1024-thread CTA per scene (bev_iou.cpp:87-112).  Per block of 64 boxes:
// thread 0 resolves the diagonal word serially in registers; the rows of the block (all
// column words to the right) were loaded by all threads BEFORE the resolve (64 rows x 16
// column lanes, 8 words in flight per thread), so once the kept bits are known the kept
// rows are OR-reduced (redux.or per warp, then one shared atomicOr) into the removal
// words without waiting for memory again.
constexpr int kSweep2Threads = 1024;
constexpr int kSweepInflight = 8;
__global__ void __launch_bounds__(kSweep2Threads)
    nms_sweep_kernel2(int boxes_num, const unsigned long long *__restrict__ mask, int *__restrict__ keep,
                      int *__restrict__ num_kept, int max_keep) {
    // max_keep < boxes_num: keep is (scenes, max_keep) and the sweep stops at the block in which the max_keep-th box is
    // kept - the caller (sb_nms_fn's slice) never looks further, and with half of the boxes kept that is 1/8 of the blocks
    extern __shared__ unsigned long long remv[];  // [cb]
    __shared__ unsigned long long sdiag[kTile];
    __shared__ unsigned long long skeep;
    const int cb = (boxes_num + kTile - 1) / kTile;
    const int scene = blockIdx.x, t = threadIdx.x, lane = t & 31;
    const int r = t & 63, g = t >> 6;  // row of the block, column lane (16 of them)
    const unsigned long long *sm = mask + (size_t)scene * boxes_num * cb;
    int *kp = keep + (size_t)scene * max_keep;
    for (int c = t; c < cb; c += kSweep2Threads) remv[c] = 0ull;
    int kept = 0;
    unsigned long long dnext = (t < kTile && t < boxes_num) ? sm[(size_t)t * cb] : 0ull;
    __syncthreads();
    for (int nb = 0; nb < cb; ++nb) {
        const int size = min(boxes_num - nb * kTile, kTile);
        const bool rvalid = r < size;
        const unsigned long long *rowp = sm + (size_t)(nb * kTile + r) * cb;
        if (t < kTile) sdiag[t] = dnext;
        // first wave of this block's row words, issued before the resolve
        unsigned long long w[kSweepInflight];
#pragma unroll
        for (int k = 0; k < kSweepInflight; ++k) {
            int c = nb + 1 + g + 16 * k;
            w[k] = (rvalid && c < cb) ? rowp[c] : 0ull;
        }
        if (t < kTile && nb + 1 < cb) {
            int row = (nb + 1) * kTile + t;
            dnext = row < boxes_num ? sm[(size_t)row * cb + nb + 1] : 0ull;
        }
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
        if (t < kTile && ((kb >> t) & 1ull)) {
            const int pos = kept + __popcll(kb & ((1ull << t) - 1ull));
            if (pos < max_keep) kp[pos] = nb * kTile + t;
        }
        kept += __popcll(kb);
        if (kept >= max_keep) {  // block uniform
            kept = max_keep;
            break;
        }
        const bool rkept = (kb >> r) & 1ull;
        for (int c0 = nb + 1; c0 < cb; c0 += 16 * kSweepInflight) {
            if (c0 != nb + 1) {
#pragma unroll
                for (int k = 0; k < kSweepInflight; ++k) {
                    int c = c0 + g + 16 * k;
                    w[k] = (rvalid && c < cb) ? rowp[c] : 0ull;
                }
            }
#pragma unroll
            for (int k = 0; k < kSweepInflight; ++k) {
                int c = c0 + g + 16 * k;  // uniform across the warp
                if (c < cb) {
                    unsigned long long v = rkept ? w[k] : 0ull;
                    unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
                    unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
                    if (lane == 0 && (lo | hi)) atomicOr(&remv[c], ((unsigned long long)hi << 32) | lo);
                }
            }
        }
        __syncthreads();
    }
    __syncthreads();          // (the early exit leaves the loop before its closing barrier)
    const int first = kp[0];  // written by this CTA before the barrier; pad with keep[0] (:110-112)
    for (int p = kept + t; p < max_keep; p += kSweep2Threads) kp[p] = first;
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

namespace hf {
namespace {
struct NmsWorkspace {
    size_t mask_off, rec_off, lite_off, perm_off, tbox_off, pairs_off, count_off, total;
    unsigned long long cap;
};
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
NmsWorkspace nms_workspace(int scenes, int boxes_num) {
    NmsWorkspace w{};
    size_t S = (size_t)scenes, N = (size_t)boxes_num, cb = (N + kTile - 1) / kTile;
    w.cap = (unsigned long long)(S * N * 64);  // listed pairs; anything beyond is clipped inside the filter kernel
    w.mask_off = 0;
    w.rec_off = align256(w.mask_off + S * N * cb * sizeof(unsigned long long));
    w.lite_off = align256(w.rec_off + S * N * kRec * sizeof(float));
    w.perm_off = align256(w.lite_off + S * N * kLite * sizeof(float));
    w.tbox_off = align256(w.perm_off + S * N * sizeof(int));
    w.pairs_off = align256(w.tbox_off + S * cb * 3 * sizeof(float4));  // tile + 2 half-tile rectangles
    w.count_off = align256(w.pairs_off + (size_t)w.cap * sizeof(unsigned long long));
    w.total = w.count_off + 256;
    return w;
}
}  // namespace
}  // namespace hf

namespace hf {
namespace {
thread_local int g_nms_spatial_sort = 1;
}
}  // namespace hf
extern "C" void hf_nms_set_spatial_sort(int on) { hf::g_nms_spatial_sort = on; }

extern "C" size_t hf_oriented_nms_workspace_bytes(int scenes, int boxes_num) {
    if (scenes <= 0 || boxes_num <= 0) return 0;
    return hf::nms_workspace(scenes, boxes_num).total;
}

// Debug/measurement aid: number of pairs the last hf_oriented_nms call on this workspace sent
// to the exact clip (read back from the workspace; synchronises the stream).
extern "C" long long hf_oriented_nms_last_pair_count(int scenes, int boxes_num, const void *workspace, hf_stream_t stream) {
    using namespace hf;
    if (scenes <= 0 || boxes_num <= 0 || !workspace) return -1;
    const NmsWorkspace w = nms_workspace(scenes, boxes_num);
    unsigned long long v = 0;
    if (cudaMemcpyAsync(&v, static_cast<const char *>(workspace) + w.count_off, sizeof(v), cudaMemcpyDeviceToHost,
                        (cudaStream_t)stream) != cudaSuccess)
        return -1;
    if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return -1;
    return (long long)v;
}

namespace hf {
namespace {
int oriented_nms_any(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int max_keep, int *keep,
                     int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream);
}
}  // namespace hf

extern "C" int hf_oriented_nms(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int *keep,
                               int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    return hf::oriented_nms_any(boxes, scenes, boxes_num, nms_overlap_thresh, boxes_num, keep, num_kept, workspace,
                                workspace_bytes, stream);
}

extern "C" int hf_oriented_nms_topk(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int max_keep,
                                    int *keep, int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(max_keep >= 1, "OrientedNMS: max_keep must be >= 1");
    return oriented_nms_any(boxes, scenes, boxes_num, nms_overlap_thresh, max_keep < boxes_num ? max_keep : boxes_num, keep,
                            num_kept, workspace, workspace_bytes, stream);
}

namespace hf {
namespace {
int oriented_nms_any(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int max_keep, int *keep,
                     int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream) {
    HF_REQUIRE(nms_overlap_thresh >= 0, "Need nms_threshold >= 0");
    HF_REQUIRE(boxes_num > 0 && scenes >= 0, "OrientendNMS expects (N, 5) boxes shape");
    HF_REQUIRE(boxes && keep, "OrientedNMS: null pointer");
    HF_REQUIRE(scenes <= 65535 && boxes_num < (1 << 22), "OrientedNMS: at most 65535 scenes of < 4194304 boxes");
    if (scenes == 0) return HF_OK;
    const NmsWorkspace w = nms_workspace(scenes, boxes_num);
    if (!workspace || workspace_bytes < w.total) {
        set_error("OrientedNMS: workspace of %zu bytes required, got %zu", w.total, workspace ? workspace_bytes : (size_t)0);
        return HF_ERR_WORKSPACE;
    }
    HF_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "OrientedNMS: workspace must be 16-byte aligned");
    const int cb = (boxes_num + kTile - 1) / kTile;
    size_t smem = (size_t)cb * sizeof(unsigned long long);
    if (smem > 200 * 1024) {
        set_error("OrientedNMS: %d boxes exceed the sweep kernel's shared memory", boxes_num);
        return HF_ERR_UNSUPPORTED;
    }
    cudaStream_t s = (cudaStream_t)stream;
    char *base = static_cast<char *>(workspace);
    unsigned long long *mask = reinterpret_cast<unsigned long long *>(base + w.mask_off);
    float *rec = reinterpret_cast<float *>(base + w.rec_off);
    float *lite = reinterpret_cast<float *>(base + w.lite_off);
    int *perm = reinterpret_cast<int *>(base + w.perm_off);
    float4 *tbox = reinterpret_cast<float4 *>(base + w.tbox_off);
    NmsLists lists;
    lists.pairs = reinterpret_cast<unsigned long long *>(base + w.pairs_off);
    lists.count = reinterpret_cast<unsigned long long *>(base + w.count_off);
    lists.cap = w.cap;
    // the clip kernel ORs bits into the mask; the pair counter starts at 0
    HF_CUDA_OK(cudaMemsetAsync(mask, 0, (size_t)scenes * boxes_num * cb * sizeof(unsigned long long), s));
    HF_CUDA_OK(cudaMemsetAsync(lists.count, 0, sizeof(unsigned long long), s));
    // spatial order: counting sort of the centres (the pair list region doubles as its scratch)
    if (g_nms_spatial_sort) {
        nms_sort_kernel<<<scenes, kSortThreads, 0, s>>>(boxes_num, boxes, perm, lists.pairs);
        HF_LAUNCH_OK("nms_sort_kernel");
    } else {
        perm = nullptr;
    }
    float4 *hbox = tbox + (size_t)scenes * cb;
    const float tq = nms_overlap_thresh * (1.f - 1e-3f) - 1e-6f;  // same expression as in the filter kernel
    nms_prepare_kernel<<<dim3(cb, scenes), kTile, 0, s>>>(boxes_num, boxes, perm, rec, lite, tbox, hbox,
                                                          tq >= 0.5f ? 0.5f * nms_rscale(tq) : 1.f);
    HF_LAUNCH_OK("nms_prepare_kernel");
    nms_filter_kernel<<<dim3(cb, scenes), 256, 0, s>>>(boxes_num, nms_overlap_thresh, rec, lite, tbox, hbox, perm, mask, lists);
    HF_LAUNCH_OK("nms_filter_kernel");
    unsigned long long want = (w.cap + 255) / 256;
    unsigned cgrid = (unsigned)(want < (unsigned long long)num_sms() * 8 ? want : (unsigned long long)num_sms() * 8);
    nms_clip_kernel<<<cgrid, 256, 0, s>>>(boxes_num, nms_overlap_thresh, rec, perm, mask, lists);
    HF_LAUNCH_OK("nms_clip_kernel");
    if (smem > 40 * 1024)
        HF_CUDA_OK(cudaFuncSetAttribute(nms_sweep_kernel2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nms_sweep_kernel2<<<scenes, kSweep2Threads, smem, s>>>(boxes_num, mask, keep, num_kept, max_keep);
    HF_LAUNCH_OK("nms_sweep_kernel2");
    return HF_OK;
}
}  // namespace
}  // namespace hf
