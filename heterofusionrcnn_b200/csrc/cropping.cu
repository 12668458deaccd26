// cropping.cu — RoI point crop with feature pooling (pc_crop_and_sample) for sm_100a.
//
// Replaces cropping/tf_cropping_g.cu:7-132 + the output memsets of
// cropping/tf_cropping.cpp:171-176 (see include/hfops.h).
//
// One CTA per box, three phases, everything the box needs staying in shared memory:
//   1. select: each warp scans a contiguous slice of the scene's points in ascending
//      order and compacts the inside-the-box indices with ballot/popc (ordered, no
//      atomics — the reference's atomicInc slot assignment, :85, is a race); a prefix over
//      the 8 warp counts concatenates the slices, first `resize` kept;
//   2. pad: slot s >= count repeats slot (s - count) mod count (:108-126) — done on the
//      index list, so padded rows are copied from the source exactly like selected rows;
//   3. copy: HBM-bound.  One warp per output row, 128-bit loads of the gathered feature
//      row and streaming 128-bit stores, four rows in flight per warp.  Every output
//      element is written exactly once (empty boxes are zero-filled here), so no memset
//      pass over the 4.8 GB output is needed.
#include <limits.h>

#include "common.cuh"

namespace hf {
namespace {

constexpr int kCropThreads = 256;
constexpr int kCropWarps = kCropThreads / 32;
constexpr int kCropRows = 4;    // consecutive output rows a warp copies per step
constexpr int kCropFlight = 9;  // 128-bit loads in flight per lane (4 rows x 72 chunks = 9 x 32)

// dot() of tf_cropping_g.cu:3-5 in the order nvcc contracts it on sm_100a.
__device__ __forceinline__ float dot3(float x1, float y1, float z1, float x2, float y2, float z2) {
    return __fmaf_rn(z1, z2, __fmaf_rn(y1, y2, __fmul_rn(x1, x2)));
}

// INDEXED: `index` is the cloud index of pts (hf_cloud_index_build: per scan the points sorted along a
// Hilbert curve as {x, y, z, original index} in buckets of 32 with their bounding boxes, spatial.cu).  Phase 1
// then tests the 512 bucket boxes of the scan against the RoI's (padded) bounding box and runs the exact
// predicate only on the points of the few buckets that can touch it, instead of on all 16384 points — the
// full scan is two thirds of the kernel's instructions.  Hits come out in curve order, so they are sorted
// by original index (bitonic, shared memory) to give the same canonical slot order as the scan.
template <bool VEC4, bool INDEXED>
__global__ void __launch_bounds__(kCropThreads)
    crop_kernel(const float *__restrict__ pts_data, const float *__restrict__ fts_data,
                const float *__restrict__ int_data, const uint8_t *__restrict__ mask_data,
                const float *__restrict__ boxes, const int *__restrict__ box_ind, int npts, int resize,
                int channel, int ichannel, float *__restrict__ crop_pts, float *__restrict__ crop_fts,
                float *__restrict__ crop_int, uint8_t *__restrict__ crop_mask, int *__restrict__ crop_ind,
                uint8_t *__restrict__ non_empty, const float4 *__restrict__ ix_sorted, const float4 *__restrict__ ix_boxes,
                int npad, const float4 *__restrict__ ix_super, int nsuper) {
    extern __shared__ int smem_i[];
    int *sel = smem_i;                // [resize] final source index per slot
    int *wlist = smem_i + resize;     // [kCropWarps][resize] per-warp ordered hits
    __shared__ int wcount[kCropWarps];
    __shared__ int hcount;
    __shared__ int slist[64], scount;  // super-boxes (groups of 32 buckets) the RoI's bounding box touches

    const int b = blockIdx.x;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const float *bx = boxes + (size_t)b * 24;
    const float p1x = bx[0], p1y = bx[8], p1z = bx[16];
    const float p2x = bx[1], p2y = bx[9], p2z = bx[17];
    const float p4x = bx[3], p4y = bx[11], p4z = bx[19];
    const float p5x = bx[4], p5y = bx[12], p5z = bx[20];
    // is_point_inside (:7-41); everything that does not depend on the point is hoisted
    const float ux = p2x - p1x, uy = p2y - p1y, uz = p2z - p1z;
    const float vx = p4x - p1x, vy = p4y - p1y, vz = p4z - p1z;
    const float wx = p5x - p1x, wy = p5y - p1y, wz = p5z - p1z;
    const float u_p1 = dot3(ux, uy, uz, p1x, p1y, p1z), u_p2 = dot3(ux, uy, uz, p2x, p2y, p2z);
    const float v_p1 = dot3(vx, vy, vz, p1x, p1y, p1z), v_p4 = dot3(vx, vy, vz, p4x, p4y, p4z);
    const float w_p1 = dot3(wx, wy, wz, p1x, p1y, p1z), w_p5 = dot3(wx, wy, wz, p5x, p5y, p5z);

    const int bch = box_ind[b];
    const float *pts = pts_data + (size_t)bch * npts * 3;

    // ---- 1. select ----
    int count = -1;  // -1: not selected yet (no index, or more hits than the list holds)
    if (INDEXED) {
        int *hits = wlist;  // [kCropWarps * resize] original indices, any order
        const int hcap = kCropWarps * resize;
        if (t == 0) hcount = 0;
        // bounding box of the 8 corners, padded far beyond the rounding of the dot products below
        float lox = bx[0], hix = bx[0], loy = bx[8], hiy = bx[8], loz = bx[16], hiz = bx[16];
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            lox = fminf(lox, bx[k]); hix = fmaxf(hix, bx[k]);
            loy = fminf(loy, bx[8 + k]); hiy = fmaxf(hiy, bx[8 + k]);
            loz = fminf(loz, bx[16 + k]); hiz = fmaxf(hiz, bx[16 + k]);
        }
        const float padx = 1e-2f + 1e-4f * fmaxf(fabsf(lox), fabsf(hix)), pady = 1e-2f + 1e-4f * fmaxf(fabsf(loy), fabsf(hiy)),
                    padz = 1e-2f + 1e-4f * fmaxf(fabsf(loz), fabsf(hiz));
        lox -= padx; hix += padx; loy -= pady; hiy += pady; loz -= padz; hiz += padz;
        const int nb = npad >> 5;
        const float4 *srt = ix_sorted + (size_t)bch * npad;
        const float4 *bxs = ix_boxes + (size_t)bch * nb * 2;
        // large clouds carry a second box level (spatial.cu:cloud_super_kernel): only the groups of 32 buckets whose
        // super-box the RoI touches are looked at, instead of every bucket box of the scan (2048 at 65536 points)
        int ngroups = (nb + 31) >> 5;
        const bool two_level = ix_super != nullptr && nsuper <= 64;
        if (two_level) {
            if (t == 0) scount = 0;
            __syncthreads();
            if (t < nsuper) {
                const float4 slo = __ldg(&ix_super[((size_t)bch * nsuper + t) * 2]), shi = __ldg(&ix_super[((size_t)bch * nsuper + t) * 2 + 1]);
                if (slo.x <= hix && shi.x >= lox && slo.y <= hiy && shi.y >= loy && slo.z <= hiz && shi.z >= loz)
                    slist[atomicAdd(&scount, 1)] = t;
            }
        }
        __syncthreads();
        if (two_level) ngroups = scount;
        for (int gi = warp; gi < ngroups; gi += kCropWarps) {
            const int g0 = (two_level ? slist[gi] : gi) * 32;
            const int cb = g0 + lane;
            bool touch = false;
            if (cb < nb) {
                const float4 blo = __ldg(&bxs[cb * 2]), bhi = __ldg(&bxs[cb * 2 + 1]);
                touch = blo.x <= hix && bhi.x >= lox && blo.y <= hiy && bhi.y >= loy && blo.z <= hiz && bhi.z >= loz;
            }
            unsigned todo = __ballot_sync(0xffffffffu, touch);
            while (todo) {
                const int l = __ffs(todo) - 1;
                todo &= todo - 1;
                const float4 c = __ldg(&srt[(size_t)(g0 + l) * 32 + lane]);
                const float ud = dot3(ux, uy, uz, c.x, c.y, c.z);
                const float vd = dot3(vx, vy, vz, c.x, c.y, c.z);
                const float wd = dot3(wx, wy, wz, c.x, c.y, c.z);
                const bool in = u_p1 < ud && ud < u_p2 && v_p1 < vd && vd < v_p4 && w_p1 < wd && wd < w_p5;  // NaN padding fails
                const unsigned bal = __ballot_sync(0xffffffffu, in);
                if (bal) {
                    int base = 0;
                    if (lane == 0) base = atomicAdd(&hcount, __popc(bal));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    const int pos = base + __popc(bal & ((1u << lane) - 1u));
                    if (in && pos < hcap) hits[pos] = __float_as_int(c.w);
                }
            }
        }
        __syncthreads();
        const int total = hcount;
        // the list is sorted over P = next power of two >= max(32, total) entries: it must hold P, not just
        // `total` (resize = 100: hcap = 800, 600 hits would pad and sort 1024 entries; resize < 4: hcap < 32)
        int P = 32;
        while (P < total) P <<= 1;
        if (P <= hcap) {
            for (int i = total + t; i < P; i += kCropThreads) hits[i] = INT_MAX;
            __syncthreads();
            for (int k2 = 2; k2 <= P; k2 <<= 1) {
                for (int j = k2 >> 1; j > 0; j >>= 1) {
                    for (int i = t; i < P; i += kCropThreads) {
                        const int q = i ^ j;
                        if (q > i) {
                            const int x = hits[i], y = hits[q];
                            if ((x > y) == ((i & k2) == 0)) {
                                hits[i] = y;
                                hits[q] = x;
                            }
                        }
                    }
                    __syncthreads();
                }
            }
            count = min(total, resize);
            for (int i = t; i < count; i += kCropThreads) sel[i] = hits[i];
            __syncthreads();
        }
    }
    if (count < 0) {  // ordered scan of the whole scan (block uniform)
        const int per_warp = (npts + kCropWarps - 1) / kCropWarps;
        const int lo = warp * per_warp;
        const int hi = min(npts, lo + per_warp);
        int *mine = wlist + warp * resize;
        int cnt = 0;
        for (int p0 = lo; p0 < hi && cnt < resize; p0 += 32) {
            int p = p0 + lane;
            bool in = false;
            if (p < hi) {
                float px = pts[(size_t)p * 3], py = pts[(size_t)p * 3 + 1], pz = pts[(size_t)p * 3 + 2];
                float ud = dot3(ux, uy, uz, px, py, pz);
                float vd = dot3(vx, vy, vz, px, py, pz);
                float wd = dot3(wx, wy, wz, px, py, pz);
                in = u_p1 < ud && ud < u_p2 && v_p1 < vd && vd < v_p4 && w_p1 < wd && wd < w_p5;
            }
            unsigned bal = __ballot_sync(0xffffffffu, in);
            int pos = cnt + __popc(bal & ((1u << lane) - 1u));
            if (in && pos < resize) mine[pos] = p;
            cnt += __popc(bal);
        }
        if (cnt > resize) cnt = resize;
        __syncthreads();  // (indexed overflow: every warp is done with the hit list before it is reused)
        if (lane == 0) wcount[warp] = cnt;
        __syncthreads();
        int offset = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kCropWarps; ++w) {
            int c = wcount[w];
            if (w < warp) offset += c;
            total += c;
        }
        count = min(total, resize);
        for (int i = lane; i < cnt; i += 32) {
            int pos = offset + i;
            if (pos < resize) sel[pos] = mine[i];
        }
        __syncthreads();
    }
    // ---- 2. pad ----
    if (count > 0) {
        for (int s = count + t; s < resize; s += kCropThreads) sel[s] = sel[(s - count) % count];
    }
    __syncthreads();
    if (t == 0) non_empty[b] = count > 0 ? 1 : 0;

    // ---- 3. copy ----
    float *o_pts = crop_pts + (size_t)b * resize * 3;
    float *o_int = crop_int + (size_t)b * resize * ichannel;
    uint8_t *o_mask = crop_mask + (size_t)b * resize;
    int *o_ind = crop_ind + (size_t)b * resize;
    const float *intens = int_data + (size_t)bch * npts * ichannel;
    const uint8_t *mask = mask_data + (size_t)bch * npts;
    for (int s = t; s < resize; s += kCropThreads) {
        if (count > 0) {
            int p = sel[s];
            o_pts[s * 3] = pts[(size_t)p * 3];
            o_pts[s * 3 + 1] = pts[(size_t)p * 3 + 1];
            o_pts[s * 3 + 2] = pts[(size_t)p * 3 + 2];
            for (int c = 0; c < ichannel; ++c) o_int[s * ichannel + c] = intens[(size_t)p * ichannel + c];
            o_mask[s] = mask[p];
            o_ind[s] = p;
        } else {
            o_pts[s * 3] = 0.f; o_pts[s * 3 + 1] = 0.f; o_pts[s * 3 + 2] = 0.f;
            for (int c = 0; c < ichannel; ++c) o_int[s * ichannel + c] = 0.f;
            o_mask[s] = 0;
            o_ind[s] = 0;
        }
    }
    if (VEC4) {
        const int c4 = channel >> 2;
        const float4 *fts = reinterpret_cast<const float4 *>(fts_data) + (size_t)bch * npts * c4;
        float4 *o = reinterpret_cast<float4 *>(crop_fts) + (size_t)b * resize * c4;
        if (count > 0) {
            // A warp copies kCropRows consecutive output rows per step as ONE flat run of kCropRows * c4 128-bit
            // chunks (the rows are contiguous in `o`): every lane's loads of the step are issued before its first store,
            // and no iteration runs with idle lanes (c4 = 72: 9 full iterations per 4 rows instead of 3 per row with the
            // last one on 8 lanes).  The (row, chunk) of a lane advances incrementally: no division by c4.
            for (int s0 = warp * kCropRows; s0 < resize; s0 += kCropWarps * kCropRows) {
                const int rows = min(kCropRows, resize - s0);
                const int total = rows * c4;
                float4 *dst = o + (size_t)s0 * c4;
                for (int e0 = 0; e0 < total; e0 += 32 * kCropFlight) {
                    float4 val[kCropFlight];
                    int u = 0, v = e0 + lane;  // chunk e = e0 + lane + 32 * k  ->  row u, offset v
                    while (v >= c4) { v -= c4; ++u; }
                    int uu = u, vv = v;
#pragma unroll
                    for (int k = 0; k < kCropFlight; ++k) {
                        if (e0 + lane + 32 * k < total) val[k] = __ldg(&fts[(size_t)sel[s0 + uu] * c4 + vv]);
                        vv += 32;
                        while (vv >= c4) { vv -= c4; ++uu; }
                    }
#pragma unroll
                    for (int k = 0; k < kCropFlight; ++k)
                        if (e0 + lane + 32 * k < total) stg_cs_f4(&dst[e0 + lane + 32 * k], val[k]);
                }
            }
        } else {
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
            const size_t tot = (size_t)resize * c4;
            for (size_t e = t; e < tot; e += kCropThreads) stg_cs_f4(&o[e], zero);
        }
    } else {
        const float *fts = fts_data + (size_t)bch * npts * channel;
        float *o = crop_fts + (size_t)b * resize * channel;
        for (int s = warp; s < resize; s += kCropWarps) {
            if (count > 0) {
                const float *src = fts + (size_t)sel[s] * channel;
                for (int c = lane; c < channel; c += 32) o[(size_t)s * channel + c] = __ldg(&src[c]);
            } else {
                for (int c = lane; c < channel; c += 32) o[(size_t)s * channel + c] = 0.f;
            }
        }
    }
}

}  // namespace
}  // namespace hf

extern "C" size_t hf_cloud_index_bytes(int b, int n);

namespace hf {
namespace {
template <bool VEC4, bool INDEXED>
int launch_crop(size_t smem, cudaStream_t s, const float *pts, const float *fts, const float *intensities, const uint8_t *mask,
                const float *boxes, const int *box_ind, int num_boxes, int npts, int resize, int channel,
                int intensity_channel, float *crop_pts, float *crop_fts, float *crop_intensities, uint8_t *crop_mask,
                int *crop_ind, uint8_t *non_empty_box, const float4 *ix_sorted, const float4 *ix_boxes, int npad,
                const float4 *ix_super, int nsuper) {
    auto kern = crop_kernel<VEC4, INDEXED>;
    if (smem > 40 * 1024) HF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<num_boxes, kCropThreads, smem, s>>>(pts, fts, intensities, mask, boxes, box_ind, npts, resize, channel,
                                               intensity_channel, crop_pts, crop_fts, crop_intensities, crop_mask, crop_ind,
                                               non_empty_box, ix_sorted, ix_boxes, npad, ix_super, nsuper);
    HF_LAUNCH_OK("crop_kernel");
    return HF_OK;
}
}  // namespace
}  // namespace hf

extern "C" int hf_pc_crop_and_sample_indexed(const float *pts, const float *fts, const float *intensities,
                                             const uint8_t *mask, const float *boxes, const int *box_ind, int num_boxes,
                                             int batch, int npts, int resize, int channel, int intensity_channel,
                                             float *crop_pts, float *crop_fts, float *crop_intensities,
                                             uint8_t *crop_mask, int *crop_ind, uint8_t *non_empty_box,
                                             const void *pts_index, hf_stream_t stream) {
    using namespace hf;
    HF_REQUIRE(resize > 0, "PcCropAndSample expects positive resize");
    HF_REQUIRE(batch >= 1 && npts > 0, "PcCropAndSample expects (B, P, 3) pts shape");
    HF_REQUIRE(channel > 0, "PcCropAndSample expects (B, P, C) fts shape");
    HF_REQUIRE(intensity_channel >= 1, "PcCropAndSample expects (B, P, 1) intensities shape");
    HF_REQUIRE(num_boxes >= 0, "PcCropAndSample expects (N, 3, 8) boxes shape");
    HF_REQUIRE(pts && fts && intensities && mask && boxes && box_ind && crop_pts && crop_fts &&
                   crop_intensities && crop_mask && crop_ind && non_empty_box,
               "PcCropAndSample: null pointer");
    if (num_boxes == 0) return HF_OK;
    size_t smem = (size_t)(1 + kCropWarps) * resize * sizeof(int);
    if (smem > 200 * 1024) {
        set_error("PcCropAndSample: resize=%d needs %zu bytes of shared memory (max 204800)", resize, smem);
        return HF_ERR_UNSUPPORTED;
    }
    const bool vec = (channel % 4 == 0) && ((reinterpret_cast<uintptr_t>(fts) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(crop_fts) & 15) == 0);
    // the layout of a cloud index belongs to spatial.cu: ask it for the view (false = no usable index, plain select)
    CloudIndexView ix = {nullptr, nullptr, nullptr, (npts + 31) & ~31, 0};
    const bool indexed = cloud_index_view(batch, npts, pts_index, &ix);
    const int npad = ix.npad, nsuper = ix.nsuper;
    const float4 *ix_sorted = ix.sorted, *ix_boxes = ix.boxes, *ix_super = ix.super;
    cudaStream_t s = (cudaStream_t)stream;
#define HF_CROP_ARGS                                                                                                     \
    smem, s, pts, fts, intensities, mask, boxes, box_ind, num_boxes, npts, resize, channel, intensity_channel, crop_pts, \
        crop_fts, crop_intensities, crop_mask, crop_ind, non_empty_box, ix_sorted, ix_boxes, npad, ix_super, nsuper
    if (vec) return indexed ? launch_crop<true, true>(HF_CROP_ARGS) : launch_crop<true, false>(HF_CROP_ARGS);
    return indexed ? launch_crop<false, true>(HF_CROP_ARGS) : launch_crop<false, false>(HF_CROP_ARGS);
#undef HF_CROP_ARGS
}

extern "C" int hf_pc_crop_and_sample(const float *pts, const float *fts, const float *intensities,
                                     const uint8_t *mask, const float *boxes, const int *box_ind, int num_boxes,
                                     int batch, int npts, int resize, int channel, int intensity_channel,
                                     float *crop_pts, float *crop_fts, float *crop_intensities,
                                     uint8_t *crop_mask, int *crop_ind, uint8_t *non_empty_box,
                                     hf_stream_t stream) {
    return hf_pc_crop_and_sample_indexed(pts, fts, intensities, mask, boxes, box_ind, num_boxes, batch, npts, resize, channel,
                                         intensity_channel, crop_pts, crop_fts, crop_intensities, crop_mask, crop_ind,
                                         non_empty_box, nullptr, stream);
}
