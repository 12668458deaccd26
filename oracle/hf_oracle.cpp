// hf_oracle.cpp — CPU ORACLE for the HeteroFusionRCNN point-cloud operator path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product package
// (heterofusionrcnn_b200/) may import, link or call this file; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// use it, and only as the checker / the timed CPU baseline.
//
// Each function is a serial restatement of one reference CUDA kernel (or of the
// reference's own CPU prototype where one exists), cited file:line against
// /root/reference.  Build: g++ -O2 -ffp-contract=off (oracle/Makefile); the
// places where nvcc 12.9 contracts the reference source into FMAs on sm_100a
// (checked with cuobjdump, see DESIGN.md "arithmetic contract") are written as
// explicit fmaf() so index outputs are bit-identical to the reference kernels
// as built for the B200.
//
// Parity pin: oracle/_ref (the reference's own sources compiled where they lie:
// interpolate.cpp, grouping/test/query_ball_point.cpp on the CPU, the five
// *_g.cu files with nvcc for the GPU box) + the demo known-answers in
// tests/golden/.  See tests/test_oracle_*.py.
//
// Optional threading (hf_oracle_set_threads): a plain parallel-for over the
// independent units the reference kernels already treat as independent (scenes,
// query rows, boxes, mask rows).  No algorithmic change; used only for the timed
// CPU baseline.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace {

int g_threads = 1;

// parallel_for over [0,count): g_threads std::threads pulling contiguous chunks off an
// atomic cursor (dynamic, so triangular workloads such as the NMS mask stay balanced).
void parallel_for(int64_t count, const std::function<void(int64_t, int64_t)> &fn) {
    int nt = g_threads;
    if (nt <= 1 || count <= 1) {
        fn(0, count);
        return;
    }
    if (nt > count) nt = (int)count;
    const int64_t chunk = std::max<int64_t>(1, count / ((int64_t)nt * 16));
    std::atomic<int64_t> cursor{0};
    std::vector<std::thread> pool;
    pool.reserve(nt);
    for (int t = 0; t < nt; ++t) {
        pool.emplace_back([&fn, &cursor, chunk, count] {
            for (;;) {
                int64_t lo = cursor.fetch_add(chunk, std::memory_order_relaxed);
                if (lo >= count) break;
                fn(lo, std::min(count, lo + chunk));
            }
        });
    }
    for (auto &th : pool) th.join();
}

// Squared distance exactly as nvcc 12.9 -O2 sm_100a emits it for
//   (x2-x1)*(x2-x1)+(y2-y1)*(y2-y1)+(z2-z1)*(z2-z1)
// in tf_sampling_g.cu:142, tf_grouping_g.cu:24 and tf_interpolate_g.cu:49:
//   t = dy*dy (FMUL); t = fma(dx,dx,t); t = fma(dz,dz,t).
inline float sqdist(float dx, float dy, float dz) {
    float t = dy * dy;
    t = std::fmaf(dx, dx, t);
    t = std::fmaf(dz, dz, t);
    return t;
}

// dot() of tf_cropping_g.cu:3-5 as contracted on sm_100a: fma(z1,z2, fma(y1,y2, x1*x2)).
inline float dot3(float x1, float y1, float z1, float x2, float y2, float z2) {
    float t = x1 * x2;
    t = std::fmaf(y1, y2, t);
    t = std::fmaf(z1, z2, t);
    return t;
}

}  // namespace

extern "C" {

void hf_oracle_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
int hf_oracle_get_threads() { return g_threads; }

// ---------------------------------------------------------------------------
// sampling/tf_sampling_g.cu:105-170  farthestpointsamplingKernel (<<<32,512>>>)
// One scene = one 512-thread block.  Thread t owns points k == t (mod 512);
// strict '>' against best=-1 keeps the lowest k per thread (:146-149); the
// pairwise tree (:153-163) keeps the LEFT slot on ties.
// ---------------------------------------------------------------------------
void hf_oracle_farthest_point_sample(int b, int n, int m, const float *inp, int *out) {
    if (m <= 0) return;
    const int BlockSize = 512;
    parallel_for(b, [&](int64_t lo, int64_t hi) {
        std::vector<float> temp(n);
        float dists[BlockSize];
        int dists_i[BlockSize];
        for (int64_t i = lo; i < hi; ++i) {
            const float *dataset = inp + i * (int64_t)n * 3;
            int *idxs = out + i * (int64_t)m;
            int old = 0;
            idxs[0] = old;
            for (int j = 0; j < n; ++j) temp[j] = 1e38f;
            for (int j = 1; j < m; ++j) {
                float x1 = dataset[old * 3 + 0];
                float y1 = dataset[old * 3 + 1];
                float z1 = dataset[old * 3 + 2];
                for (int t = 0; t < BlockSize; ++t) {
                    int besti = 0;
                    float best = -1;
                    for (int k = t; k < n; k += BlockSize) {
                        float td = temp[k];
                        float x2 = dataset[k * 3 + 0];
                        float y2 = dataset[k * 3 + 1];
                        float z2 = dataset[k * 3 + 2];
                        float d = sqdist(x2 - x1, y2 - y1, z2 - z1);
                        float d2 = std::fmin(d, td);
                        if (d2 != td) temp[k] = d2;
                        if (d2 > best) {
                            best = d2;
                            besti = k;
                        }
                    }
                    dists[t] = best;
                    dists_i[t] = besti;
                }
                for (int u = 0; (1 << u) < BlockSize; ++u) {
                    for (int t = 0; t < (BlockSize >> (u + 1)); ++t) {
                        int i1 = (t * 2) << u;
                        int i2 = (t * 2 + 1) << u;
                        if (dists[i1] < dists[i2]) {
                            dists[i1] = dists[i2];
                            dists_i[i1] = dists_i[i2];
                        }
                    }
                }
                old = dists_i[0];
                idxs[j] = old;
            }
        }
    });
}

// sampling/tf_sampling_g.cu:172-181  gatherpointKernel
void hf_oracle_gather_point(int b, int n, int m, const float *inp, const int *idx, float *out) {
    for (int64_t i = 0; i < b; ++i)
        for (int64_t j = 0; j < m; ++j) {
            int a = idx[i * m + j];
            for (int c = 0; c < 3; ++c) out[(i * m + j) * 3 + c] = inp[(i * n + a) * 3 + c];
        }
}

// ---------------------------------------------------------------------------
// grouping/tf_grouping_g.cu:3-36  query_ball_point_gpu
// (== grouping/test/query_ball_point.cpp:19-47 plus pts_cnt, :34 of the .cu)
// idx rows with no hit are left untouched (the reference leaves them
// uninitialised, tf_grouping.cpp:87-100); callers pre-zero.
// ---------------------------------------------------------------------------
void hf_oracle_query_ball_point(int b, int n, int m, float radius, int nsample, const float *xyz1,
                                const float *xyz2, int *idx, int *pts_cnt) {
    parallel_for((int64_t)b * m, [&](int64_t lo, int64_t hi) {
        for (int64_t q = lo; q < hi; ++q) {
            int64_t i = q / m;
            const float *p1 = xyz1 + i * (int64_t)n * 3;
            const float *p2 = xyz2 + q * 3;
            int *row = idx + q * nsample;
            int cnt = 0;
            float x2 = p2[0], y2 = p2[1], z2 = p2[2];
            for (int k = 0; k < n; ++k) {
                if (cnt == nsample) break;
                float x1 = p1[k * 3 + 0], y1 = p1[k * 3 + 1], z1 = p1[k * 3 + 2];
                float d = std::fmax(std::sqrt(sqdist(x2 - x1, y2 - y1, z2 - z1)), 1e-20f);
                if (d < radius) {
                    if (cnt == 0)
                        for (int l = 0; l < nsample; ++l) row[l] = k;
                    row[cnt] = k;
                    cnt += 1;
                }
            }
            pts_cnt[q] = cnt;
        }
    });
}

// grouping/tf_grouping_g.cu:40-57  group_point_gpu (== query_ball_point.cpp:52-66)
void hf_oracle_group_point(int b, int n, int c, int m, int nsample, const float *points,
                           const int *idx, float *out) {
    parallel_for((int64_t)b * m, [&](int64_t lo, int64_t hi) {
        for (int64_t q = lo; q < hi; ++q) {
            int64_t i = q / m;
            const float *pts = points + i * (int64_t)n * c;
            for (int k = 0; k < nsample; ++k) {
                int ii = idx[q * nsample + k];
                std::memcpy(out + (q * nsample + k) * c, pts + (int64_t)ii * c, sizeof(float) * c);
            }
        }
    });
}

// ---------------------------------------------------------------------------
// interpolate/tf_interpolate_g.cu:22-65  three_nn_kernel_fast
// d in float, bests in double starting at 1e40, strict '<' cascade.
// (interpolate/interpolate.cpp:21-64 is a stub that ignores the query; unused.)
// ---------------------------------------------------------------------------
void hf_oracle_three_nn(int b, int n, int m, const float *unknown, const float *known,
                        float *dist2, int *idx) {
    parallel_for((int64_t)b * n, [&](int64_t lo, int64_t hi) {
        for (int64_t q = lo; q < hi; ++q) {
            int64_t i = q / n;
            const float *kn = known + i * (int64_t)m * 3;
            float ux = unknown[q * 3 + 0], uy = unknown[q * 3 + 1], uz = unknown[q * 3 + 2];
            double best1 = 1e40, best2 = 1e40, best3 = 1e40;
            int besti1 = 0, besti2 = 0, besti3 = 0;
            for (int k = 0; k < m; ++k) {
                float x = kn[k * 3 + 0], y = kn[k * 3 + 1], z = kn[k * 3 + 2];
                float d = sqdist(ux - x, uy - y, uz - z);
                if (d < best1) {
                    best3 = best2; besti3 = besti2;
                    best2 = best1; besti2 = besti1;
                    best1 = d; besti1 = k;
                } else if (d < best2) {
                    best3 = best2; besti3 = besti2;
                    best2 = d; besti2 = k;
                } else if (d < best3) {
                    best3 = d; besti3 = k;
                }
            }
            dist2[q * 3 + 0] = (float)best1; dist2[q * 3 + 1] = (float)best2; dist2[q * 3 + 2] = (float)best3;
            idx[q * 3 + 0] = besti1; idx[q * 3 + 1] = besti2; idx[q * 3 + 2] = besti3;
        }
    });
}

// ---------------------------------------------------------------------------
// three_interpolate, channel-last (b,m,c)->(b,n,c): the Python contract of
// interpolate/tf_interpolate.py:26-37 and the layout of the CPU prototype
// interpolate/interpolate.cpp:84-104.  Arithmetic follows the GPU kernel
// tf_interpolate_g.cu:109 as contracted on sm_100a:
//   fma(w2,p2, fma(w0,p0, w1*p1)).
// ---------------------------------------------------------------------------
void hf_oracle_three_interpolate(int b, int m, int c, int n, const float *points, const int *idx,
                                 const float *weight, float *out) {
    parallel_for((int64_t)b * n, [&](int64_t lo, int64_t hi) {
        for (int64_t q = lo; q < hi; ++q) {
            int64_t i = q / n;
            const float *pts = points + i * (int64_t)m * c;
            float w0 = weight[q * 3], w1 = weight[q * 3 + 1], w2 = weight[q * 3 + 2];
            const float *p0 = pts + (int64_t)idx[q * 3] * c;
            const float *p1 = pts + (int64_t)idx[q * 3 + 1] * c;
            const float *p2 = pts + (int64_t)idx[q * 3 + 2] * c;
            float *o = out + q * c;
            for (int l = 0; l < c; ++l) {
                float t = w1 * p1[l];
                t = std::fmaf(w0, p0[l], t);
                t = std::fmaf(w2, p2[l], t);
                o[l] = t;
            }
        }
    });
}

// Channel-first form of the native launcher three_interpolate_gpu
// (tf_interpolate_g.cu:90-124): points (b,c,m) -> out (b,c,n).
void hf_oracle_three_interpolate_cf(int b, int c, int m, int n, const float *points, const int *idx,
                                    const float *weight, float *out) {
    for (int64_t i = 0; i < b; ++i)
        for (int64_t l = 0; l < c; ++l) {
            const float *pts = points + (i * c + l) * m;
            float *o = out + (i * c + l) * n;
            for (int64_t j = 0; j < n; ++j) {
                const float *w = weight + (i * n + j) * 3;
                const int *id = idx + (i * n + j) * 3;
                float t = w[1] * pts[id[1]];
                t = std::fmaf(w[0], pts[id[0]], t);
                t = std::fmaf(w[2], pts[id[2]], t);
                o[j] = t;
            }
        }
}

// ---------------------------------------------------------------------------
// cropping/tf_cropping_g.cu:7-41 is_point_inside, :43-132 pccropandsampleKernel,
// wrapper memsets cropping/tf_cropping.cpp:171-176 (outputs 0, non_empty 1).
// The reference slot order is an atomicInc race (:85); the canonical order
// defined for oracle and product is a serial execution of the same loop:
// inside points in ascending point index, first `resize` kept.
// bool tensors are 1 byte (TF bool), represented as uint8_t here.
// ---------------------------------------------------------------------------
static bool is_point_inside(float px, float py, float pz, float p1x, float p1y, float p1z,
                            float p2x, float p2y, float p2z, float p4x, float p4y, float p4z,
                            float p5x, float p5y, float p5z) {
    float ux = p2x - p1x, uy = p2y - p1y, uz = p2z - p1z;
    float vx = p4x - p1x, vy = p4y - p1y, vz = p4z - p1z;
    float wx = p5x - p1x, wy = p5y - p1y, wz = p5z - p1z;
    float u_dot_x = dot3(ux, uy, uz, px, py, pz);
    float u_dot_p1 = dot3(ux, uy, uz, p1x, p1y, p1z);
    float u_dot_p2 = dot3(ux, uy, uz, p2x, p2y, p2z);
    float v_dot_x = dot3(vx, vy, vz, px, py, pz);
    float v_dot_p1 = dot3(vx, vy, vz, p1x, p1y, p1z);
    float v_dot_p4 = dot3(vx, vy, vz, p4x, p4y, p4z);
    float w_dot_x = dot3(wx, wy, wz, px, py, pz);
    float w_dot_p1 = dot3(wx, wy, wz, p1x, p1y, p1z);
    float w_dot_p5 = dot3(wx, wy, wz, p5x, p5y, p5z);
    return u_dot_p1 < u_dot_x && u_dot_x < u_dot_p2 && v_dot_p1 < v_dot_x && v_dot_x < v_dot_p4 &&
           w_dot_p1 < w_dot_x && w_dot_x < w_dot_p5;
}

void hf_oracle_pc_crop_and_sample(const float *pts_data, const float *fts_data,
                                  const float *intensities_data, const uint8_t *mask_data,
                                  const float *boxes_data, const int *box_ind_data, int num_boxes,
                                  int batch, int npts, int resize, int channel,
                                  int intensity_channel, float *crop_pts_data, float *crop_fts_data,
                                  float *crop_intensities_data, uint8_t *crop_mask_data,
                                  int *crop_ind_data, uint8_t *non_empty_box_data) {
    (void)batch;
    // wrapper memsets, tf_cropping.cpp:171-176
    std::memset(crop_pts_data, 0, sizeof(float) * (size_t)num_boxes * resize * 3);
    std::memset(crop_fts_data, 0, sizeof(float) * (size_t)num_boxes * resize * channel);
    std::memset(crop_intensities_data, 0, sizeof(float) * (size_t)num_boxes * resize * intensity_channel);
    std::memset(crop_mask_data, 0, (size_t)num_boxes * resize);
    std::memset(crop_ind_data, 0, sizeof(int) * (size_t)num_boxes * resize);
    std::memset(non_empty_box_data, 1, (size_t)num_boxes);
    parallel_for(num_boxes, [&](int64_t lo, int64_t hi) {
        for (int64_t b = lo; b < hi; ++b) {
            const float *bx = boxes_data + b * 24;
            float p1x = bx[0], p1y = bx[8], p1z = bx[16];
            float p2x = bx[1], p2y = bx[9], p2z = bx[17];
            float p4x = bx[3], p4y = bx[11], p4z = bx[19];
            float p5x = bx[4], p5y = bx[12], p5z = bx[20];
            int64_t bch = box_ind_data[b];
            const float *pts = pts_data + bch * npts * 3;
            const float *fts = fts_data + bch * npts * channel;
            const float *intens = intensities_data + bch * npts * intensity_channel;
            const uint8_t *mask = mask_data + bch * npts;
            float *crop_pts = crop_pts_data + b * resize * 3;
            float *crop_fts = crop_fts_data + b * resize * channel;
            float *crop_int = crop_intensities_data + b * resize * intensity_channel;
            uint8_t *crop_mask = crop_mask_data + b * resize;
            int *crop_ind = crop_ind_data + b * resize;
            unsigned box_pts_num = 0;
            for (int p = 0; p < npts; ++p) {
                float px = pts[p * 3], py = pts[p * 3 + 1], pz = pts[p * 3 + 2];
                if (is_point_inside(px, py, pz, p1x, p1y, p1z, p2x, p2y, p2z, p4x, p4y, p4z, p5x, p5y, p5z)) {
                    unsigned pos = box_pts_num++;
                    if (pos < (unsigned)resize) {
                        crop_pts[pos * 3] = px; crop_pts[pos * 3 + 1] = py; crop_pts[pos * 3 + 2] = pz;
                        for (int c = 0; c < channel; ++c) crop_fts[(int64_t)pos * channel + c] = fts[(int64_t)p * channel + c];
                        for (int c = 0; c < intensity_channel; ++c)
                            crop_int[pos * intensity_channel + c] = intens[p * intensity_channel + c];
                        crop_mask[pos] = mask[p];
                        crop_ind[pos] = p;
                    } else {
                        break;
                    }
                }
            }
            if (box_pts_num <= 0) {
                non_empty_box_data[b] = 0;
            } else if (box_pts_num < (unsigned)resize) {
                int counter = 0;
                int init_num = box_pts_num;
                while (box_pts_num < (unsigned)resize) {
                    int s = counter++ % init_num;
                    crop_pts[box_pts_num * 3] = crop_pts[s * 3];
                    crop_pts[box_pts_num * 3 + 1] = crop_pts[s * 3 + 1];
                    crop_pts[box_pts_num * 3 + 2] = crop_pts[s * 3 + 2];
                    for (int c = 0; c < channel; ++c)
                        crop_fts[(int64_t)box_pts_num * channel + c] = crop_fts[(int64_t)s * channel + c];
                    for (int c = 0; c < intensity_channel; ++c)
                        crop_int[box_pts_num * intensity_channel + c] = crop_int[s * intensity_channel + c];
                    crop_mask[box_pts_num] = crop_mask[s];
                    crop_ind[box_pts_num] = crop_ind[s];
                    box_pts_num++;
                }
            }
        }
    });
}

// ---------------------------------------------------------------------------
// bev_iou/bev_iou_g.cu:8-224  rotated-rectangle overlap / IoU (device math
// restated for the host; float arithmetic, the same double->float narrowing
// points: Point(double,double) ctor :11-13 and fabs(area)/2.0 :205).
// ---------------------------------------------------------------------------
}  // extern "C" (the bev:: helpers below are overloaded C++)

namespace bev {
const float EPS = 1e-8;
struct Point {
    float x, y;
    Point() {}
    Point(double _x, double _y) { x = _x, y = _y; }
    void set(float _x, float _y) { x = _x; y = _y; }
    Point operator+(const Point &b) const { return Point(x + b.x, y + b.y); }
    Point operator-(const Point &b) const { return Point(x - b.x, y - b.y); }
};
inline float cross(const Point &a, const Point &b) { return a.x * b.y - a.y * b.x; }
inline float cross(const Point &p1, const Point &p2, const Point &p0) {
    return (p1.x - p0.x) * (p2.y - p0.y) - (p2.x - p0.x) * (p1.y - p0.y);
}
inline int check_rect_cross(const Point &p1, const Point &p2, const Point &q1, const Point &q2) {
    return std::min(p1.x, p2.x) <= std::max(q1.x, q2.x) && std::min(q1.x, q2.x) <= std::max(p1.x, p2.x) &&
           std::min(p1.y, p2.y) <= std::max(q1.y, q2.y) && std::min(q1.y, q2.y) <= std::max(p1.y, p2.y);
}
inline int check_in_box2d(const float *box, const Point &p) {
    const float MARGIN = 1e-5;
    float center_x = (box[0] + box[2]) / 2;
    float center_y = (box[1] + box[3]) / 2;
    float angle_cos = std::cos(-box[4]), angle_sin = std::sin(-box[4]);
    float rot_x = (p.x - center_x) * angle_cos + (p.y - center_y) * angle_sin + center_x;
    float rot_y = -(p.x - center_x) * angle_sin + (p.y - center_y) * angle_cos + center_y;
    return (rot_x > box[0] - MARGIN && rot_x < box[2] + MARGIN && rot_y > box[1] - MARGIN &&
            rot_y < box[3] + MARGIN);
}
inline int intersection(const Point &p1, const Point &p0, const Point &q1, const Point &q0, Point &ans) {
    if (check_rect_cross(p0, p1, q0, q1) == 0) return 0;
    float s1 = cross(q0, p1, p0);
    float s2 = cross(p1, q1, p0);
    float s3 = cross(p0, q1, q0);
    float s4 = cross(q1, p1, q0);
    if (!(s1 * s2 > 0 && s3 * s4 > 0)) return 0;
    float s5 = cross(q1, p1, p0);
    if (std::fabs(s5 - s1) > EPS) {
        ans.x = (s5 * q0.x - s1 * q1.x) / (s5 - s1);
        ans.y = (s5 * q0.y - s1 * q1.y) / (s5 - s1);
    } else {
        float a0 = p0.y - p1.y, b0 = p1.x - p0.x, c0 = p0.x * p1.y - p1.x * p0.y;
        float a1 = q0.y - q1.y, b1 = q1.x - q0.x, c1 = q0.x * q1.y - q1.x * q0.y;
        float D = a0 * b1 - a1 * b0;
        ans.x = (b0 * c1 - b1 * c0) / D;
        ans.y = (a1 * c0 - a0 * c1) / D;
    }
    return 1;
}
inline void rotate_around_center(const Point &center, const float angle_cos, const float angle_sin, Point &p) {
    float new_x = (p.x - center.x) * angle_cos + (p.y - center.y) * angle_sin + center.x;
    float new_y = -(p.x - center.x) * angle_sin + (p.y - center.y) * angle_cos + center.y;
    p.set(new_x, new_y);
}
inline int point_cmp(const Point &a, const Point &b, const Point &center) {
    return std::atan2(a.y - center.y, a.x - center.x) > std::atan2(b.y - center.y, b.x - center.x);
}
// bev_iou_g.cu:102-206.  cross_points is sized 24 here (16 edge pairs + 8
// corners) so the degenerate >16 case the reference would overflow on is
// defined behaviour in the oracle.
inline float box_overlap(const float *box_a, const float *box_b) {
    float a_x1 = box_a[0], a_y1 = box_a[1], a_x2 = box_a[2], a_y2 = box_a[3], a_angle = box_a[4];
    float b_x1 = box_b[0], b_y1 = box_b[1], b_x2 = box_b[2], b_y2 = box_b[3], b_angle = box_b[4];
    Point center_a((a_x1 + a_x2) / 2, (a_y1 + a_y2) / 2);
    Point center_b((b_x1 + b_x2) / 2, (b_y1 + b_y2) / 2);
    Point box_a_corners[5];
    box_a_corners[0].set(a_x1, a_y1);
    box_a_corners[1].set(a_x2, a_y1);
    box_a_corners[2].set(a_x2, a_y2);
    box_a_corners[3].set(a_x1, a_y2);
    Point box_b_corners[5];
    box_b_corners[0].set(b_x1, b_y1);
    box_b_corners[1].set(b_x2, b_y1);
    box_b_corners[2].set(b_x2, b_y2);
    box_b_corners[3].set(b_x1, b_y2);
    float a_angle_cos = std::cos(a_angle), a_angle_sin = std::sin(a_angle);
    float b_angle_cos = std::cos(b_angle), b_angle_sin = std::sin(b_angle);
    for (int k = 0; k < 4; k++) {
        rotate_around_center(center_a, a_angle_cos, a_angle_sin, box_a_corners[k]);
        rotate_around_center(center_b, b_angle_cos, b_angle_sin, box_b_corners[k]);
    }
    box_a_corners[4] = box_a_corners[0];
    box_b_corners[4] = box_b_corners[0];
    Point cross_points[24];
    Point poly_center;
    int cnt = 0, flag = 0;
    poly_center.set(0, 0);
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            flag = intersection(box_a_corners[i + 1], box_a_corners[i], box_b_corners[j + 1], box_b_corners[j],
                                cross_points[cnt]);
            if (flag) {
                poly_center = poly_center + cross_points[cnt];
                cnt++;
            }
        }
    for (int k = 0; k < 4; k++) {
        if (check_in_box2d(box_a, box_b_corners[k])) {
            poly_center = poly_center + box_b_corners[k];
            cross_points[cnt] = box_b_corners[k];
            cnt++;
        }
        if (check_in_box2d(box_b, box_a_corners[k])) {
            poly_center = poly_center + box_a_corners[k];
            cross_points[cnt] = box_a_corners[k];
            cnt++;
        }
    }
    poly_center.x /= cnt;
    poly_center.y /= cnt;
    Point temp;
    for (int j = 0; j < cnt - 1; j++)
        for (int i = 0; i < cnt - j - 1; i++)
            if (point_cmp(cross_points[i], cross_points[i + 1], poly_center)) {
                temp = cross_points[i];
                cross_points[i] = cross_points[i + 1];
                cross_points[i + 1] = temp;
            }
    float area = 0;
    for (int k = 0; k < cnt - 1; k++)
        area += cross(cross_points[k] - cross_points[0], cross_points[k + 1] - cross_points[0]);
    return std::fabs(area) / 2.0;
}
inline float iou_bev(const float *box_a, const float *box_b, const float s_overlap) {
    float sa = (box_a[2] - box_a[0]) * (box_a[3] - box_a[1]);
    float sb = (box_b[2] - box_b[0]) * (box_b[3] - box_b[1]);
    return s_overlap / std::fmax(sa + sb - s_overlap, EPS);
}
}  // namespace bev

extern "C" {

// bev_iou_g.cu:240-254 boxes_iou_bev_kernel; wrapper bev_iou.cpp:147-179
void hf_oracle_compute_bev_iou(int num_a, const float *boxes_a, int num_b, const float *boxes_b,
                               float *ans_overlap, float *ans_iou) {
    parallel_for(num_a, [&](int64_t lo, int64_t hi) {
        for (int64_t a = lo; a < hi; ++a)
            for (int64_t bq = 0; bq < num_b; ++bq) {
                float s = bev::box_overlap(boxes_a + a * 5, boxes_b + bq * 5);
                ans_overlap[a * num_b + bq] = s;
                ans_iou[a * num_b + bq] = bev::iou_bev(boxes_a + a * 5, boxes_b + bq * 5, s);
            }
    });
}

// bev_iou_g.cu:256-298 nms_kernel: full N x ceil(N/64) mask (all tiles; diagonal
// tiles only j>i).
void hf_oracle_oriented_nms_mask(const float *boxes, unsigned long long *mask, int boxes_num,
                                 float nms_overlap_thresh) {
    const int T = 64;
    const int col_blocks = (boxes_num + T - 1) / T;
    parallel_for(boxes_num, [&](int64_t lo, int64_t hi) {
        for (int64_t i = lo; i < hi; ++i) {
            int row_start = (int)(i / T), tx = (int)(i % T);
            for (int col_start = 0; col_start < col_blocks; ++col_start) {
                int col_size = std::min(boxes_num - col_start * T, T);
                unsigned long long t = 0;
                int start = (row_start == col_start) ? tx + 1 : 0;
                for (int j = start; j < col_size; ++j) {
                    const float *bj = boxes + (int64_t)(col_start * T + j) * 5;
                    float s = bev::box_overlap(boxes + i * 5, bj);
                    if (bev::iou_bev(boxes + i * 5, bj, s) > nms_overlap_thresh) t |= 1ULL << j;
                }
                mask[i * col_blocks + col_start] = t;
            }
        }
    });
}

// Host sweep + padding of OrientedNMSOp::Compute, bev_iou/bev_iou.cpp:87-112.
// Returns the number of boxes kept before padding.
int hf_oracle_nms_sweep(const unsigned long long *mask, int boxes_num, int *keep) {
    const int T = 64;
    const int col_blocks = (boxes_num + T - 1) / T;
    std::vector<unsigned long long> remv(col_blocks, 0ULL);
    int num_to_keep = 0;
    for (int i = 0; i < boxes_num; i++) {
        int nblock = i / T, inblock = i % T;
        if (!(remv[nblock] & (1ULL << inblock))) {
            keep[num_to_keep++] = i;
            const unsigned long long *p = mask + (int64_t)i * col_blocks;
            for (int j = nblock; j < col_blocks; j++) remv[j] |= p[j];
        }
    }
    int kept = num_to_keep;
    for (; num_to_keep < boxes_num; num_to_keep++) keep[num_to_keep] = keep[0];
    return kept;
}

// oriented_nms end to end (bev_iou.py:30-39 -> bev_iou.cpp:54-118).
// Only the upper-triangular tiles the sweep reads are evaluated; the result is
// identical to mask+sweep (tests check that).  min_margin (optional) receives
// min |iou - thresh| over the evaluated pairs, the quantity SURVEY A.7 asks the
// tests to report.
int hf_oracle_oriented_nms(const float *boxes, int boxes_num, float thresh, int *keep, float *min_margin) {
    const int T = 64;
    const int col_blocks = (boxes_num + T - 1) / T;
    std::vector<unsigned long long> mask((size_t)boxes_num * col_blocks, 0ULL);
    std::mutex mu;
    float global_mm = 3.0e38f;
    parallel_for(boxes_num, [&](int64_t lo, int64_t hi) {
        float mm = 3.0e38f;
        for (int64_t i = lo; i < hi; ++i) {
            int row_start = (int)(i / T);
            for (int col_start = row_start; col_start < col_blocks; ++col_start) {
                int col_size = std::min(boxes_num - col_start * T, T);
                unsigned long long t = 0;
                int start = (row_start == col_start) ? (int)(i % T) + 1 : 0;
                for (int j = start; j < col_size; ++j) {
                    const float *bj = boxes + (int64_t)(col_start * T + j) * 5;
                    float s = bev::box_overlap(boxes + i * 5, bj);
                    float iou = bev::iou_bev(boxes + i * 5, bj, s);
                    if (iou > thresh) t |= 1ULL << j;
                    float mg = std::fabs(iou - thresh);
                    if (mg < mm) mm = mg;
                }
                mask[i * col_blocks + col_start] = t;
            }
        }
        std::lock_guard<std::mutex> g(mu);
        if (mm < global_mm) global_mm = mm;
    });
    if (min_margin) *min_margin = global_mm;
    return hf_oracle_nms_sweep(mask.data(), boxes_num, keep);
}

// ---------------------------------------------------------------------------
// Gradients.  The reference kernels scatter with float
// atomicAdd, so their summation order is whatever the hardware schedules; the oracle
// runs the same loops SERIALLY (ascending entry order), which is also the order the
// CUDA path guarantees.  The outputs are zeroed first, as the TF wrappers do
// (tf_sampling.cpp:174, tf_interpolate.cpp:178).  Out-of-range indices are skipped.
// ---------------------------------------------------------------------------
// sampling/tf_sampling_g.cu:183-192  scatteraddpointKernel
void hf_oracle_gather_point_grad(int b, int n, int m, const float *out_g, const int *idx, float *inp_g) {
    std::memset(inp_g, 0, sizeof(float) * (size_t)b * n * 3);
    for (int64_t i = 0; i < b; ++i)
        for (int64_t j = 0; j < m; ++j) {
            int a = idx[i * m + j];
            if (a < 0 || a >= n) continue;
            for (int l = 0; l < 3; ++l) inp_g[(i * n + a) * 3 + l] += out_g[(i * m + j) * 3 + l];
        }
}

// grouping/tf_grouping_g.cu:61-78  group_point_grad_gpu
void hf_oracle_group_point_grad(int b, int n, int c, int m, int nsample, const float *grad_out, const int *idx,
                                float *grad_points) {
    std::memset(grad_points, 0, sizeof(float) * (size_t)b * n * c);
    for (int64_t i = 0; i < b; ++i)
        for (int64_t j = 0; j < m; ++j)
            for (int k = 0; k < nsample; ++k) {
                int ii = idx[(i * m + j) * nsample + k];
                if (ii < 0 || ii >= n) continue;
                for (int l = 0; l < c; ++l)
                    grad_points[(i * n + ii) * c + l] += grad_out[((i * m + j) * nsample + k) * c + l];
            }
}

// interpolate/tf_interpolate_g.cu:133-155  three_interpolate_grad_kernel_fast, channel-first:
// grad_out (b,c,n), idx / weight (b,n,3) -> grad_points (b,c,m); the product is rounded, then added.
void hf_oracle_three_interpolate_grad(int b, int c, int n, int m, const float *grad_out, const int *idx,
                                      const float *weight, float *grad_points) {
    std::memset(grad_points, 0, sizeof(float) * (size_t)b * c * m);
    for (int64_t i = 0; i < b; ++i)
        for (int64_t ch = 0; ch < c; ++ch)
            for (int64_t p = 0; p < n; ++p)
                for (int k = 0; k < 3; ++k) {
                    int d = idx[(i * n + p) * 3 + k];
                    if (d < 0 || d >= m) continue;
                    volatile float prod = grad_out[(i * c + ch) * n + p] * weight[(i * n + p) * 3 + k];
                    grad_points[(i * c + ch) * m + d] += prod;
                }
}

// cropping/tf_cropping_g.cu:134-150  pccropandsamplegradftsKernel
void hf_oracle_pc_crop_and_sample_grad_fts(const int *box_ind, const int *crop_ind, const float *grad_crop_fts,
                                           int num_boxes, int batch, int npts, int resize, int channel,
                                           float *grad_fts) {
    std::memset(grad_fts, 0, sizeof(float) * (size_t)batch * npts * channel);
    for (int64_t bx = 0; bx < num_boxes; ++bx) {
        int bch = box_ind[bx];
        if (bch < 0 || bch >= batch) continue;
        for (int64_t p = 0; p < resize; ++p) {
            int id = crop_ind[bx * resize + p];
            if (id < 0 || id >= npts) continue;
            for (int ch = 0; ch < channel; ++ch)
                grad_fts[((int64_t)bch * npts + id) * channel + ch] += grad_crop_fts[(bx * resize + p) * channel + ch];
        }
    }
}

// ---------------------------------------------------------------------------
// KNN grouping
// ---------------------------------------------------------------------------
// grouping/tf_grouping_g.cu:83-123  selection_sort_gpu, one row at a time
void hf_oracle_select_top_k(int b, int n, int m, int k, const float *dist, int *outi, float *out) {
    parallel_for((int64_t)b * m, [&](int64_t lo, int64_t hi) {
        for (int64_t r = lo; r < hi; ++r) {
            float *p = out + r * n;
            int *pi = outi + r * n;
            for (int s = 0; s < n; ++s) {
                p[s] = dist[r * n + s];
                pi[s] = s;
            }
            for (int s = 0; s < k && s < n; ++s) {
                int mn = s;
                for (int t = s + 1; t < n; ++t)
                    if (p[t] < p[mn]) mn = t;
                if (mn != s) {
                    std::swap(p[mn], p[s]);
                    std::swap(pi[mn], pi[s]);
                }
            }
        }
    });
}

// grouping/tf_grouping.py:62-95  knn_point: squared distances, k smallest ascending, ties to the lower
// index (tf.nn.top_k).  Distances in the direct form every other op of the path uses (sqdist), not the
// reference's r2 - 2*x2.x1 + r1 expansion, whose rounding depends on TensorFlow's matmul (absent here).
void hf_oracle_knn_point(int b, int n, int m, int k, const float *xyz1, const float *xyz2, float *val, int *idx) {
    parallel_for((int64_t)b * m, [&](int64_t lo, int64_t hi) {
        std::vector<std::pair<float, int>> d(n);
        for (int64_t q = lo; q < hi; ++q) {
            int64_t i = q / m;
            const float *c = xyz1 + i * (int64_t)n * 3;
            const float *p = xyz2 + q * 3;
            for (int j = 0; j < n; ++j) d[j] = {sqdist(p[0] - c[3 * j], p[1] - c[3 * j + 1], p[2] - c[3 * j + 2]), j};
            std::partial_sort(d.begin(), d.begin() + k, d.end());
            for (int s = 0; s < k; ++s) {
                val[q * k + s] = d[s].first;
                idx[q * k + s] = d[s].second;
            }
        }
    });
}

// ---------------------------------------------------------------------------
// prob_sample: sampling/tf_sampling_g.cu:7-104,198-201.
// The sampled index is the result of a binary search in a FLOAT prefix sum, so the prefix sum's
// association order is part of the contract.  cumsumKernel's order, restated serially: rows are cut
// into chunks of 8192; inside a chunk, quads (a,b,c,d) become (a, a+b, c+(a+b), (c+d)+(a+b)); the quad
// totals go through a work-efficient up-sweep / down-sweep scan (tree bounds exactly as :46-65); element
// = (quad prefix + scanned total of the quads before it) + running sum; the running sum is carried
// between chunks with a two-float compensated add (:79-82).
// ---------------------------------------------------------------------------
static void oracle_cumsum_row(int n, const float *inp, float *out) {
    const int CH = 8192;
    std::vector<float> q(CH), tot(CH / 4);
    float running = 0.f, running2 = 0.f;
    for (int j = 0; j < n; j += CH) {
        const int ni = std::min(n - j, CH), n24 = (ni + 3) & ~3, n2 = n24 >> 2;
        for (int k = 0; k < ni; k += 4) {
            if (k + 3 < ni) {
                const float a = inp[j + k], b = inp[j + k + 1], c = inp[j + k + 2], d = inp[j + k + 3];
                const float ab = b + a, cd = d + c;
                q[k] = a; q[k + 1] = ab; q[k + 2] = c + ab; q[k + 3] = cd + ab;
                tot[k >> 2] = q[k + 3];
            } else {
                float v = 0.f;
                for (int k2 = k; k2 < ni; ++k2) { v += inp[j + k2]; q[k2] = v; }
                for (int k2 = ni; k2 < n24; ++k2) q[k2] = v;
                tot[k >> 2] = v;
            }
        }
        int u = 0;
        for (; (2 << u) <= n2; ++u)
            for (int k = 0; k < (n2 >> (u + 1)); ++k) tot[(((k << 1) + 2) << u) - 1] += tot[(((k << 1) + 1) << u) - 1];
        for (--u; u >= 0; --u)
            for (int k = 0; k < ((n2 - (1 << u)) >> (u + 1)); ++k) tot[(((k << 1) + 3) << u) - 1] += tot[(((k << 1) + 2) << u) - 1];
        for (int k = 0; k < ni; ++k) {
            float v = q[k];
            if ((k >> 2) != 0) v += tot[(k >> 2) - 1];
            out[j + k] = v + running;
        }
        volatile float t = tot[n2 - 1] + running2;
        volatile float r2 = running + t;
        volatile float diff = r2 - running;
        running2 = t - diff;
        running = r2;
    }
}

void hf_oracle_cumsum(int b, int n, const float *inp, float *out) {
    for (int i = 0; i < b; ++i) oracle_cumsum_row(n, inp + (int64_t)i * n, out + (int64_t)i * n);
}

// binarysearchKernel, tf_sampling_g.cu:86-104.  inp_p (b,n) weights, inp_r (b,m) uniform numbers -> out (b,m).
void hf_oracle_prob_sample(int b, int n, int m, const float *inp_p, const float *inp_r, int *out) {
    std::vector<float> cs(n);
    int base = 1;
    while (base < n) base <<= 1;
    for (int64_t i = 0; i < b; ++i) {
        oracle_cumsum_row(n, inp_p + i * n, cs.data());
        for (int64_t j = 0; j < m; ++j) {
            volatile float qv = inp_r[i * m + j] * cs[n - 1];
            const float q = qv;
            int r = n - 1;
            for (int k = base; k >= 1; k >>= 1)
                if (r >= k && cs[r - k] >= q) r -= k;
            out[i * m + j] = r;
        }
    }
}

}  // extern "C"
