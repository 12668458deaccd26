/* hfops.h — C-ABI of libhfops.so: the B200 (sm_100a) implementation of the
 * HeteroFusionRCNN point-cloud operator path.
 *
 * Every entry point replaces one native launcher of the reference (the functions
 * its TF OpKernels forward-declare and its *_g.cu files define); the reference
 * interface each one stands in for is cited as file:line against the reference
 * tree.  Differences from the reference launchers, uniform across the API:
 *   - extern "C", plain pointers and ints only (no torch / TF types);
 *   - a trailing `hf_stream_t stream` (a cudaStream_t; NULL = legacy default
 *     stream, which is what the reference always used);
 *   - an int status is returned (HF_OK or a negative HF_ERR_*); nothing ever
 *     calls exit() (reference: tf_interpolate_g.cu:82-86, bev_iou.cpp:14-22);
 *   - no hidden allocation: scratch is passed in by the caller and sized by a
 *     *_workspace_bytes() query;
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - re-entrant: the only process-wide state is a launch counter (atomic) and
 *     per-device caches of device attributes; the last-error string and the
 *     tuning overrides (hf_fps_set_*, hf_spatial_set_mode, hf_nms_set_spatial_sort)
 *     are PER CALLING THREAD, so an override never changes what another thread
 *     launches.
 * Layouts are the reference's: row-major contiguous (batch, points, channels),
 * fp32 / int32, bool = 1 byte.
 */
#ifndef HFOPS_H_
#define HFOPS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void *hf_stream_t; /* cudaStream_t */

#define HF_OK 0
#define HF_ERR_INVALID_ARG (-1) /* the reference's OP_REQUIRES -> InvalidArgument cases */
#define HF_ERR_CUDA (-2)        /* a CUDA runtime call / launch failed */
#define HF_ERR_WORKSPACE (-3)   /* workspace NULL or too small */
#define HF_ERR_UNSUPPORTED (-4) /* size outside what the kernels cover */

/* ---- library info ------------------------------------------------------ */
const char *hf_version(void);
/* Message for the last non-HF_OK status returned on the calling thread. */
const char *hf_last_error(void);
/* Number of kernels this library has launched in this process (all threads). */
unsigned long long hf_launch_count(void);

/* Measurement aid (roofline denominator): runs `blocks` CTAs of 1024 threads, each thread `iters` rounds of 8
 * independent FMA chains and nothing else (packed != 0: fma.rn.f32x2, two FMAs per instruction).  Time the launch
 * with CUDA events and divide hf_fp32_probe_flops(blocks, iters, packed) by it = the FP32 non-tensor peak of
 * the device.  out: hf_fp32_probe_out_floats(blocks) floats. */
size_t hf_fp32_probe_out_floats(int blocks);
double hf_fp32_probe_flops(int blocks, int iters, int packed);
int hf_fp32_probe(int blocks, int iters, int packed, float *out, hf_stream_t stream);

/* ---- sampling ----------------------------------------------------------- */
/* Replaces farthestpointsamplingLauncher(b,n,m,inp,temp,out), sampling/tf_sampling_g.cu:203
 * (op: sampling/tf_sampling.cpp:95-123).  inp (b,n,3) -> out (b,m) int32, out[:,0]=0.
 * Tie-break is the reference's: max min-distance, then lowest (index mod 512), then
 * lowest index (tf_sampling_g.cu:146-163).  `temp` is the reference's 32*n float
 * scratch; it is not needed here (the running distances live in registers) and may
 * be NULL — the parameter is kept so the call shape matches the reference launcher.
 * Requires m >= 1 (tf_sampling.cpp:99). */
int hf_farthest_point_sample(int b, int n, int m, const float *inp, float *temp, int *out,
                             hf_stream_t stream);
/* farthest_point_sample + gather_point of one SA level (hf/core/feature_extractors/pointnet_util.py:42-44) in one
 * launch: out (b,m) as above and new_xyz (b,m,3) = inp[out].  The sampling kernel holds the coordinates of every
 * pick when it writes the index, so the separate gatherpointLauncher launch (tf_sampling_g.cu:206) disappears. */
int hf_farthest_point_sample_xyz(int b, int n, int m, const float *inp, float *temp, int *out, float *new_xyz,
                                 hf_stream_t stream);
/* The same with the cloud's spatial index (hf_cloud_index_build over inp; NULL = none): the kernel reads the cloud in
 * curve order, a warp holds a compact region of it and skips every accepted centre whose exact lower-bound distance
 * to that region cannot lower any of its running distances.  Indices and points are identical to
 * hf_farthest_point_sample_xyz (new_xyz may be NULL); clouds of 4097..65536 points, others take the plain kernel. */
int hf_farthest_point_sample_indexed(int b, int n, int m, const float *inp, const void *index, int *out, float *new_xyz,
                                     hf_stream_t stream);
/* `levels` (1..HF_FPS_PYRAMID_MAX_LEVELS) consecutive SA sampling levels in ONE launch: level j samples m[j] points
 * from the picks of level j-1 (level 0 from inp (b,n,3)), exactly like calling hf_farthest_point_sample_xyz level by
 * level (pointnet_util.py:42-44 per set-abstraction module).  m: host array of `levels` ints; idx / new_xyz: HOST
 * arrays of `levels` device pointers, (b,m[j]) int32 and (b,m[j],3) float.  Clouds of up to 16384 points with
 * m[j+1] <= m[j] run as one cluster kernel that keeps each level's picks in shared memory; other shapes are
 * issued level by level. */
#define HF_FPS_PYRAMID_MAX_LEVELS 3
int hf_fps_pyramid(int b, int levels, int n, const int *m, const float *inp, int *const *idx, float *const *new_xyz,
                   hf_stream_t stream);
/* Tuning override for the FPS launch shape (cluster size 1/2/4/6/8, 0 = heuristic). */
void hf_fps_set_cluster(int cluster);
/* Tuning override for the FPS CTA size (256 or 512 threads, 0 = heuristic). */
void hf_fps_set_threads(int threads);
/* Tuning override for the FPS algorithm.  Results are identical for every setting.
 *   algo 0 = heuristic (the multi-pick cluster kernel of csrc/fps_multi.cu where it has a shape, else algo 1);
 *   algo 1 = plain on-chip kernel (every point every iteration, csrc/sampling.cu);
 *   algo 2 = bucket-pruned kernel (csrc/fps_bucket.cu) with `warps` x `buckets_per_warp` buckets of 32 points per
 *            CTA (0,0 = heuristic shape);
 *   algo 3 = the multi-pick cluster kernel with a forced CTA shape: `warps` = warps per CTA (4, 6 or 8; 0 = the
 *            heuristic's), and for the indexed form (hf_farthest_point_sample_indexed) `buckets_per_warp` selects how
 *            the curve-ordered cloud is dealt out: 1 = a compact region per warp, any other value = a compact region per
 *            thread (the default).  A shape the kernel has no instantiation for returns HF_ERR_UNSUPPORTED.
 * Like every hf_*_set_* override (hf_fps_set_cluster / _threads / _variant, hf_gather_set_mode, hf_spatial_set_mode,
 * hf_spatial_set_pair_counter, hf_nms_set_spatial_sort) the setting belongs to the CALLING THREAD: it changes what
 * that thread launches from then on and is invisible to every other thread, so set it on the thread that makes the
 * calls.  hf_fps_get_variant reads the calling thread's setting back (any pointer may be NULL) so that a scoped
 * override can restore what was there before. */
void hf_fps_set_variant(int algo, int warps, int buckets_per_warp);
void hf_fps_get_variant(int *algo, int *warps, int *buckets_per_warp);

/* Replaces gatherpointLauncher(b,n,m,inp,idx,out), sampling/tf_sampling_g.cu:206.
 * inp (b,n,3), idx (b,m) -> out (b,m,3). */
int hf_gather_point(int b, int n, int m, const float *inp, const int *idx, float *out,
                    hf_stream_t stream);
/* Replaces probsampleLauncher(b,n,m,inp_p,inp_r,temp,out), sampling/tf_sampling_g.cu:198-201 (and
 * cumsumLauncher, :194-196: temp (b,n) receives the float32 prefix sums, same rounding as the reference's
 * scan tree; m = 0 computes only them).  inp_p (b,n) weights, inp_r (b,m) uniform numbers in [0,1)
 * -> out (b,m): smallest index whose prefix sum is >= r * total. */
int hf_prob_sample(int b, int n, int m, const float *inp_p, const float *inp_r, float *temp, int *out,
                   hf_stream_t stream);

/* The fixed-size sampling of the KITTI loader (hf/datasets/kitti/kitti_dataset.py:341-371) on the device, fed by
 * hf_lidar_fov_filter: pts (S,n_cap,4) [x,y,z,intensity] with counts (S) valid rows per scan (device memory: the scan
 * does not return to the host) -> out (S,num_out,4) [x,y,z,intensity - intensity_shift].  n > num_out: every point
 * with depth z >= depth_split (40 m) and a draw without replacement from the nearer ones; n < num_out: every point and
 * num_out - n repeats (without replacement up to 2n, with replacement beyond); then a shuffle.  numpy's RNG stream is
 * replaced by host-drawn inputs that need no knowledge of the scan:
 *   keys (S,n_cap) uint32: choice(a, k, replace=False) := the k candidates with the smallest (keys[i], i), ascending i;
 *   extra (S,num_out) uint32 (may be NULL unless num_out > 2n): choice(a, k, replace=True) := a[extra[e] % len(a)];
 *   slot_of (S,num_out) int32, a permutation per scan: shuffle(choice) := new[slot_of[j]] = old[j].
 * choice (S,num_out) int32 (may be NULL) receives the source row of every output row; status (S): 1, or 0 where the
 * reference raises (empty scan; more far points than num_out) and the rows are zero.  A slot_of entry outside
 * [0, num_out) also gives status 0 and zero rows for that scan; that slot_of holds no value twice is the caller's
 * precondition and is not checked (a duplicate leaves one row unwritten).  n_cap <= 131072. */
int hf_pc_sample_points(int scans, int n_cap, int num_out, const float *pts, const int *counts, const unsigned *keys,
                        const int *slot_of, const unsigned *extra, float depth_split, float intensity_shift, float *out,
                        int *choice, int *status, hf_stream_t stream);

/* ---- grouping ----------------------------------------------------------- */
/* Replaces queryBallPointLauncher(b,n,m,radius,nsample,xyz1,xyz2,idx,pts_cnt),
 * grouping/tf_grouping_g.cu:125.  xyz1 (b,n,3) dataset, xyz2 (b,m,3) queries ->
 * idx (b,m,nsample), pts_cnt (b,m).  First `nsample` dataset points in index order
 * with max(sqrtf(d2),1e-20f) < radius; remaining slots repeat the first hit.
 * Rows with no hit are written as 0 (the reference leaves them uninitialised,
 * tf_grouping.cpp:87-100).  Requires radius > 0, nsample > 0 (tf_grouping.cpp:71,74). */
int hf_query_ball_point(int b, int n, int m, float radius, int nsample, const float *xyz1,
                        const float *xyz2, int *idx, int *pts_cnt, hf_stream_t stream);

/* Spatially pruned forms of hf_query_ball_point / hf_three_nn (csrc/spatial.cu): identical
 * results, but both clouds are first Morton-sorted into `workspace` (caller-provided,
 * 16-byte aligned, >= hf_spatial_workspace_bytes(b, n, m) bytes, no hidden allocation) and
 * a warp of 32 neighbouring queries only visits the candidate buckets whose bounding box
 * is within reach.  With a null / short workspace, clouds beyond 16384 points, or calls too
 * small to pay for the sort they run the every-pair kernels above.
 * hf_spatial_set_mode: 0 = auto, 1 = never pruned, 2 = pruned whenever a shape exists (tests). */
size_t hf_spatial_workspace_bytes(int b, int n, int m);
void hf_spatial_set_mode(int mode);
/* Measurement aid (per calling thread): while device_counter != NULL the pruned ball query / three_nn kernels launched
 * by this thread add the (query, candidate) distance evaluations they actually execute to *device_counter (one
 * atomicAdd per warp; 1024 per visited bucket of 32 candidates x 32 queries).  bench.py divides them, not the
 * every-pair count, by the launch time for the FP32 fraction of these kernels. */
void hf_spatial_set_pair_counter(unsigned long long *device_counter);
int hf_query_ball_point_ws(int b, int n, int m, float radius, int nsample, const float *xyz1,
                           const float *xyz2, int *idx, int *pts_cnt, void *workspace,
                           size_t workspace_bytes, hf_stream_t stream);
int hf_three_nn_ws(int b, int n, int m, const float *unknown, const float *known, float *dist2,
                   int *idx, void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* The same kernels with the sort factored out, for callers that run several ops on one cloud
 * (a PointNet++ level runs ball query and three_nn on the same two clouds): build the index of a
 * cloud (b,n,3) once into caller memory (16-byte aligned, hf_cloud_index_bytes(b, n) bytes; 0 =
 * cloud too large to index) and pass it to any number of calls that take that cloud.  index1 /
 * index_unknown belong to xyz1 / unknown, index2 / index_known to xyz2 / known; a null index
 * (or a call too small to profit) runs the every-pair kernel. */
size_t hf_cloud_index_bytes(int b, int n);
/* 1 if a call with nq query points and nc candidate points would take the pruned kernels. */
int hf_spatial_applies(int nq, int nc);
int hf_cloud_index_build(int b, int n, const float *xyz, void *index, size_t index_bytes,
                         hf_stream_t stream);
int hf_query_ball_point_indexed(int b, int n, int m, float radius, int nsample, const float *xyz1,
                                const float *xyz2, int *idx, int *pts_cnt, const void *index1,
                                const void *index2, hf_stream_t stream);
int hf_three_nn_indexed(int b, int n, int m, const float *unknown, const float *known,
                        float *dist2, int *idx, const void *index_unknown,
                        const void *index_known, hf_stream_t stream);
/* three_nn fused with the inverse-distance weights its only caller computes next
 * (hf/core/feature_extractors/pointnet_util.py:303-308: d = max(d, 1e-10), w = (1/d) / sum(1/d)):
 * also writes weight (b,n,3).  dist2 is the raw squared distance as in hf_three_nn.  Both indices
 * may be NULL (every-pair kernel). */
int hf_three_nn_weights(int b, int n, int m, const float *unknown, const float *known,
                        float *dist2, int *idx, float *weight, const void *index_unknown,
                        const void *index_known, hf_stream_t stream);

/* Replaces groupPointLauncher(b,n,c,m,nsample,points,idx,out), grouping/tf_grouping_g.cu:133.
 * points (b,n,c), idx (b,m,nsample) -> out (b,m,nsample,c). */
int hf_group_point(int b, int n, int c, int m, int nsample, const float *points, const int *idx,
                   float *out, hf_stream_t stream);
/* The two group_point calls of sample_and_group (hf/core/feature_extractors/pointnet_util.py:49,56 — the
 * coordinates and the features of the same neighbourhoods) in one pass over idx: xyz (b,n,3) -> out_xyz (b,m,nsample,3)
 * and points (b,n,c) -> out_points (b,m,nsample,c).  Identical to two hf_group_point calls. */
int hf_group_point_pair(int b, int n, int c, int m, int nsample, const float *xyz, const float *points, const int *idx,
                        float *out_xyz, float *out_points, hf_stream_t stream);
/* Tuning override (per calling thread): 1 = group_point with c % 4 == 0 moves its rows with TMA bulk copies
 * (cp.async.bulk global -> shared per gathered row, one bulk copy shared -> global per stage of rows); 0 (default) =
 * the register-path kernel, which measures faster for rows of 0.4-2 KB (profiles/r2_gather_bench_tma_rows.txt).
 * Results are identical. */
void hf_gather_set_mode(int bulk);
/* The tail of sample_and_group (hf/core/feature_extractors/pointnet_util.py:49-60) in one pass:
 * out (b,m,nsample,3+c) = concat(xyz[idx] - new_xyz (translation normalisation), points[idx]); c = 0 (points
 * NULL) gives the translated grouped_xyz alone.  Replaces two group_point calls, a tile + subtract and a concat. */
int hf_group_point_concat(int b, int n, int c, int m, int nsample, const float *xyz,
                          const float *points, const int *idx, const float *new_xyz, float *out,
                          hf_stream_t stream);

/* The same pass in the column order of the multi-scale-grouping module, pointnet_sa_module_msg
 * (hf/core/feature_extractors/pointnet_util.py:253-264: per scale, group_point on xyz, subtract the seed, group_point on
 * the features, tf.concat([grouped_points, grouped_xyz])): out (b,m,nsample,c+3) = concat(points[idx], xyz[idx] -
 * new_xyz).  c = 0 (points NULL) gives the translated grouped_xyz alone, as that module does without features. */
int hf_group_point_concat_msg(int b, int n, int c, int m, int nsample, const float *xyz,
                              const float *points, const int *idx, const float *new_xyz, float *out,
                              hf_stream_t stream);

/* ---- interpolate -------------------------------------------------------- */
/* Replaces three_nn_gpu(b,n,m,unknown,known,dist2,idx), interpolate/tf_interpolate_g.cu:68.
 * unknown (b,n,3), known (b,m,3) -> dist2 (b,n,3) squared distances ascending,
 * idx (b,n,3); equal distances keep the lower index.  With m < 3 the unused
 * distances are +inf and their indices 0, as in the reference. */
int hf_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2,
                int *idx, hf_stream_t stream);

/* Replaces three_interpolate_gpu(b,c,m,n,points,idx,weight,out),
 * interpolate/tf_interpolate_g.cu:112 — CHANNEL-FIRST like that launcher:
 * points (b,c,m), idx/weight (b,n,3) -> out (b,c,n). */
int hf_three_interpolate(int b, int c, int m, int n, const float *points, const int *idx,
                         const float *weight, float *out, hf_stream_t stream);
/* Channel-last form = the Python contract of interpolate/tf_interpolate.py:26-37
 * (which wraps the launcher in two transposes; this entry makes them unnecessary):
 * points (b,m,c), idx/weight (b,n,3) -> out (b,n,c). */
int hf_three_interpolate_nlc(int b, int m, int c, int n, const float *points, const int *idx,
                             const float *weight, float *out, hf_stream_t stream);


/* ---- cropping ----------------------------------------------------------- */
/* Replaces pccropandsample_gpu(...), cropping/tf_cropping_g.cu:152 together with the
 * output memsets of its wrapper (cropping/tf_cropping.cpp:171-176): every output
 * element is written by this call, no pre-initialisation is needed.
 * pts (B,P,3), fts (B,P,C), intensities (B,P,IC), mask (B,P) bool, boxes (N,3,8)
 * corners [axis][corner], box_ind (N) -> crop_pts (N,R,3), crop_fts (N,R,C),
 * crop_intensities (N,R,IC), crop_mask (N,R) bool, crop_ind (N,R), non_empty_box (N) bool.
 * Slot order is canonical (ascending point index, first R kept); the reference's
 * order is an atomicInc race (tf_cropping_g.cu:85).  Requires resize > 0 and
 * 0 <= box_ind < batch (out-of-range box_ind is the caller's error, unchecked, as in
 * the reference). */
int hf_pc_crop_and_sample(const float *pts, const float *fts, const float *intensities,
                          const uint8_t *mask, const float *boxes, const int *box_ind,
                          int num_boxes, int batch, int npts, int resize, int channel,
                          int intensity_channel, float *crop_pts, float *crop_fts,
                          float *crop_intensities, uint8_t *crop_mask, int *crop_ind,
                          uint8_t *non_empty_box, hf_stream_t stream);
/* The same with the cloud index of pts (hf_cloud_index_build(batch, npts, pts, ...)): the select phase
 * tests bucket boxes against the RoI instead of every point.  Results are identical; a NULL index
 * (or a scan too large to index) is the plain call. */
int hf_pc_crop_and_sample_indexed(const float *pts, const float *fts, const float *intensities,
                                  const uint8_t *mask, const float *boxes, const int *box_ind,
                                  int num_boxes, int batch, int npts, int resize, int channel,
                                  int intensity_channel, float *crop_pts, float *crop_fts,
                                  float *crop_intensities, uint8_t *crop_mask, int *crop_ind,
                                  uint8_t *non_empty_box, const void *pts_index, hf_stream_t stream);

/* ---- bev_iou ------------------------------------------------------------ */
/* Replaces compute_bev_iou_gpu(num_a,boxes_a,num_b,boxes_b,ans_overlap,ans_iou),
 * bev_iou/bev_iou_g.cu:308 (+ the two memsets of bev_iou.cpp:175-176).
 * boxes (N,5) [x1,y1,x2,y2,ry] -> overlap (N,M), iou (N,M). */
int hf_compute_bev_iou(int num_a, const float *boxes_a, int num_b, const float *boxes_b,
                       float *ans_overlap, float *ans_iou, hf_stream_t stream);

/* Replaces oriented_nms_gpu(boxes,mask,boxes_num,thresh), bev_iou/bev_iou_g.cu:317.
 * mask is (N, ceil(N/64)) uint64, bit j of mask[i][cb] set iff iou(i, 64*cb+j) > thresh,
 * diagonal tiles only j > i.  upper_only != 0 skips the tiles below the diagonal
 * (written as 0) — the reference sweep never reads them (bev_iou.cpp:101-103). */
int hf_oriented_nms_mask(const float *boxes, unsigned long long *mask, int boxes_num,
                         float nms_overlap_thresh, int upper_only, hf_stream_t stream);

/* Whole of OrientedNMSOp::Compute (bev_iou/bev_iou.cpp:54-118) on the device: mask,
 * greedy sweep and keep[0]-padding, with no host round trip.  Boxes must be
 * score-sorted by the caller, as in the reference (compute_iou.py:77-80).
 * Batched over `scenes` independent box sets (the reference maps scenes with
 * tf.map_fn, rpn_model.py:683): boxes (S,N,5) -> keep (S,N) int32, num_kept (S)
 * (may be NULL).  Workspace: hf_oriented_nms_workspace_bytes(S,N). */
size_t hf_oriented_nms_workspace_bytes(int scenes, int boxes_num);
int hf_oriented_nms(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh,
                    int *keep, int *num_kept, void *workspace, size_t workspace_bytes,
                    hf_stream_t stream);
/* The same for callers that only use the first max_keep entries of the keep list (sb_nms_fn's slice,
 * hf/core/models/model_util.py:126): keep is (S,max_keep) and equals hf_oriented_nms's keep[:, :max_keep]; num_kept (S) =
 * min(boxes kept, max_keep).  The greedy sweep stops in the block where the max_keep-th box is kept. */
int hf_oriented_nms_topk(const float *boxes, int scenes, int boxes_num, float nms_overlap_thresh, int max_keep, int *keep,
                         int *num_kept, void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* Tuning override: 0 = enumerate pairs in score order (no spatial sort, no tile culling to
 * speak of), 1 = Morton order (default).  Results are identical. */
void hf_nms_set_spatial_sort(int on);
/* Measurement aid: how many box pairs the last hf_oriented_nms call that used `workspace`
 * sent to the exact polygon clip (the rest were settled by the bounding tests).
 * Synchronises `stream`; -1 on error. */
long long hf_oriented_nms_last_pair_count(int scenes, int boxes_num, const void *workspace,
                                          hf_stream_t stream);

/* ---- gradients ------------------------------ */
/* The four scatter-add gradients, deterministic: each output row is the sum of its
 * contributions in ascending entry order (= the reference's loops run serially), written
 * exactly once — no float atomics, no memset of the output (the reference wrappers issue one:
 * tf_sampling.cpp:174, tf_interpolate.cpp:178).  Entries whose index is out of range are
 * skipped.  All take caller workspace of hf_scatter_workspace_bytes(scenes, dst_rows, entries)
 * bytes, 16-byte aligned:
 *   gather_point_grad          (b, n, m)                    group_point_grad (b, n, m*nsample)
 *   three_interpolate_grad*    (b, m, 3*n)                  crop grad_fts    (1, batch*npts, num_boxes*resize) */
size_t hf_scatter_workspace_bytes(int scenes, int dst_rows, int entries);
/* Replaces scatteraddpointLauncher(b,n,m,out_g,idx,inp_g), sampling/tf_sampling_g.cu:209.
 * out_g (b,m,3), idx (b,m) -> inp_g (b,n,3). */
int hf_gather_point_grad(int b, int n, int m, const float *out_g, const int *idx, float *inp_g,
                         void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* Replaces groupPointGradLauncher(b,n,c,m,nsample,grad_out,idx,grad_points),
 * grouping/tf_grouping_g.cu:137.  grad_out (b,m,nsample,c), idx (b,m,nsample) -> grad_points (b,n,c). */
int hf_group_point_grad(int b, int n, int c, int m, int nsample, const float *grad_out,
                        const int *idx, float *grad_points, void *workspace,
                        size_t workspace_bytes, hf_stream_t stream);
/* Replaces three_interpolate_grad_gpu(b,c,n,m,grad_out,idx,weight,grad_points),
 * interpolate/tf_interpolate_g.cu:157 (channel-first): grad_out (b,c,n), idx / weight (b,n,3)
 * -> grad_points (b,c,m).  _nlc is the channel-last form that matches the Python API's
 * tensors: grad_out (b,n,c) -> grad_points (b,m,c). */
int hf_three_interpolate_grad(int b, int c, int n, int m, const float *grad_out, const int *idx,
                              const float *weight, float *grad_points, void *workspace,
                              size_t workspace_bytes, hf_stream_t stream);
int hf_three_interpolate_grad_nlc(int b, int n, int c, int m, const float *grad_out,
                                  const int *idx, const float *weight, float *grad_points,
                                  void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* Replaces pccropandsamplegradfts_gpu(box_ind,crop_ind,grad_crop_fts,num_boxes,npts,resize,
 * channel,grad_fts), cropping/tf_cropping_g.cu:161 (+ `batch`, which the reference takes from
 * the fts tensor, tf_cropping.cpp:225).  box_ind (N), crop_ind (N,R), grad_crop_fts (N,R,C)
 * -> grad_fts (batch,npts,C). */
int hf_pc_crop_and_sample_grad_fts(const int *box_ind, const int *crop_ind,
                                   const float *grad_crop_fts, int num_boxes, int batch, int npts,
                                   int resize, int channel, float *grad_fts, void *workspace,
                                   size_t workspace_bytes, hf_stream_t stream);

/* ---- KNN grouping --------------------------- */
/* knn_point (grouping/tf_grouping.py:62-95): the k nearest points of xyz1 (b,n,3) for every query
 * of xyz2 (b,m,3) -> val (b,m,k) squared distances ascending, idx (b,m,k); ties go to the lower
 * index (tf.nn.top_k's rule).  The reference builds the whole (b,m,n) matrix as r2 - 2*x2.x1^T + r1
 * with a TF matmul and runs top_k on the CPU; here distance and selection are fused and the matrix
 * never exists.  val is the directly computed squared distance (the expanded form's cancellation
 * error, ~1e-7 * |x|^2, is not reproduced; tests pin the op on the reference formula restated in float32 within
 * that error).  Requires 1 <= k <= min(n, 128); clouds of up to 65536 points take the pruned search.  workspace =
 * hf_knn_workspace_bytes(b, n, m) bytes (0 = the call is too small to profit from sorting) or NULL
 * (then every pair is tested). */
size_t hf_knn_workspace_bytes(int b, int n, int m);
int hf_knn_point(int b, int n, int m, int k, const float *xyz1, const float *xyz2, float *val,
                 int *idx, void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* knn_indices_general(queries, points, k, sort=True, unique=True) of PointCNN (hf/core/pointfly.py:185-214, called from
 * hf/core/feature_extractors/pointcnn.py:72): the k nearest `points` (b,n,3) of every query (b,p,3), ascending ->
 * distances (b,p,k) and indices (b,p,k,2) int32 = (batch index, point index) pairs, the layout tf.gather_nd takes.
 * `unique` has no effect in the reference (prepare_for_unique_top_k rebinds a local, pointfly.py:162-164).  Same
 * kernel, workspace (hf_knn_workspace_bytes(b, n, p)) and value definition as hf_knn_point. */
int hf_knn_indices_general(int b, int p, int n, int k, const float *queries, const float *points, float *distances,
                           int *indices, void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* Replaces selectionSortLauncher(b,n,m,k,dist,outi,out), grouping/tf_grouping_g.cu:129: k steps of
 * selection sort on every row of dist (b,m,n) -> outi (b,m,n) int32, out (b,m,n); the first k
 * entries are the k smallest in order, the tail is the permutation the swaps leave (bit-exact,
 * tail included).  Rows of up to 25600 elements. */
int hf_select_top_k(int b, int n, int m, int k, const float *dist, int *outi, float *out,
                    hf_stream_t stream);

/* ---- glue of hf/core/compute_iou.py --------- */
/* boxes3d_to_bev_tf, compute_iou.py:7-20: (N,7) [x,y,z,l,w,h,ry] -> (N,5) [x-l/2, z-w/2, x+l/2, z+w/2, ry]. */
int hf_boxes3d_to_bev(long long num, const float *boxes3d, float *boxes_bev, hf_stream_t stream);
/* tf_box_3d_to_box_8co, hf/core/box_8c_encoder.py:101-184, with the proposal expansion its caller applies
 * first (hf/core/models/rcnn_model.py:462-476: y += e, l/w/h += 2e; e = 0 for the plain conversion):
 * boxes3d (N,7) -> corners (N,3,8), the box input of hf_pc_crop_and_sample. */
int hf_boxes3d_to_corners(long long num, const float *boxes3d, float expand_length, float *corners,
                          hf_stream_t stream);
/* box3d_iou_tf, compute_iou.py:23-64: boxes_a (N,7), boxes_b (M,7) -> iou_3d (N,M), iou_2d (N,M)
 * (BEV overlap x height overlap over the union of the volumes, clipped at 1e-7). */
size_t hf_box3d_iou_workspace_bytes(int num_a, int num_b);
int hf_box3d_iou(int num_a, const float *boxes_a, int num_b, const float *boxes_b, float *iou_3d,
                 float *iou_2d, void *workspace, size_t workspace_bytes, hf_stream_t stream);
/* hf_box3d_iou for callers that have no device memory to lend and no way to get any.  This entry has NO counterpart
 * in the reference; it exists because hf_box3d_iou needs a staging buffer for the two BEV box lists, and some hosts
 * cannot supply one: a plugin running inside another framework's executor that forbids cudaMalloc on its worker
 * threads (an allocation synchronises the whole device and stalls every other stream), a real-time loop whose memory
 * is fixed at start-up, a process whose memory pool is already handed out in full.  Here the staging buffer is a fixed array inside the library's own device image (2 x 8192 x 20 bytes =
 * 320 KiB; every GPU that loads libhfops.so keeps it until the process exits, whether or not this entry is ever
 * called).  The call makes no cudaMalloc / cudaFree and calls nothing that allocates; inputs and outputs are still
 * the caller's, and the results are those of hf_box3d_iou.
 *   - Limit: num_a and num_b <= HF_BOX3D_IOU_STATIC_MAX_BOXES each; more returns HF_ERR_UNSUPPORTED (hf_last_error
 *     names the limit) and launches nothing.  Larger problems: hf_box3d_iou with a workspace.
 *   - First call on each device: the library looks up the device address of its array (cudaGetSymbolAddress, with
 *     that device current) and remembers it per device; a failed look-up returns HF_ERR_CUDA.  Make that first call
 *     during set-up, before any stream capture begins.
 *   - One array per device, shared by every call and every thread: two calls whose kernels can be on the GPU at the
 *     same time (different streams, or different host threads on different streams) overwrite each other's staging
 *     data and give wrong answers.  Keep calls on one stream, or order them with events.  The array keeps the
 *     previous call's boxes; nothing in it has to start at zero, because each call writes the num_a + num_b entries
 *     it is about to read before it reads them. */
#define HF_BOX3D_IOU_STATIC_MAX_BOXES 8192
int hf_box3d_iou_static(int num_a, const float *boxes_a, int num_b, const float *boxes_b, float *iou_3d, float *iou_2d,
                        hf_stream_t stream);
/* oriented_nms_tf, compute_iou.py:67-80, batched over scenes (the reference maps scenes with
 * tf.map_fn, rpn_model.py:683): boxes3d (S,N,7), scores (S,N) in any order -> keep (S,N) indices
 * into the caller's box order (descending score, suppressed boxes removed, tail padded with
 * keep[0] like bev_iou.cpp:104-112), num_kept (S) (may be NULL).  Equal scores keep the lower
 * index first (tf.nn.top_k).  At most 25600 boxes per scene. */
size_t hf_oriented_nms_scored_workspace_bytes(int scenes, int boxes_num);
int hf_oriented_nms_scored(const float *boxes3d, const float *scores, int scenes, int boxes_num,
                           float nms_overlap_thresh, int *keep, int *num_kept, void *workspace,
                           size_t workspace_bytes, hf_stream_t stream);
/* sb_nms_fn (hf/core/models/model_util.py:101-142, mapped over the scenes of a batch by rpn_model.py:683): scored
 * oriented NMS as above, then the first nms_size entries of the keep list; with fixed_num_proposal_nms = 0 the
 * duplicate padding of the keep list is dropped (tf.unique) and the selection is padded with -1.
 * nms_indices (S,nms_size) int32 indices into the caller's boxes or -1; num_selected (S) = entries before the -1
 * padding (min(nms_size, N) when fixed, min(nms_size, kept) otherwise).  workspace: hf_sb_nms_workspace_bytes. */
size_t hf_sb_nms_workspace_bytes(int scenes, int boxes_num);
int hf_sb_nms(const float *boxes3d, const float *scores, int scenes, int boxes_num, float thresh, int nms_size,
              int fixed_num_proposal_nms, int *nms_indices, int *num_selected, void *workspace, size_t workspace_bytes,
              hf_stream_t stream);

/* ---- input side --------------------------- */
/* get_lidar_point_cloud with im_size, hf/core/obj_utils.py:221-277 (lidar_to_cam_frame, hf/core/calib_utils.py:
 * 370-407; project_to_image, :280-296): points (S,n,4) raw velodyne [x,y,z,i]; calib (S,24) float64 = the first
 * three rows of r0_rect(4x4) @ tr_velodyne_to_cam(4x4) followed by p2 (3,4), row major -> out (S,n,4)
 * [x,y,z,i] in the rectified camera frame, the points with z > 0 that project inside (0,w)x(0,h), in scan
 * order, count (S) of them (rows past count[s] are not written).  Arithmetic in float64 like the reference. */
size_t hf_lidar_fov_filter_workspace_bytes(int scans, int n);
int hf_lidar_fov_filter(int scans, int n, const float *points, const double *calib, int im_w, int im_h,
                        float *out, int *count, void *workspace, size_t workspace_bytes, hf_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HFOPS_H_ */
