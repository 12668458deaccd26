"""numpy/ctypes bindings of the CPU oracle (oracle/hf_oracle.cpp) and of oracle/_ref.

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.  Every function takes and returns
host numpy arrays; shapes and dtypes follow the reference's Python API.
"""
import ctypes
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_DIR, "libhf_oracle.so")
REF_DIR = os.path.join(_DIR, "_ref")

_i, _f, _p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference is present) via oracle/Makefile."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(_DIR, "hf_oracle.cpp")):
        subprocess.run(["make", "-C", _DIR, "libhf_oracle.so"], check=True, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT)
    if os.path.isdir("/root/reference"):
        subprocess.run(["make", "-C", _DIR, "ref"], check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return ORACLE_SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build()
        _lib = ctypes.CDLL(ORACLE_SO)
        _lib.hf_oracle_oriented_nms.restype = _i
        _lib.hf_oracle_nms_sweep.restype = _i
        _lib.hf_oracle_get_threads.restype = _i
    return _lib


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a):
    return a.ctypes.data_as(_p)


def set_threads(n):
    lib().hf_oracle_set_threads(_i(int(n)))


# ---- oracle ops ------------------------------------------------------------------
def farthest_point_sample(npoint, inp):
    inp = _c(inp, np.float32)
    b, n, _ = inp.shape
    out = np.zeros((b, npoint), np.int32)
    lib().hf_oracle_farthest_point_sample(_i(b), _i(n), _i(npoint), _ptr(inp), _ptr(out))
    return out


def gather_point(inp, idx):
    inp, idx = _c(inp, np.float32), _c(idx, np.int32)
    b, n, _ = inp.shape
    m = idx.shape[1]
    out = np.zeros((b, m, 3), np.float32)
    lib().hf_oracle_gather_point(_i(b), _i(n), _i(m), _ptr(inp), _ptr(idx), _ptr(out))
    return out


def query_ball_point(radius, nsample, xyz1, xyz2):
    xyz1, xyz2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    idx = np.zeros((b, m, nsample), np.int32)
    cnt = np.zeros((b, m), np.int32)
    lib().hf_oracle_query_ball_point(_i(b), _i(n), _i(m), _f(radius), _i(nsample), _ptr(xyz1), _ptr(xyz2), _ptr(idx),
                                     _ptr(cnt))
    return idx, cnt


def group_point(points, idx):
    points, idx = _c(points, np.float32), _c(idx, np.int32)
    b, n, c = points.shape
    m, ns = idx.shape[1], idx.shape[2]
    out = np.zeros((b, m, ns, c), np.float32)
    lib().hf_oracle_group_point(_i(b), _i(n), _i(c), _i(m), _i(ns), _ptr(points), _ptr(idx), _ptr(out))
    return out


def three_nn(xyz1, xyz2):
    xyz1, xyz2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    dist = np.zeros((b, n, 3), np.float32)
    idx = np.zeros((b, n, 3), np.int32)
    lib().hf_oracle_three_nn(_i(b), _i(n), _i(m), _ptr(xyz1), _ptr(xyz2), _ptr(dist), _ptr(idx))
    return dist, idx


def three_interpolate(points, idx, weight):
    points, idx, weight = _c(points, np.float32), _c(idx, np.int32), _c(weight, np.float32)
    b, m, c = points.shape
    n = idx.shape[1]
    out = np.zeros((b, n, c), np.float32)
    lib().hf_oracle_three_interpolate(_i(b), _i(m), _i(c), _i(n), _ptr(points), _ptr(idx), _ptr(weight), _ptr(out))
    return out


def three_interpolate_channel_first(points, idx, weight):
    points, idx, weight = _c(points, np.float32), _c(idx, np.int32), _c(weight, np.float32)
    b, c, m = points.shape
    n = idx.shape[1]
    out = np.zeros((b, c, n), np.float32)
    lib().hf_oracle_three_interpolate_cf(_i(b), _i(c), _i(m), _i(n), _ptr(points), _ptr(idx), _ptr(weight), _ptr(out))
    return out


def pc_crop_and_sample(pts, fts, intensities, mask, boxes, box_ind, resize):
    pts, fts, intensities = _c(pts, np.float32), _c(fts, np.float32), _c(intensities, np.float32)
    mask = _c(mask, np.uint8)
    boxes, box_ind = _c(boxes, np.float32), _c(box_ind, np.int32)
    B, P, C = fts.shape
    IC = intensities.shape[2]
    N, R = boxes.shape[0], int(resize)
    crop_pts = np.zeros((N, R, 3), np.float32)
    crop_fts = np.zeros((N, R, C), np.float32)
    crop_int = np.zeros((N, R, IC), np.float32)
    crop_mask = np.zeros((N, R), np.uint8)
    crop_ind = np.zeros((N, R), np.int32)
    non_empty = np.zeros((N,), np.uint8)
    lib().hf_oracle_pc_crop_and_sample(_ptr(pts), _ptr(fts), _ptr(intensities), _ptr(mask), _ptr(boxes), _ptr(box_ind),
                                       _i(N), _i(B), _i(P), _i(R), _i(C), _i(IC), _ptr(crop_pts), _ptr(crop_fts),
                                       _ptr(crop_int), _ptr(crop_mask), _ptr(crop_ind), _ptr(non_empty))
    return crop_pts, crop_fts, crop_int, crop_mask.astype(bool), crop_ind, non_empty.astype(bool)


def compute_bev_iou(proposals, gt_bboxes):
    a, g = _c(proposals, np.float32), _c(gt_bboxes, np.float32)
    n, m = a.shape[0], g.shape[0]
    overlap = np.zeros((n, m), np.float32)
    iou = np.zeros((n, m), np.float32)
    lib().hf_oracle_compute_bev_iou(_i(n), _ptr(a), _i(m), _ptr(g), _ptr(overlap), _ptr(iou))
    return overlap, iou


def oriented_nms_mask(boxes, thresh):
    boxes = _c(boxes, np.float32)
    n = boxes.shape[0]
    mask = np.zeros((n, (n + 63) // 64), np.uint64)
    lib().hf_oracle_oriented_nms_mask(_ptr(boxes), _ptr(mask), _i(n), _f(thresh))
    return mask


def nms_sweep(mask):
    mask = _c(mask, np.uint64)
    n = mask.shape[0]
    keep = np.zeros((n,), np.int32)
    kept = lib().hf_oracle_nms_sweep(_ptr(mask), _i(n), _ptr(keep))
    return keep, kept


def oriented_nms(boxes, thresh, return_info=False):
    """-> keep (N,) int32 [, (num_kept, min |iou - thresh| over evaluated pairs)]."""
    boxes = _c(boxes, np.float32)
    n = boxes.shape[0]
    keep = np.zeros((n,), np.int32)
    margin = ctypes.c_float(0)
    kept = lib().hf_oracle_oriented_nms(_ptr(boxes), _i(n), _f(thresh), _ptr(keep), ctypes.byref(margin))
    return (keep, (kept, float(margin.value))) if return_info else keep


# ---- oracle/_ref: the reference's own sources, compiled where they lie --------------

# ---- gradients (serial = ascending entry order) ----------------------------------------
def gather_point_grad(inp, idx, out_g):
    idx, out_g = _c(idx, np.int32), _c(out_g, np.float32)
    b, n, _ = inp.shape
    m = idx.shape[1]
    inp_g = np.zeros((b, n, 3), np.float32)
    lib().hf_oracle_gather_point_grad(_i(b), _i(n), _i(m), _ptr(out_g), _ptr(idx), _ptr(inp_g))
    return inp_g


def group_point_grad(points, idx, grad_out):
    idx, grad_out = _c(idx, np.int32), _c(grad_out, np.float32)
    b, n, c = points.shape
    m, ns = idx.shape[1], idx.shape[2]
    g = np.zeros((b, n, c), np.float32)
    lib().hf_oracle_group_point_grad(_i(b), _i(n), _i(c), _i(m), _i(ns), _ptr(grad_out), _ptr(idx), _ptr(g))
    return g


def three_interpolate_grad_channel_first(points, idx, weight, grad_out):
    """points (b,c,m), idx / weight (b,n,3), grad_out (b,c,n) -> (b,c,m): the reference op's layout."""
    idx, weight, grad_out = _c(idx, np.int32), _c(weight, np.float32), _c(grad_out, np.float32)
    b, c, m = points.shape
    n = idx.shape[1]
    g = np.zeros((b, c, m), np.float32)
    lib().hf_oracle_three_interpolate_grad(_i(b), _i(c), _i(n), _i(m), _ptr(grad_out), _ptr(idx), _ptr(weight), _ptr(g))
    return g


def three_interpolate_grad(points, idx, weight, grad_out):
    """Channel-last form of the Python API: points (b,m,c), grad_out (b,n,c) -> (b,m,c)."""
    g = three_interpolate_grad_channel_first(np.transpose(points, (0, 2, 1)), idx, weight,
                                             np.transpose(grad_out, (0, 2, 1)))
    return np.ascontiguousarray(np.transpose(g, (0, 2, 1)))


def pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, grad_crop_fts):
    box_ind, crop_ind, grad_crop_fts = _c(box_ind, np.int32), _c(crop_ind, np.int32), _c(grad_crop_fts, np.float32)
    batch, npts, c = fts.shape
    nb, r = crop_ind.shape
    g = np.zeros((batch, npts, c), np.float32)
    lib().hf_oracle_pc_crop_and_sample_grad_fts(_ptr(box_ind), _ptr(crop_ind), _ptr(grad_crop_fts), _i(nb), _i(batch),
                                                _i(npts), _i(r), _i(c), _ptr(g))
    return g



# ---- KNN grouping -------------------------------------------------------------------------
def select_top_k(k, dist):
    dist = _c(dist, np.float32)
    b, m, n = dist.shape
    outi = np.zeros((b, m, n), np.int32)
    out = np.zeros((b, m, n), np.float32)
    lib().hf_oracle_select_top_k(_i(b), _i(n), _i(m), _i(int(k)), _ptr(dist), _ptr(outi), _ptr(out))
    return outi, out


def knn_point(k, xyz1, xyz2):
    xyz1, xyz2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    val = np.zeros((b, m, k), np.float32)
    idx = np.zeros((b, m, k), np.int32)
    lib().hf_oracle_knn_point(_i(b), _i(n), _i(m), _i(int(k)), _ptr(xyz1), _ptr(xyz2), _ptr(val), _ptr(idx))
    return val, idx


def knn_point_reference_form(k, xyz1, xyz2):
    """grouping/tf_grouping.py:78-93 restated operation for operation in numpy float32 (TensorFlow is not installable
    here): r1 = sum(xyz1*xyz1), r2 = sum(xyz2*xyz2), mul = xyz2 @ xyz1^T, dist = r2 - 2*mul + r1^T, then
    tf.nn.top_k(-dist, k) = the k smallest, equal values in ascending index order.  Returns (val, idx, dist, tol):
    dist is the whole (b,m,n) matrix and tol (b,m,1) bounds |dist - exact squared distance| — the expanded form
    cancels, and TF's matmul may add its three products in another order than numpy's, so both the reference's own
    values and any comparison with them are only defined up to tol = 8 eps (r2 + max r1)."""
    xyz1, xyz2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    r1 = np.sum(xyz1 * xyz1, axis=2, keepdims=True, dtype=np.float32)  # (b,n,1)
    r2 = np.sum(xyz2 * xyz2, axis=2, keepdims=True, dtype=np.float32)  # (b,m,1)
    mul = np.matmul(xyz2, np.transpose(xyz1, (0, 2, 1))).astype(np.float32)  # (b,m,n)
    dist = (r2 - np.float32(2) * mul + np.transpose(r1, (0, 2, 1))).astype(np.float32)
    order = np.argsort(dist, axis=2, kind="stable")[:, :, :k]
    val = np.take_along_axis(dist, order, axis=2)
    tol = (8 * np.finfo(np.float32).eps) * (r2 + r1.max(axis=1, keepdims=True))
    return val, order.astype(np.int32), dist, tol.astype(np.float32)


def check_knn_against_reference_form(k, xyz1, xyz2, val, idx):
    """The pin of knn_point on the reference's own formula.  (val, idx) = the implementation under test (direct
    squared distances).  Asserts, per query: every value within tol of the reference-form value of the SAME index;
    the values ascending; every reference-form neighbour that is missing from idx (and every extra one) lies within
    2 tol of the k-th distance; and wherever consecutive reference-form distances are more than 2 tol apart on both
    sides, the index at that rank is the reference's.  Returns the fraction of ranks that were compared exactly."""
    rv, ri, dist, tol = knn_point_reference_form(k, xyz1, xyz2)
    val, idx = np.asarray(val), np.asarray(idx).astype(np.int64)
    same = np.take_along_axis(dist, idx, axis=2)
    assert np.all(np.abs(same - val) <= tol), "values differ from the reference form by more than its own error"
    assert np.all(np.diff(val, axis=2) >= 0)
    n = dist.shape[2]
    srt = np.sort(dist, axis=2)
    kth = srt[:, :, k - 1:k]
    inside = dist <= kth - 2 * tol  # certainly among the k nearest
    outside = dist >= kth + 2 * tol  # certainly not
    got = np.zeros(dist.shape, bool)
    np.put_along_axis(got, idx, True, axis=2)
    assert not np.any(inside & ~got), "a certain neighbour is missing"
    assert not np.any(outside & got), "a certain non-neighbour was returned"
    lo = np.concatenate([np.full_like(srt[:, :, :1], -np.inf), srt[:, :, :k - 1]], axis=2)
    hi = srt[:, :, 1:k + 1] if n > k else np.concatenate([srt[:, :, 1:k], np.full_like(srt[:, :, :1], np.inf)], axis=2)
    cur = srt[:, :, :k]
    decided = (cur - lo > 2 * tol) & (hi - cur > 2 * tol)
    assert np.array_equal(idx[decided], ri.astype(np.int64)[decided]), "index differs at a rank the reference form decides"
    return float(decided.mean())


def _mangled(so, name):
    out = subprocess.run(["nm", "-D", "--defined-only", so], check=True, stdout=subprocess.PIPE, text=True).stdout
    for line in out.splitlines():
        sym = line.split()[-1]
        if sym.startswith("_Z%d%s" % (len(name), name)):
            return sym
    raise KeyError("%s not exported by %s" % (name, so))


def ref_available(which):
    return os.path.exists(os.path.join(REF_DIR, which))


_ref_cache = {}


def _ref(which):
    if which not in _ref_cache:
        _ref_cache[which] = ctypes.CDLL(os.path.join(REF_DIR, which))
    return _ref_cache[which]


def ref_interpolate_cpu(points, idx, weight):
    """interpolate/interpolate.cpp:84-104 interpolate_cpu, the reference's own CPU code."""
    so = os.path.join(REF_DIR, "libref_interpolate_cpu.so")
    fn = getattr(_ref("libref_interpolate_cpu.so"), _mangled(so, "interpolate_cpu"))
    points, idx, weight = _c(points, np.float32), _c(idx, np.int32), _c(weight, np.float32)
    b, m, c = points.shape
    n = idx.shape[1]
    out = np.zeros((b, n, c), np.float32)
    fn(_i(b), _i(m), _i(c), _i(n), _ptr(points), _ptr(idx), _ptr(weight), _ptr(out))
    return out


def ref_query_ball_point_cpu(radius, nsample, xyz1, xyz2):
    """grouping/test/query_ball_point.cpp:19-47 (no pts_cnt output)."""
    so = os.path.join(REF_DIR, "libref_grouping_cpu.so")
    fn = getattr(_ref("libref_grouping_cpu.so"), _mangled(so, "query_ball_point_cpu"))
    xyz1, xyz2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    idx = np.zeros((b, m, nsample), np.int32)
    fn(_i(b), _i(n), _i(m), _f(radius), _i(nsample), _ptr(xyz1), _ptr(xyz2), _ptr(idx))
    return idx


def ref_group_point_cpu(points, idx):
    """grouping/test/query_ball_point.cpp:52-66."""
    so = os.path.join(REF_DIR, "libref_grouping_cpu.so")
    fn = getattr(_ref("libref_grouping_cpu.so"), _mangled(so, "group_point_cpu"))
    points, idx = _c(points, np.float32), _c(idx, np.int32)
    b, n, c = points.shape
    m, ns = idx.shape[1], idx.shape[2]
    out = np.zeros((b, m, ns, c), np.float32)
    fn(_i(b), _i(n), _i(c), _i(m), _i(ns), _ptr(points), _ptr(idx), _ptr(out))
    return out



def ref_selection_sort_cpu(k, dist):
    """grouping/test/selection_sort.cpp:19-63 (the reference's own CPU prototype; it prints every row
    it sorts, so keep the inputs small)."""
    so = os.path.join(REF_DIR, "libref_selection_sort_cpu.so")
    fn = getattr(_ref("libref_selection_sort_cpu.so"), _mangled(so, "selection_sort_cpu"))
    dist = _c(dist, np.float32)
    b, m, n = dist.shape
    idx = np.zeros((b, m, n), np.int32)
    val = np.zeros((b, m, n), np.float32)
    fn(_i(b), _i(n), _i(m), _i(int(k)), _ptr(dist), _ptr(idx), _ptr(val))
    return idx, val


def ref_bev_host(boxes_a, boxes_b):
    """bev_iou/bev_iou_g.cu:102-224 box_overlap / iou_bev compiled for the host (oracle/ref_bev_host.cu)."""
    L = _ref("libref_bev_host.so")
    L.ref_host_box_overlap.restype = _f
    L.ref_host_iou_bev.restype = _f
    a, g = _c(boxes_a, np.float32), _c(boxes_b, np.float32)
    overlap = np.zeros((a.shape[0], g.shape[0]), np.float32)
    iou = np.zeros_like(overlap)
    for i in range(a.shape[0]):
        for j in range(g.shape[0]):
            pa, pb = _ptr(a[i]), _ptr(g[j])
            overlap[i, j] = L.ref_host_box_overlap(pa, pb)
            iou[i, j] = L.ref_host_iou_bev(pa, pb)
    return overlap, iou


# ---- glue of hf/core/compute_iou.py (numpy restatement, float32 throughout) ----------------
def boxes3d_to_bev(boxes3d):
    """compute_iou.py:7-20."""
    b = np.asarray(boxes3d, np.float32)
    hl, hw = b[:, 3] / np.float32(2), b[:, 4] / np.float32(2)
    return np.stack([b[:, 0] - hl, b[:, 2] - hw, b[:, 0] + hl, b[:, 2] + hw, b[:, 6]], axis=1).astype(np.float32)


def box3d_iou(boxes_a, boxes_b):
    """compute_iou.py:23-64."""
    a, b = np.asarray(boxes_a, np.float32), np.asarray(boxes_b, np.float32)
    overlap, iou2d = compute_bev_iou(boxes3d_to_bev(a), boxes3d_to_bev(b))
    amin, amax = (a[:, 1] - a[:, 5])[:, None], a[:, 1][:, None]
    bmin, bmax = (b[:, 1] - b[:, 5])[None, :], b[:, 1][None, :]
    oh = np.clip(np.minimum(amax, bmax) - np.maximum(amin, bmin), 0, np.finfo(np.float32).max).astype(np.float32)
    o3 = overlap * oh
    va, vb = (a[:, 3] * a[:, 4] * a[:, 5])[:, None], (b[:, 3] * b[:, 4] * b[:, 5])[None, :]
    iou3d = o3 / np.clip(va + vb - o3, np.float32(1e-7), np.finfo(np.float32).max)
    return iou3d.astype(np.float32), iou2d


def oriented_nms_scored(boxes3d, scores, thresh):
    """compute_iou.py:67-80: tf.nn.top_k over all scores = stable descending sort."""
    order = np.argsort(-np.asarray(scores, np.float32), kind="stable").astype(np.int32)
    keep_sorted = oriented_nms(boxes3d_to_bev(np.asarray(boxes3d, np.float32)[order]), thresh)
    return order[keep_sorted]


def pc_sample_points(pts, count, num_out, keys, slot_of, extra=None, depth_split=40.0, intensity_shift=0.5):
    """hf/datasets/kitti/kitti_dataset.py:341-371 restated for ONE scan with the np.random calls replaced by the
    host-input definitions of include/hfops.h:hf_pc_sample_points.  pts (n_cap,4) float32 [x,y,z,intensity], count =
    valid rows.  Returns (sampled (P,4) float32, choice (P) int32, ok); ok = 0 (zeros, -1) where the reference raises."""
    pts = np.asarray(pts, np.float32)
    n, P = int(count), int(num_out)
    keys = np.asarray(keys).astype(np.uint32)[:n]

    def draw(cand, k):  # choice(cand, k, replace=False): k smallest (key, index), ascending index
        order = np.lexsort((cand, keys[cand]))[:k]
        return np.sort(cand[order])

    bad = (np.zeros((P, 4), np.float32), np.full((P,), -1, np.int32), 0)
    if P == 0:
        return np.zeros((0, 4), np.float32), np.zeros((0,), np.int32), 1
    if n == 0:
        return bad
    idx = np.arange(n)
    if P < n:
        near = pts[:n, 2] < np.float32(depth_split)
        far_idx, near_idx = idx[~near], idx[near]
        if P - len(far_idx) < 0:
            return bad
        choice = np.concatenate([draw(near_idx, P - len(far_idx)), far_idx])
    else:
        choice = idx
        if P > n:
            if P <= 2 * n:
                more = draw(idx, P - n)
            else:
                more = (np.asarray(extra).astype(np.uint32)[: P - n] % np.uint32(n)).astype(np.int64)
            choice = np.concatenate([choice, more])
    shuffled = np.empty_like(choice)
    shuffled[np.asarray(slot_of, np.int64)] = choice
    out = pts[shuffled].copy()
    out[:, 3] = out[:, 3] - np.float32(intensity_shift)
    return out, shuffled.astype(np.int32), 1


def sb_nms_fn(boxes3d, scores, thresh, nms_size, fixed_num_proposal_nms):
    """hf/core/models/model_util.py:124-142 restated: slice, tf.unique (first occurrences, in order), -1 padding."""
    idx = oriented_nms_scored(boxes3d, scores, thresh)
    idx = idx[: min(nms_size, len(idx))]
    if not fixed_num_proposal_nms:
        _, first = np.unique(idx, return_index=True)
        idx = idx[np.sort(first)]
    num = len(idx)
    if nms_size > num:
        idx = np.concatenate([idx, np.full((nms_size - num,), -1, idx.dtype)])
    return idx.astype(np.int32), num


def box_3d_to_box_8co(boxes_3d, expand_length=0.0):
    """hf/core/box_8c_encoder.py:101-184 after the expansion of rcnn_model.py:462-476 -> (N,3,8), float32."""
    b = np.asarray(boxes_3d, np.float32).copy()
    e = np.float32(expand_length)
    b[:, 1] += e
    b[:, 3:6] += np.float32(2) * e
    hl, hw, h = b[:, 3] / np.float32(2), b[:, 4] / np.float32(2), b[:, 5]
    sx = np.array([1, 1, -1, -1, 1, 1, -1, -1], np.float32)
    sz = np.array([1, -1, -1, 1, 1, -1, -1, 1], np.float32)
    top = np.array([0, 0, 0, 0, 1, 1, 1, 1], np.float32)
    xc, zc, yc = hl[:, None] * sx[None], hw[:, None] * sz[None], -h[:, None] * top[None]
    c, s = np.cos(b[:, 6]).astype(np.float32)[:, None], np.sin(b[:, 6]).astype(np.float32)[:, None]
    x = c * xc + s * zc + b[:, 0:1]
    z = -s * xc + c * zc + b[:, 2:3]
    return np.stack([x, yc + b[:, 1:2], z], axis=1).astype(np.float32)


def cumsum(inp):
    inp = _c(inp, np.float32)
    b, n = inp.shape
    out = np.zeros((b, n), np.float32)
    lib().hf_oracle_cumsum(_i(b), _i(n), _ptr(inp), _ptr(out))
    return out


def prob_sample(inp, inpr):
    """sampling/tf_sampling.py:18-27: inp (b,ncategory) weights, inpr (b,npoints) uniform numbers -> (b,npoints) int32."""
    inp, inpr = _c(inp, np.float32), _c(inpr, np.float32)
    b, n = inp.shape
    m = inpr.shape[1]
    out = np.zeros((b, m), np.int32)
    lib().hf_oracle_prob_sample(_i(b), _i(n), _i(m), _ptr(inp), _ptr(inpr), _ptr(out))
    return out


def lidar_fov_filter(points, r0_rect, tr_velodyne_to_cam, p2, im_size):
    """CPU restatement of get_lidar_point_cloud with im_size (hf/core/obj_utils.py:242-272; lidar_to_cam_frame,
    hf/core/calib_utils.py:388-407; project_to_image, :289-296): raw (n,4) float32 [x,y,z,i] -> float64 (N,4)
    [x,y,z,i] in the rectified camera frame of the points with z > 0 that project into (0,w) x (0,h), scan order.
    Also returns the indices of the kept points."""
    points = np.asarray(points, dtype=np.float32)
    r0 = np.zeros((4, 4))
    r0[:3, :3] = np.asarray(r0_rect, dtype=np.float64)
    r0[3, 3] = 1.0
    tf = np.zeros((4, 4))
    tf[:3, :4] = np.asarray(tr_velodyne_to_cam, dtype=np.float64)
    tf[3, 3] = 1.0
    homog = np.concatenate([points[:, :3].astype(np.float64), np.ones((len(points), 1))], axis=1)
    cam = np.dot(np.dot(r0, tf), homog.T)[:3].T
    front = np.where(cam[:, 2] > 0)[0]
    cam = cam[front]
    uvw = np.dot(np.asarray(p2, dtype=np.float64), np.concatenate([cam.T, np.ones((1, len(cam)))], axis=0))
    with np.errstate(divide="ignore", invalid="ignore"):
        u, v = uvw[0] / uvw[2], uvw[1] / uvw[2]
    inside = (u > 0) & (u < im_size[0]) & (v > 0) & (v < im_size[1])
    kept = front[inside]
    return np.hstack([cam[inside], points[kept, 3:4].astype(np.float64)]), kept
