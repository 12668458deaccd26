"""GPU: the compute_iou.py glue (csrc/glue.cu) through the Python mirror -> ctypes -> C-ABI against the oracle.
Bar: keep lists bit-exact; BEV boxes bit-exact; IoU values within 1e-5 relative (sinf/cosf differ by ulps
between CUDA and glibc, as for compute_bev_iou)."""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.compute_iou import box3d_iou_tf, boxes3d_to_bev_tf, oriented_nms_tf

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def N(t):
    return t.detach().cpu().numpy()


def clustered_boxes3d(seed, num, clusters=25):
    rng = np.random.default_rng(seed)
    xyz, _ = synthetic.make_scan(seed, 4096)
    base = synthetic.make_boxes3d(seed, xyz, clusters)
    b = base[rng.integers(0, clusters, num)].copy()
    b[:, [0, 2]] += rng.normal(0, 0.3, (num, 2)).astype(np.float32)
    b[:, 6] += rng.normal(0, 0.1, num).astype(np.float32)
    return b.astype(np.float32)


def test_boxes3d_to_bev_and_box3d_iou(cuda, oracle):
    a, b = clustered_boxes3d(1, 300), clustered_boxes3d(2, 40)
    b[:5] = a[:5]  # identical boxes: IoU 1
    assert np.array_equal(N(boxes3d_to_bev_tf(T(a, cuda))), oracle.boxes3d_to_bev(a))
    want3, want2 = oracle.box3d_iou(a, b)
    got3, got2 = box3d_iou_tf(T(a, cuda), T(b, cuda))
    assert np.allclose(N(got2), want2, rtol=1e-5, atol=1e-6)
    assert np.allclose(N(got3), want3, rtol=1e-5, atol=1e-6)
    assert np.allclose(np.diag(N(got3)[:5, :5]), 1.0, atol=1e-4) and (N(got3) > 0).sum() > 100


@pytest.mark.parametrize("s,n,thresh,ties", [(1, 700, 0.5, False), (3, 2000, 0.8, False), (2, 9000, 0.85, False),
                                             (2, 1500, 0.01, True), (1, 33, 0.3, True)])
def test_oriented_nms_tf_matches_oracle(cuda, oracle, s, n, thresh, ties):
    rng = np.random.default_rng(n)
    boxes = np.stack([clustered_boxes3d(10 + i, n) for i in range(s)])
    scores = rng.random((s, n)).astype(np.float32)
    if ties:
        scores = np.round(scores * 20) / np.float32(20)  # many equal scores: the lower index goes first
        scores[:, ::17] = -0.0
    keep, num = oriented_nms_tf(T(boxes, cuda), T(scores, cuda), thresh, return_num_kept=True)
    for i in range(s):
        want = oracle.oriented_nms_scored(boxes[i], scores[i], thresh)
        assert np.array_equal(N(keep[i]), want), i
    k0, n0 = N(keep[0]), int(N(num)[0])
    assert len(set(k0[:n0].tolist())) == n0 and (np.diff(scores[0][k0[:n0]]) <= 0).all()
    single = oriented_nms_tf(T(boxes[0], cuda), T(scores[0], cuda), thresh)
    assert torch.equal(single, keep[0])


@pytest.mark.parametrize("s,n,thresh,nms_size", [(2, 2000, 0.8, 300), (1, 700, 0.3, 512), (3, 1500, 0.01, 1024),
                                                 (2, 9000, 0.85, 512), (1, 100, 0.5, 300), (1, 64, 0.7, 0)])
def test_sb_nms_fn_matches_oracle(cuda, oracle, s, n, thresh, nms_size):
    """hf/core/models/model_util.py:101-142: slice to nms_size, tf.unique when the proposal count is not fixed, -1
    padding, count before padding; both settings, batched and single-scene calls."""
    from heterofusionrcnn_b200.model_util import sb_nms_fn

    rng = np.random.default_rng(n + nms_size)
    boxes = np.stack([clustered_boxes3d(20 + i, n) for i in range(s)])
    scores = rng.random((s, n)).astype(np.float32)
    for fixed in (True, False):
        got, num = sb_nms_fn((T(boxes, cuda), T(scores, cuda)), thresh, nms_size, fixed, None)
        assert tuple(got.shape) == (s, nms_size) and got.dtype == torch.int32
        for i in range(s):
            want, wnum = oracle.sb_nms_fn(boxes[i], scores[i], thresh, nms_size, fixed)
            assert np.array_equal(N(got[i]), want), (i, fixed)
            assert int(N(num)[i]) == wnum, (i, fixed)
        one, onum = sb_nms_fn((T(boxes[0], cuda), T(scores[0], cuda)), thresh, nms_size, fixed, None)
        assert torch.equal(one, got[0]) and int(onum) == int(num[0])


def test_sb_nms_fn_matches_the_reference_modules(cuda):
    """The CUDA path against tests/golden/sb_nms_ref.npz, the output of the reference's own model_util.sb_nms_fn /
    compute_iou.oriented_nms_tf (tests/golden/make_golden_sb_nms.py)."""
    import os

    from heterofusionrcnn_b200.model_util import sb_nms_fn

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sb_nms_ref.npz"))
    for c in sorted({k.split(".")[0] for k in g.files}):
        boxes, scores = T(g[c + ".boxes"], cuda), T(g[c + ".scores"], cuda)
        for fixed, tag in ((True, "fixed"), (False, "unique")):
            idx, num = sb_nms_fn((boxes, scores), float(g[c + ".thresh"]), int(g[c + ".nms_size"]), fixed, None)
            assert int(num) == int(g[c + ".num_" + tag]), (c, tag)
            assert np.array_equal(N(idx), g[c + ".idx_" + tag]), (c, tag)


def test_box_3d_to_box_8co_feeds_the_crop(cuda, oracle):
    """tf_box_3d_to_box_8co (+ the proposal expansion) against the oracle, and as the crop's box input."""
    from heterofusionrcnn_b200.box_8c_encoder import tf_box_3d_to_box_8co
    from heterofusionrcnn_b200.cropping.tf_cropping import pc_crop_and_sample

    xyz, inten = synthetic.make_scans(5, 2, 8192)
    b3 = np.concatenate([synthetic.make_boxes3d(70 + i, xyz[i], 20) for i in range(2)])
    for e in (0.0, 1.0):
        want = oracle.box_3d_to_box_8co(b3, e)
        got = tf_box_3d_to_box_8co(T(b3, cuda), e)
        assert got.shape == (40, 3, 8)
        assert np.allclose(N(got), want, rtol=1e-5, atol=2e-5)
    rng = np.random.default_rng(1)
    fts = rng.standard_normal((2, 8192, 8)).astype(np.float32)
    mask = rng.random((2, 8192)) < 0.3
    box_ind = np.repeat(np.arange(2, dtype=np.int32), 20)
    corners = tf_box_3d_to_box_8co(T(b3, cuda), 1.0)
    got = pc_crop_and_sample(T(xyz, cuda), T(fts, cuda), T(inten, cuda), T(mask, cuda), corners, T(box_ind, cuda), 64)
    want = oracle.pc_crop_and_sample(xyz, fts, inten, mask, N(corners), box_ind, 64)
    for g, w in zip(got, want):
        assert np.array_equal(N(g), w)
    assert N(got[5]).all()


@pytest.mark.parametrize("knn,use_xyz,c", [(False, True, 5), (False, True, 96), (True, True, 7), (False, False, 4), (False, True, 0)])
def test_sample_and_group_and_fp_interpolation(cuda, oracle, knn, use_xyz, c):
    """pointnet_util.sample_and_group / interpolate_points against the same composition of oracle ops."""
    from heterofusionrcnn_b200 import pointnet_util

    rng = np.random.default_rng(c + 3)
    b, n, m, ns, radius = 2, 4096, 512, 16, 1.0
    xyz, _ = synthetic.make_scans(23, b, n)
    pts = rng.standard_normal((b, n, c)).astype(np.float32) if c else None
    new_xyz, new_points, idx, grouped_xyz = pointnet_util.sample_and_group(
        m, radius, ns, T(xyz, cuda), T(pts, cuda) if c else None, knn=knn, use_xyz=use_xyz)
    w_new = oracle.gather_point(xyz, oracle.farthest_point_sample(m, xyz))
    w_idx = oracle.knn_point(ns, xyz, w_new)[1] if knn else oracle.query_ball_point(radius, ns, xyz, w_new)[0]
    w_gxyz = oracle.group_point(xyz, w_idx) - w_new[:, :, None, :]
    assert np.array_equal(N(new_xyz), w_new) and np.array_equal(N(idx), w_idx)
    assert np.array_equal(N(grouped_xyz), w_gxyz)
    if c == 0:
        want = w_gxyz
    elif use_xyz:
        want = np.concatenate([w_gxyz, oracle.group_point(pts, w_idx)], axis=-1)
    else:
        want = oracle.group_point(pts, w_idx)
    assert N(new_points).shape == want.shape and np.array_equal(N(new_points), want)
    if c:
        feats = rng.standard_normal((b, m, c)).astype(np.float32)
        got = pointnet_util.interpolate_points(T(xyz, cuda), new_xyz, T(feats, cuda))
        wd, wi = oracle.three_nn(xyz, w_new)
        want = oracle.three_interpolate(feats, wi, synthetic.fp_weights(wd))
        assert np.allclose(N(got), want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,m,radii,nsamples,c,use_xyz", [(16384, 4096, [0.1, 0.5], [16, 32], 1, True),
                                                          (4096, 1024, [0.5, 1.0], [16, 32], 96, True),
                                                          (1024, 256, [1.0, 2.0], [16, 32], 32, False),
                                                          (512, 64, [2.0, 4.0], [16, 32], 0, True)])
def test_sample_and_group_msg_matches_the_reference_composition(cuda, oracle, n, m, radii, nsamples, c, use_xyz):
    """pointnet_util.sample_and_group_msg = the operator part of pointnet_sa_module_msg (reference pointnet_util.py:
    223-286) at the radii / nsample of rpn_cars_pointnet_paper.config:65-111, against the same composition of oracle
    ops: one FPS, then per scale ball query, two group_points, seed subtraction and the [points, xyz] concat.  Bit-exact."""
    from heterofusionrcnn_b200 import pointnet_util

    rng = np.random.default_rng(n + c)
    b = 2
    xyz, _ = synthetic.make_scans(31, b, n)
    pts = rng.standard_normal((b, n, c)).astype(np.float32) if c else None
    new_xyz, grouped = pointnet_util.sample_and_group_msg(m, radii, nsamples, T(xyz, cuda), T(pts, cuda) if c else None,
                                                          use_xyz=use_xyz)
    w_new = oracle.gather_point(xyz, oracle.farthest_point_sample(m, xyz))
    assert np.array_equal(N(new_xyz), w_new) and len(grouped) == len(radii)
    for r, ns, got in zip(radii, nsamples, grouped):
        w_idx, _ = oracle.query_ball_point(r, ns, xyz, w_new)
        w_gxyz = oracle.group_point(xyz, w_idx) - w_new[:, :, None, :]
        if c == 0:
            want = w_gxyz
        elif use_xyz:
            want = np.concatenate([oracle.group_point(pts, w_idx), w_gxyz], axis=-1)  # features first (:262-264)
        else:
            want = oracle.group_point(pts, w_idx)
        assert N(got).shape == want.shape and np.array_equal(N(got), want), r
