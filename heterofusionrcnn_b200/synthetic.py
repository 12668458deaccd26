"""Deterministic KITTI-shaped synthetic inputs for the operator path (numpy, host side).

Shapes and statistics follow the reference's data path (camera frame: x right, y down,
z forward; 16384 points per scan, hf/datasets/kitti/kitti_dataset.py:337-370) and its
input formats for the crop / IoU / NMS ops:
  * corner tensor (N,3,8), corner order P1..P8 of hf/core/box_8c_encoder.py:8-98;
  * BEV 5-tuple [x-l/2, z-w/2, x+l/2, z+w/2, ry] of hf/core/compute_iou.py:7-20.
Used by tests/ and bench.py; there are no datasets in this environment.
"""
import numpy as np


def make_scan(seed, n=16384, duplicate_pad=0):
    """One scan: xyz (n,3) float32, intensity (n,1) float32 in [-0.5, 0.5).

    Dense near the sensor, sparse far away, ground heavy, randomly permuted like the
    loader's shuffle.  duplicate_pad > 0 replaces the last `duplicate_pad` points with
    copies of random earlier ones, as the loader does for short scans
    (kitti_dataset.py:357-365) — this creates exact distance ties.
    """
    rng = np.random.default_rng(seed)
    u = rng.random(n)
    z = 4.0 + 66.0 * u * u
    x = np.clip(z * rng.uniform(-0.6, 0.6, n), -40.0, 40.0)
    y = np.clip(rng.normal(1.0, 0.6, n), -3.0, 3.0)
    xyz = np.stack([x, y, z], axis=1).astype(np.float32)
    inten = (rng.random((n, 1)) - 0.5).astype(np.float32)
    perm = rng.permutation(n)
    xyz, inten = xyz[perm], inten[perm]
    if duplicate_pad > 0:
        k = n - duplicate_pad
        src = rng.integers(0, k, duplicate_pad)
        xyz[k:] = xyz[src]
        inten[k:] = inten[src]
    return xyz, inten


def make_scans(seed, batch, n=16384, duplicate_pad=0):
    xs, its = zip(*[make_scan(seed * 1000 + i, n, duplicate_pad) for i in range(batch)])
    return np.stack(xs), np.stack(its)


def make_boxes3d(seed, xyz, num):
    """(num,7) [x,y,z,l,w,h,ry] car-sized boxes centred on scene points (y = bottom face)."""
    rng = np.random.default_rng(seed)
    c = xyz[rng.integers(0, xyz.shape[0], num)]
    l = rng.uniform(3.0, 5.0, num)
    w = rng.uniform(1.4, 2.0, num)
    h = rng.uniform(1.3, 1.9, num)
    ry = rng.uniform(-np.pi, np.pi, num)
    y = c[:, 1] + h * rng.uniform(0.2, 0.8, num)  # bottom below the seed point
    return np.stack([c[:, 0], y, c[:, 2], l, w, h, ry], axis=1).astype(np.float32)


def boxes3d_to_corners(boxes3d):
    """(N,7) -> (N,3,8): rows are x/y/z, columns the corners P1..P8.

    P1=(+l/2,0,+w/2) P2=(+l/2,0,-w/2) P3=(-l/2,0,-w/2) P4=(-l/2,0,+w/2), P5..P8 the same at
    y=-h; rotated about y by ry (x' = x cos + z sin, z' = -x sin + z cos) and translated.
    """
    b = np.asarray(boxes3d, np.float32)
    l, w, h, ry = b[:, 3], b[:, 4], b[:, 5], b[:, 6]
    sx = np.array([1, 1, -1, -1, 1, 1, -1, -1], np.float32) * 0.5
    sz = np.array([1, -1, -1, 1, 1, -1, -1, 1], np.float32) * 0.5
    top = np.array([0, 0, 0, 0, 1, 1, 1, 1], np.float32)
    xc = l[:, None] * sx[None]
    zc = w[:, None] * sz[None]
    yc = -h[:, None] * top[None]
    c, s = np.cos(ry)[:, None].astype(np.float32), np.sin(ry)[:, None].astype(np.float32)
    xr = xc * c + zc * s
    zr = -xc * s + zc * c
    out = np.stack([b[:, 0:1] + xr, b[:, 1:2] + yc, b[:, 2:3] + zr], axis=1)
    return out.astype(np.float32)


def boxes3d_to_bev(boxes3d):
    """(N,7) -> (N,5) [x - l/2, z - w/2, x + l/2, z + w/2, ry]."""
    b = np.asarray(boxes3d, np.float32)
    hl, hw = b[:, 3] / 2, b[:, 4] / 2
    return np.stack([b[:, 0] - hl, b[:, 2] - hw, b[:, 0] + hl, b[:, 2] + hw, b[:, 6]], axis=1).astype(np.float32)


def make_nms_boxes(seed, num=9000, clusters=40):
    """(num,5) BEV boxes, already sorted by descending score: jittered copies of `clusters`
    car-sized boxes (sigma 0.3 m, 0.1 rad), the shape of an RPN's 9000 pre-NMS proposals."""
    rng = np.random.default_rng(seed)
    cz = 4.0 + 66.0 * rng.random(clusters) ** 2
    cx = cz * rng.uniform(-0.5, 0.5, clusters)
    cl = rng.uniform(3.0, 5.0, clusters)
    cw = rng.uniform(1.4, 2.0, clusters)
    cr = rng.uniform(-np.pi, np.pi, clusters)
    k = rng.integers(0, clusters, num)
    x = cx[k] + rng.normal(0, 0.3, num)
    z = cz[k] + rng.normal(0, 0.3, num)
    l = cl[k] * rng.uniform(0.9, 1.1, num)
    w = cw[k] * rng.uniform(0.9, 1.1, num)
    ry = cr[k] + rng.normal(0, 0.1, num)
    boxes3d = np.stack([x, np.zeros(num), z, l, w, np.full(num, 1.5), ry], axis=1)
    scores = rng.random(num)
    order = np.argsort(-scores, kind="stable")
    return boxes3d_to_bev(boxes3d[order])


def fp_weights(dist):
    """PointNet++ FP weights from three_nn's squared distances
    (hf/core/feature_extractors/pointnet_util.py:304-307), float32 numpy."""
    d = np.maximum(dist.astype(np.float32), np.float32(1e-10))
    inv = np.float32(1.0) / d
    return (inv / inv.sum(axis=2, keepdims=True)).astype(np.float32)
