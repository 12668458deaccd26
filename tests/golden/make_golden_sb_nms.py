#!/usr/bin/env python
"""Generates tests/golden/sb_nms_ref.npz by IMPORTING THE REFERENCE'S OWN MODULES hf/core/models/model_util.py and
hf/core/compute_iou.py (run in the build container, where /root/reference exists) and calling sb_nms_fn
(model_util.py:101-142) -> oriented_nms_tf (compute_iou.py:67-80) on seeded proposals.  Stand-ins registered before the
import, for the two dependencies that cannot be loaded here:
  * `tensorflow` -> a numpy shim with the names those functions use (stack, shape, minimum, greater, unique, cond, pad,
    gather, nn.top_k as a stable descending sort — TF's documented tie rule: the lower index first);
  * `bev_iou.bev_iou.oriented_nms` -> the suppression mask from the reference's own iou_bev (bev_iou/bev_iou_g.cu
    compiled unmodified for the host, oracle/_ref/libref_bev_host.so; `iou > thresh` as in :317-338) swept by the
    restatement of bev_iou/bev_iou.cpp:87-112 (that file needs TensorFlow headers and cannot be compiled).
The slice, the tf.unique, the -1 padding and the count are the reference's own lines.

    python tests/golden/make_golden_sb_nms.py
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import binding as O  # noqa: E402

O.build()
assert O.ref_available("libref_bev_host.so")


def top_k(x, k):
    order = np.argsort(-np.asarray(x, np.float32), kind="stable")[:k]
    return np.asarray(x)[order], order.astype(np.int32)


def unique(x):
    x = np.asarray(x)
    _, first = np.unique(x, return_index=True)
    return x[np.sort(first)], None


tf = types.ModuleType("tensorflow")
tf.stack = lambda xs, axis=0: np.stack([np.asarray(v, np.float32) for v in xs], axis=axis)
tf.shape = lambda x: np.asarray(x).shape
tf.minimum = lambda a, b: min(int(a), int(b))
tf.greater = lambda a, b: a > b
tf.unique = unique
tf.cond = lambda c, true_fn, false_fn: true_fn() if c else false_fn()
tf.pad = lambda x, pads, mode="CONSTANT", constant_values=0: np.pad(np.asarray(x), pads, mode="constant",
                                                                    constant_values=constant_values)
tf.gather = lambda x, idx: np.asarray(x)[np.asarray(idx)]
tf.nn = types.SimpleNamespace(top_k=top_k)
tf.float32 = types.SimpleNamespace(max=np.finfo(np.float32).max)
sys.modules["tensorflow"] = tf


def ref_oriented_nms(boxes_bev, thresh):
    b = np.asarray(boxes_bev, np.float32)
    n = len(b)
    _, iou = O.ref_bev_host(b, b)
    words = (n + 63) // 64
    mask = np.zeros((n, words), np.uint64)
    for i in range(n):
        for j in np.nonzero(iou[i] > np.float32(thresh))[0]:
            mask[i, j // 64] |= np.uint64(1) << np.uint64(j % 64)
    keep, _ = O.nms_sweep(mask)
    return keep


pkg = types.ModuleType("bev_iou")
mod = types.ModuleType("bev_iou.bev_iou")
mod.oriented_nms = ref_oriented_nms
mod.compute_bev_iou = O.ref_bev_host
pkg.bev_iou = mod
sys.modules["bev_iou"] = pkg
sys.modules["bev_iou.bev_iou"] = mod
sys.path.insert(0, "/root/reference")
from hf.core.models import model_util  # noqa: E402  (the reference's module, unmodified)

from heterofusionrcnn_b200 import synthetic  # noqa: E402


def proposals(seed, num, clusters):
    rng = np.random.default_rng(seed)
    xyz, _ = synthetic.make_scan(seed, 4096)
    base = synthetic.make_boxes3d(seed, xyz, clusters)
    b = base[rng.integers(0, clusters, num)].copy()
    b[:, [0, 2]] += rng.normal(0, 0.3, (num, 2)).astype(np.float32)
    b[:, 6] += rng.normal(0, 0.1, num).astype(np.float32)
    return b.astype(np.float32), rng.random(num).astype(np.float32)


out = {}
for ci, (n, clusters, thresh, nms_size, ties) in enumerate([(300, 12, 0.5, 64, False), (300, 12, 0.5, 400, False),
                                                             (200, 5, 0.1, 32, True), (150, 40, 0.8, 256, False),
                                                             (64, 4, 0.3, 64, True)]):
    boxes, scores = proposals(100 + ci, n, clusters)
    if ties:
        scores = (np.round(scores * 10) / np.float32(10)).astype(np.float32)
    out["c%d.boxes" % ci], out["c%d.scores" % ci] = boxes, scores
    out["c%d.thresh" % ci], out["c%d.nms_size" % ci] = np.float32(thresh), np.int32(nms_size)
    for fixed in (True, False):
        idx, num = model_util.sb_nms_fn((boxes, scores), thresh, nms_size, fixed, None)
        out["c%d.idx_%s" % (ci, "fixed" if fixed else "unique")] = np.asarray(idx, np.int32)
        out["c%d.num_%s" % (ci, "fixed" if fixed else "unique")] = np.int32(num)
        print("case %d n=%d thresh=%.2f nms_size=%d fixed=%s -> %d before padding" % (ci, n, thresh, nms_size, fixed, num))
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sb_nms_ref.npz")
np.savez_compressed(path, **out)
print("wrote", path)
