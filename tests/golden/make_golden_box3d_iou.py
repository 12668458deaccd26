#!/usr/bin/env python
"""Generates tests/golden/box3d_iou_ref.npz by IMPORTING THE REFERENCE'S OWN MODULE hf/core/compute_iou.py (run in the
build container, where /root/reference exists) with two stand-ins registered before the import, because the module's
two dependencies cannot be loaded here:
  * `tensorflow` -> a numpy shim with exactly the 9 names box3d_iou_tf / boxes3d_to_bev_tf use (stack, reshape, tile,
    shape, maximum, minimum, clip_by_value, float32.max): element-wise float32 numpy equivalents;
  * `bev_iou.bev_iou.compute_bev_iou` -> oracle/_ref/libref_bev_host.so, the device math of the reference's
    bev_iou/bev_iou_g.cu compiled unmodified for the host (oracle/Makefile).
Every line of compute_iou.py:7-64 — the BEV conversion, the height overlap, the volumes, the clipped union — is then the
reference's own code running on seeded boxes.  Stored: the boxes and the reference's (iou_3d, iou_2d).

    python tests/golden/make_golden_box3d_iou.py
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import binding as O  # noqa: E402

O.build()
assert O.ref_available("libref_bev_host.so"), "oracle/_ref/libref_bev_host.so missing: run python __graft_entry__.py"

tf = types.ModuleType("tensorflow")
tf.stack = lambda xs, axis=0: np.stack([np.asarray(x, np.float32) for x in xs], axis=axis)
tf.reshape = lambda x, shape: np.reshape(x, shape)
tf.tile = lambda x, reps: np.tile(x, reps)
tf.shape = lambda x: np.asarray(x).shape
tf.maximum = np.maximum
tf.minimum = np.minimum
tf.clip_by_value = lambda x, lo, hi: np.clip(x, np.float32(lo), np.float32(hi)).astype(np.float32)
tf.float32 = types.SimpleNamespace(max=np.finfo(np.float32).max)
sys.modules["tensorflow"] = tf

pkg = types.ModuleType("bev_iou")
mod = types.ModuleType("bev_iou.bev_iou")
mod.compute_bev_iou = lambda a, b: O.ref_bev_host(np.asarray(a, np.float32), np.asarray(b, np.float32))
pkg.bev_iou = mod
sys.modules["bev_iou"] = pkg
sys.modules["bev_iou.bev_iou"] = mod

sys.path.insert(0, "/root/reference")
from hf.core import compute_iou  # noqa: E402  (the reference's module, unmodified)

rng = np.random.default_rng(77)


def boxes(n):
    b = np.stack([rng.uniform(-6, 6, n), rng.uniform(-1, 2.5, n), rng.uniform(10, 22, n), rng.uniform(3, 5, n),
                  rng.uniform(1.4, 2.2, n), rng.uniform(1.3, 2.0, n), rng.uniform(-np.pi, np.pi, n)], axis=1).astype(np.float32)
    return b


a, b = boxes(48), boxes(40)
b[:5] = a[:5]                      # identical boxes: IoU 1
b[5, :] = a[5, :]; b[5, 1] += 0.5  # same footprint, shifted in height
b[6, :] = a[6, :]; b[6, 1] += 5.0  # same footprint, no height overlap
iou3d, iou2d = compute_iou.box3d_iou_tf(a, b)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "box3d_iou_ref.npz")
np.savez_compressed(out, boxes_a=a, boxes_b=b, iou_3d=np.asarray(iou3d, np.float32), iou_2d=np.asarray(iou2d, np.float32))
print("wrote", out, iou3d.shape, "pairs with overlap:", int((iou3d > 0).sum()), "max", float(iou3d.max()))
