#!/usr/bin/env python
"""Generates tests/golden/kitti_000000_16384.npz with the REFERENCE's own input code (run in the build container,
where /root/reference exists): fixture scan 000000 of hf/tests/datasets -> hf/core/obj_utils.get_lidar_point_cloud
(lidar -> camera frame, in-front-of-camera and image-FOV filter) -> the 16384-point sampling of
hf/datasets/kitti/kitti_dataset.py:341-370 (near points sampled, far points kept, shuffled, intensity - 0.5),
restated here with a seeded numpy RNG because the dataset class needs TensorFlow.  This is BASELINE.json
configs[0]'s input ("one hf/tests/datasets KITTI velodyne scan, 16384 -> 4096 pts").  An empty stand-in module is
registered for the unused `import tensorflow`.

    python tests/golden/make_golden_kitti.py
"""
import os
import struct
import sys
import types

import numpy as np

sys.modules.setdefault("tensorflow", types.ModuleType("tensorflow"))
sys.path.insert(0, "/root/reference")
from hf.core import obj_utils  # noqa: E402

root = "/root/reference/hf/tests/datasets/Kitti/object/training"
with open(root + "/image_2/000000.png", "rb") as f:
    w, h = struct.unpack(">II", f.read(32)[16:24])
pc = obj_utils.get_lidar_point_cloud(0, root + "/calib", root + "/velodyne", im_size=[w, h])  # (N,4) float64
pts_rect, pts_intensity = pc[:, :3], pc[:, 3]
n = 16384
rng = np.random.RandomState(0)
assert n < len(pts_rect)
near = pts_rect[:, 2] < 40.0  # kitti_dataset.py:342-356
far_choice = np.where(~near)[0]
near_choice = rng.choice(np.where(near)[0], n - len(far_choice), replace=False)
choice = np.concatenate((near_choice, far_choice)) if len(far_choice) else near_choice
rng.shuffle(choice)
xyz = pts_rect[choice].astype(np.float32)
intensity = (pts_intensity[choice] - 0.5).astype(np.float32)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kitti_000000_16384.npz")
np.savez_compressed(out, xyz=xyz, intensity=intensity, image_size=np.array([w, h]), in_fov=np.array(len(pts_rect)))
print("wrote", out, xyz.shape, os.path.getsize(out), "bytes;", len(pts_rect), "points in the image FOV")
