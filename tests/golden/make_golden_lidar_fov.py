#!/usr/bin/env python
"""Generates tests/golden/lidar_fov_000000.npz with the REFERENCE's own input code (run in the build container,
where /root/reference exists): the first 49152 records of fixture scan 000000 of hf/tests/datasets are written to a
scratch velodyne directory and passed through hf/core/obj_utils.get_lidar_point_cloud(im_size=[w, h]) (lidar ->
rectified camera frame, z > 0, image field of view).  Stored: the raw records, the three calibration matrices the
reference parsed, the image size and the reference's float64 (N,4) result.  An empty stand-in module is registered
for the unused `import tensorflow`.

    python tests/golden/make_golden_lidar_fov.py
"""
import os
import struct
import sys
import tempfile
import types

import numpy as np

sys.modules.setdefault("tensorflow", types.ModuleType("tensorflow"))
sys.path.insert(0, "/root/reference")
from hf.core import calib_utils, obj_utils  # noqa: E402

root = "/root/reference/hf/tests/datasets/Kitti/object/training"
with open(root + "/image_2/000000.png", "rb") as f:
    w, h = struct.unpack(">II", f.read(32)[16:24])
raw = np.fromfile(root + "/velodyne/000000.bin", dtype=np.float32).reshape(-1, 4)[:49152]
with tempfile.TemporaryDirectory() as tmp:
    raw.tofile(tmp + "/000000.bin")
    expect = obj_utils.get_lidar_point_cloud(0, root + "/calib", tmp, im_size=[w, h])
calib = calib_utils.read_calibration(root + "/calib", 0)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lidar_fov_000000.npz")
np.savez_compressed(
    out, raw=raw, p2=np.asarray(calib.p2, np.float64), r0_rect=np.asarray(calib.r0_rect, np.float64),
    tr_velodyne_to_cam=np.asarray(calib.tr_velodyne_to_cam, np.float64), image_size=np.array([w, h]), expect=expect,
)
print("wrote", out, raw.shape, "->", expect.shape, expect.dtype, os.path.getsize(out), "bytes")
