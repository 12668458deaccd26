#!/usr/bin/env python
"""Generates tests/golden/box8co_ref.npz from the REFERENCE's own numpy code (run in the build container, where
/root/reference exists): hf/core/box_8c_encoder.py:np_box_3d_to_box_8co on seeded boxes.  The module imports
tensorflow at the top (unused by the numpy function), which is not installed: an empty stand-in module is
registered so that the import statement succeeds; nothing of TensorFlow is called.

    python tests/golden/make_golden_box8co.py
"""
import os
import sys
import types

import numpy as np

sys.modules.setdefault("tensorflow", types.ModuleType("tensorflow"))
sys.path.insert(0, "/root/reference")
from hf.core import box_8c_encoder  # noqa: E402

rng = np.random.default_rng(2024)
n = 64
boxes = np.stack([rng.uniform(-30, 30, n), rng.uniform(-2, 3, n), rng.uniform(4, 70, n), rng.uniform(3, 5, n),
                  rng.uniform(1.4, 2, n), rng.uniform(1.3, 1.9, n), rng.uniform(-np.pi, np.pi, n)], axis=1).astype(np.float32)
boxes[0, 6] = 0.0
boxes[1, 6] = np.float32(np.pi / 2)
corners = box_8c_encoder.np_box_3d_to_box_8co(boxes)  # (N, 8, 3)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "box8co_ref.npz")
np.savez_compressed(out, boxes=boxes, corners_n83=corners)
print("wrote", out, corners.shape)
