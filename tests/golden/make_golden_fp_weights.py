#!/usr/bin/env python
"""Generates tests/golden/fp_weights_ref.npz by EXECUTING THE REFERENCE'S OWN LINES (run in the build container): the
inverse-distance weights of pointnet_fp_module, hf/core/feature_extractors/pointnet_util.py:304-307, cut out of the file
and exec'd with `tf` bound to a numpy float32 shim (maximum, reduce_sum, tile) on distances that include exact zeros
(the 1e-10 clamp), duplicates and KITTI-range values.

    python tests/golden/make_golden_fp_weights.py
"""
import os
import textwrap
import types

import numpy as np

src = open("/root/reference/hf/core/feature_extractors/pointnet_util.py").read().split("\n")
start = next(i for i, l in enumerate(src) if l.strip() == "dist = tf.maximum(dist, 1e-10)")
end = next(i for i in range(start, len(src)) if src[i].strip() == "weight = (1.0 / dist) / norm")
assert (start + 1, end + 1) == (304, 307), (start + 1, end + 1)
block = textwrap.dedent("\n".join(src[start:end + 1]))
tf = types.SimpleNamespace(
    maximum=lambda a, b: np.maximum(np.asarray(a, np.float32), np.float32(b)),
    reduce_sum=lambda x, axis=None, keep_dims=False: np.sum(np.asarray(x, np.float32), axis=axis, keepdims=keep_dims, dtype=np.float32),
    tile=lambda x, r: np.tile(x, r))
rng = np.random.default_rng(5)
dist = np.sort(rng.gamma(2.0, 0.5, (3, 2000, 3)).astype(np.float32) ** 2, axis=2)
dist[0, :50, 0] = 0.0           # the query is a known point
dist[0, 50:60, :2] = 0.0        # duplicated known points
dist[1, :20] = np.float32(1e-12)  # below the clamp
dist[2, :30] *= np.float32(400.0)
ns = {"tf": tf, "dist": dist.copy()}
exec(block, ns)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fp_weights_ref.npz")
np.savez_compressed(out, dist=dist, weight=np.asarray(ns["weight"], np.float32))
print("wrote", out, ns["weight"].shape, ns["weight"].dtype)
