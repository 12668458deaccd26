#!/usr/bin/env python
"""Generates tests/golden/knn_pointfly_ref.npz by IMPORTING THE REFERENCE'S OWN MODULES (run in the build container, where
/root/reference exists): hf/core/pointfly.py:knn_indices_general(queries, points, k, sort=True, unique=False)
(:185-214 -> batch_distance_matrix_general :135-147) and grouping/tf_grouping.py's knn_point formula (:82-93, cut out of
the file and executed, because the module loads a TensorFlow op library at import), with `tensorflow` replaced by a numpy
float32 shim for the calls they make (reduce_sum, matmul, transpose, nn.top_k, shape, range, reshape, tile, concat,
expand_dims); transforms3d (unused by these functions) gets an empty stand-in.  unique=True is not run: in a TF graph its
py_func is never executed (prepare_for_unique_top_k rebinds a local, :162-164), eagerly it would call the non-existent
np.fill.  Stored: the clouds and both functions' (distances, indices).

    python tests/golden/make_golden_knn_pointfly.py
"""
import os
import sys
import textwrap
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from heterofusionrcnn_b200 import synthetic  # noqa: E402


def top_k(x, k, sorted=True):  # noqa: A002
    order = np.argsort(-np.asarray(x, np.float32), axis=-1, kind="stable")[..., :k]
    return np.take_along_axis(np.asarray(x, np.float32), order, axis=-1), order.astype(np.int32)


tf = types.ModuleType("tensorflow")
tf.reduce_sum = lambda x, axis=None, keep_dims=False: np.sum(np.asarray(x, np.float32), axis=axis, keepdims=keep_dims, dtype=np.float32)
tf.matmul = lambda a, b: np.matmul(np.asarray(a, np.float32), np.asarray(b, np.float32)).astype(np.float32)
tf.transpose = lambda x, perm=None: np.transpose(x, perm)
tf.shape = lambda x: np.asarray(x).shape
tf.range = lambda n: np.arange(n, dtype=np.int32)
tf.reshape = lambda x, s: np.reshape(x, s)
tf.tile = lambda x, r: np.tile(x, r)
tf.concat = lambda xs, axis: np.concatenate(xs, axis=axis)
tf.expand_dims = lambda x, axis: np.expand_dims(x, axis)


class _Lazy(types.SimpleNamespace):
    """tf.nn / tf.layers ...: names the module only mentions in default arguments of functions that are not called."""

    def __getattr__(self, name):
        return None


tf.nn = _Lazy(top_k=top_k)
tf.layers = _Lazy()
tf.contrib = _Lazy()
sys.modules["tensorflow"] = tf
t3 = types.ModuleType("transforms3d")
t3e = types.ModuleType("transforms3d.euler")
t3e.euler2mat = None
sys.modules["transforms3d"], sys.modules["transforms3d.euler"] = t3, t3e
sys.path.insert(0, "/root/reference")
from hf.core import pointfly  # noqa: E402  (the reference's module, unmodified)

# knn_point's formula lines of grouping/tf_grouping.py, executed as they stand
src = open("/root/reference/grouping/tf_grouping.py").read().split("\n")
start = next(i for i, l in enumerate(src) if l.strip().startswith("r_xyz1 = tf.reduce_sum"))
end = next(i for i in range(start, len(src)) if src[i].strip().startswith("val, idx = tf.nn.top_k"))
assert (start + 1, end + 1) == (82, 93), (start + 1, end + 1)
block = textwrap.dedent("\n".join(l for l in src[start:end + 1] if not l.strip().startswith("#")))

xyz, _ = synthetic.make_scans(123, 2, 1500)
points = xyz
queries = xyz[:, ::9][:, :120]
k = 24
dist_pf, ind_pf = pointfly.knn_indices_general(queries, points, k, True, False)
ns = {"tf": tf, "xyz1": points, "xyz2": queries, "k": k}
exec(block, ns)
val_kp, idx_kp = -ns["val"], ns["idx"]
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "knn_pointfly_ref.npz")
np.savez_compressed(out, points=points, queries=queries, k=np.int32(k), pointfly_dist=np.asarray(dist_pf, np.float32),
                    pointfly_ind=np.asarray(ind_pf, np.int32), knn_point_val=np.asarray(val_kp, np.float32),
                    knn_point_idx=np.asarray(idx_kp, np.int32))
print("wrote", out, dist_pf.shape, ind_pf.shape, val_kp.shape)
