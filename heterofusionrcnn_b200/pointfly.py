"""KNN of PointCNN — mirror of the operator part of hf/core/pointfly.py (knn_indices_general, :185-214), the one
function of that module that is a point-cloud search; the rest of pointfly.py is dense TF layers (out of scope).

Native code: csrc/spatial.cu (knn_kernel) via include/hfops.h.
"""
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


def knn_indices_general(queries, points, k, sort=True, unique=True):
    """Find indices of k-nearest neighbors given query points.

    Inputs:
      queries: (B, P, 3). Query points.
      points: (B, N, 3). The whole set of points.
      k: The number of k-nearest points.
      sort: the neighbours come out in ascending distance either way (tf.nn.top_k(sorted=False) leaves the order
            unspecified, so the sorted order is a valid result for both settings).
      unique: no effect, as in the reference (prepare_for_unique_top_k rebinds a local name, pointfly.py:162-164).
    Return:
      distances: (B, P, K) float32 squared distances.
      indices: (B, P, K, 2) int32 (batch index, point index) — what tf.gather_nd(points, indices) takes.

    Reference: hf/core/pointfly.py:185-214; caller hf/core/feature_extractors/pointcnn.py:72 (k = K * D, the dilated
    neighbourhood).  Values are direct squared distances (see grouping.tf_grouping.knn_point).
    """
    require(isinstance(k, int) and k > 0, "knn_indices_general expects positive k")
    queries = dev_tensor(queries, torch.float32, "queries")
    points = dev_tensor(points, torch.float32, "points")
    require(points.dim() == 3 and points.shape[2] == 3, "knn_indices_general expects (B, N, 3) points")
    require(queries.dim() == 3 and queries.shape[2] == 3 and queries.shape[0] == points.shape[0],
            "knn_indices_general expects (B, P, 3) queries")
    b, n = points.shape[0], points.shape[1]
    p = queries.shape[1]
    require(k <= n, "knn_indices_general: input must have at least k columns")
    dist = torch.empty((b, p, k), dtype=torch.float32, device=points.device)
    ind = torch.empty((b, p, k, 2), dtype=torch.int32, device=points.device)
    lib = _lib.load()
    ws_bytes = int(lib.hf_knn_workspace_bytes(b, n, p))
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=points.device) if ws_bytes else None
    with torch.cuda.device(points.device):
        rc = lib.hf_knn_indices_general(b, p, n, k, queries.data_ptr(), points.data_ptr(), dist.data_ptr(), ind.data_ptr(),
                                        ws.data_ptr() if ws is not None else None, ws_bytes, stream_ptr(points.device))
    _lib.check(rc, "knn_indices_general")
    return dist, ind
