"""Ball query / group_point — mirror of the reference's grouping/tf_grouping.py.

Native code: csrc/grouping.cu via include/hfops.h.
"""
import torch

from .. import _lib
from .._tensors import dev_tensor, require, scatter_workspace, stream_ptr


def query_ball_point(radius, nsample, xyz1, xyz2, index1=None, index2=None):
    """
    Input:
        radius: float32, ball search radius
        nsample: int32, number of points selected in each ball region
        xyz1: (batch_size, ndataset, 3) float32 array, input points
        xyz2: (batch_size, npoint, 3) float32 array, query points
    Output:
        idx: (batch_size, npoint, nsample) int32 array, indices to input points
        pts_cnt: (batch_size, npoint) int32 array, number of unique points in each local region

    Reference: grouping/tf_grouping.py:11-23, op checks grouping/tf_grouping.cpp:67-106.
    Rows with no point in the ball are returned as 0 (uninitialised in the reference).
    index1 / index2 (optional, not in the reference): spatial.CloudIndex of xyz1 / xyz2, to share
    one sort between several ops on the same clouds; results do not depend on them.
    """
    require(float(radius) > 0, "QueryBallPoint expects positive radius")
    require(int(nsample) > 0, "QueryBallPoint expects positive nsample")
    xyz1 = dev_tensor(xyz1, torch.float32, "xyz1")
    xyz2 = dev_tensor(xyz2, torch.float32, "xyz2")
    require(xyz1.dim() == 3 and xyz1.shape[2] == 3, "QueryBallPoint expects (batch_size, ndataset, 3) xyz1 shape.")
    require(xyz2.dim() == 3 and xyz2.shape[2] == 3, "QueryBallPoint expects (batch_size, npoint, 3) xyz2 shape.")
    require(xyz2.shape[0] == xyz1.shape[0], "QueryBallPoint expects xyz1 and xyz2 to share batch_size")
    b, n, m = xyz1.shape[0], xyz1.shape[1], xyz2.shape[1]
    idx = torch.empty((b, m, int(nsample)), dtype=torch.int32, device=xyz1.device)
    pts_cnt = torch.empty((b, m), dtype=torch.int32, device=xyz1.device)
    lib = _lib.load()
    p1 = index1.ptr_for(xyz1) if index1 is not None else None
    p2 = index2.ptr_for(xyz2) if index2 is not None else None
    with torch.cuda.device(xyz1.device):
        if p1 is not None and p2 is not None:
            rc = lib.hf_query_ball_point_indexed(
                b, n, m, float(radius), int(nsample), xyz1.data_ptr(), xyz2.data_ptr(), idx.data_ptr(),
                pts_cnt.data_ptr(), p1, p2, stream_ptr(xyz1.device))
        else:
            # scratch for the spatially pruned kernels (csrc/spatial.cu); 0 bytes = every-pair kernel
            ws_bytes = int(lib.hf_spatial_workspace_bytes(b, n, m)) if lib.hf_spatial_applies(m, n) else 0
            ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=xyz1.device) if ws_bytes else None
            rc = lib.hf_query_ball_point_ws(
                b, n, m, float(radius), int(nsample), xyz1.data_ptr(), xyz2.data_ptr(), idx.data_ptr(),
                pts_cnt.data_ptr(), ws.data_ptr() if ws is not None else None, ws_bytes, stream_ptr(xyz1.device))
    _lib.check(rc, "query_ball_point")
    return idx, pts_cnt


def group_point(points, idx):
    """
    Input:
        points: (batch_size, ndataset, channel) float32 array, points to sample from
        idx: (batch_size, npoint, nsample) int32 array, indices to points
    Output:
        out: (batch_size, npoint, nsample, channel) float32 array, values sampled from points

    Reference: grouping/tf_grouping.py:44-52, op checks grouping/tf_grouping.cpp:143-171.
    """
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    require(points.dim() == 3, "GroupPoint expects (batch_size, num_points, channel) points shape")
    b, n, c = points.shape
    require(idx.dim() == 3 and idx.shape[0] == b, "GroupPoint expects (batch_size, npoints, nsample) idx shape")
    m, nsample = idx.shape[1], idx.shape[2]
    out = torch.empty((b, m, nsample, c), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        rc = _lib.load().hf_group_point(b, n, c, m, nsample, points.data_ptr(), idx.data_ptr(), out.data_ptr(),
                                        stream_ptr(points.device))
    _lib.check(rc, "group_point")
    return out


def group_point_pair(xyz, points, idx):
    """The two group_point calls of sample_and_group (hf/core/feature_extractors/pointnet_util.py:49,56) in one pass:

        grouped_xyz, grouped_points = group_point(xyz, idx), group_point(points, idx)

    xyz (b,n,3), points (b,n,c), idx (b,m,nsample) int32 -> (b,m,nsample,3), (b,m,nsample,c); identical results.
    """
    xyz = dev_tensor(xyz, torch.float32, "xyz")
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    require(xyz.dim() == 3 and xyz.shape[2] == 3, "GroupPoint expects (batch_size, num_points, 3) xyz shape")
    require(points.dim() == 3 and points.shape[0] == xyz.shape[0] and points.shape[1] == xyz.shape[1],
            "GroupPoint expects (batch_size, num_points, channel) points shape")
    b, n, c = points.shape
    require(idx.dim() == 3 and idx.shape[0] == b, "GroupPoint expects (batch_size, npoints, nsample) idx shape")
    m, nsample = idx.shape[1], idx.shape[2]
    out_xyz = torch.empty((b, m, nsample, 3), dtype=torch.float32, device=points.device)
    out = torch.empty((b, m, nsample, c), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        rc = _lib.load().hf_group_point_pair(b, n, c, m, nsample, xyz.data_ptr(), points.data_ptr(), idx.data_ptr(),
                                             out_xyz.data_ptr(), out.data_ptr(), stream_ptr(points.device))
    _lib.check(rc, "group_point")
    return out_xyz, out


def group_point_grad(points, idx, grad_out):
    """Gradient of group_point w.r.t. points (the reference's GroupPointGrad op,
    grouping/tf_grouping.py:55-59, op checks grouping/tf_grouping.cpp:173-208).

    points (b,n,c) [shape only], idx (b,m,nsample) int32, grad_out (b,m,nsample,c) -> grad_points (b,n,c).
    Deterministic: a point's contributions are added in ascending (row, sample) order.
    """
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    grad_out = dev_tensor(grad_out, torch.float32, "grad_out")
    require(points.dim() == 3, "GroupPointGrad expects (batch_size, num_points, channel) points shape")
    b, n, c = points.shape
    require(idx.dim() == 3 and idx.shape[0] == b, "GroupPointGrad expects (batch_size, npoints, nsample) idx shape")
    m, nsample = idx.shape[1], idx.shape[2]
    require(grad_out.dim() == 4 and tuple(grad_out.shape) == (b, m, nsample, c),
            "GroupPointGrad expects (batch_size, npoints, nsample, channel) grad_out shape")
    grad_points = torch.empty((b, n, c), dtype=torch.float32, device=points.device)
    lib = _lib.load()
    ws, ws_bytes = scatter_workspace(lib, b, n, m * nsample, points.device)
    with torch.cuda.device(points.device):
        rc = lib.hf_group_point_grad(b, n, c, m, nsample, grad_out.data_ptr(), idx.data_ptr(), grad_points.data_ptr(),
                                     ws.data_ptr(), ws_bytes, stream_ptr(points.device))
    _lib.check(rc, "group_point_grad")
    return grad_points


def select_top_k(k, dist):
    """
    Input:
        k: int32, number of k SMALLEST elements selected
        dist: (b,m,n) float32 array, distance matrix, m query points, n dataset points
    Output:
        idx: (b,m,n) int32 array, first k in n are indices to the top k
        dist_out: (b,m,n) float32 array, first k in n are the top k

    Reference: grouping/tf_grouping.py:29-39 (SelectionSort op, grouping/tf_grouping.cpp:108-141);
    the whole output, tail included, equals the reference kernel's (tf_grouping_g.cu:83-123).
    """
    require(isinstance(k, int) and k > 0, "SelectionSort expects positive k")
    dist = dev_tensor(dist, torch.float32, "dist")
    require(dist.dim() == 3, "SelectionSort expects (b,m,n) dist shape.")
    b, m, n = dist.shape
    require(n >= 1, "SelectionSort expects (b,m,n) dist shape.")
    idx = torch.empty((b, m, n), dtype=torch.int32, device=dist.device)
    dist_out = torch.empty((b, m, n), dtype=torch.float32, device=dist.device)
    with torch.cuda.device(dist.device):
        rc = _lib.load().hf_select_top_k(b, n, m, k, dist.data_ptr(), idx.data_ptr(), dist_out.data_ptr(),
                                         stream_ptr(dist.device))
    _lib.check(rc, "select_top_k")
    return idx, dist_out


def knn_point(k, xyz1, xyz2):
    """
    Input:
        k: int32, number of k in k-nn search
        xyz1: (batch_size, ndataset, c) float32 array, input points
        xyz2: (batch_size, npoint, c) float32 array, query points
    Output:
        val: (batch_size, npoint, k) float32 array, L2 distances
        idx: (batch_size, npoint, k) int32 array, indices to input points

    Reference: grouping/tf_grouping.py:62-95.  As there, val holds SQUARED distances, ascending, and
    equal distances keep the lower index first (tf.nn.top_k).  c must be 3.  The (b,m,n) distance
    matrix of the reference is never formed, and val is the directly computed squared distance rather
    than r2 - 2*x2.x1 + r1 (which carries a cancellation error of ~1e-7 * |x|^2).
    """
    require(isinstance(k, int) and k > 0, "KnnPoint expects positive k")
    xyz1 = dev_tensor(xyz1, torch.float32, "xyz1")
    xyz2 = dev_tensor(xyz2, torch.float32, "xyz2")
    require(xyz1.dim() == 3 and xyz1.shape[2] == 3, "KnnPoint expects (batch_size, ndataset, 3) xyz1 shape")
    require(xyz2.dim() == 3 and xyz2.shape[2] == 3 and xyz2.shape[0] == xyz1.shape[0],
            "KnnPoint expects (batch_size, npoint, 3) xyz2 shape")
    b, n = xyz1.shape[0], xyz1.shape[1]
    m = xyz2.shape[1]
    require(k <= n, "KnnPoint: input must have at least k columns")
    val = torch.empty((b, m, k), dtype=torch.float32, device=xyz1.device)
    idx = torch.empty((b, m, k), dtype=torch.int32, device=xyz1.device)
    lib = _lib.load()
    ws_bytes = int(lib.hf_knn_workspace_bytes(b, n, m))
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=xyz1.device) if ws_bytes else None
    with torch.cuda.device(xyz1.device):
        rc = lib.hf_knn_point(b, n, m, k, xyz1.data_ptr(), xyz2.data_ptr(), val.data_ptr(), idx.data_ptr(),
                              ws.data_ptr() if ws is not None else None, ws_bytes, stream_ptr(xyz1.device))
    _lib.check(rc, "knn_point")
    return val, idx


def expanded_sqdist(xyz1, xyz2):
    """The (b, m, n) matrix knn_point_old sorts: r_xyz2 - 2 * xyz2 . xyz1^T + r_xyz1^T (grouping/tf_grouping.py:
    117-121), float32.  The batched product is a plain library GEMM, as in the reference (tf.matmul)."""
    r1 = (xyz1 * xyz1).sum(dim=2, keepdim=True)  # (b, n, 1)
    r2 = (xyz2 * xyz2).sum(dim=2, keepdim=True)  # (b, m, 1)
    mul = torch.bmm(xyz2, xyz1.transpose(1, 2))  # (b, m, n)
    return (r2 - 2 * mul + r1.transpose(1, 2)).contiguous()


def knn_point_old(k, xyz1, xyz2):
    """
    Input:
        k: int32, number of k in k-nn search
        xyz1: (batch_size, ndataset, c) float32 array, input points
        xyz2: (batch_size, npoint, c) float32 array, query points
    Output:
        val: (batch_size, npoint, k) float32 array, L2 distances
        idx: (batch_size, npoint, k) int32 array, indices to input points

    Reference: grouping/tf_grouping.py:98-131 - the expanded-form distance matrix, select_top_k (the SelectionSort
    op) and a slice of its first k columns.  Kept for completeness of the module's public names (the reference uses
    it only in its __main__ demo); knn_point is the fused form that never builds the matrix.  val carries the
    cancellation error of the expanded form, exactly like the reference.  Any c.
    """
    require(isinstance(k, int) and k > 0, "SelectionSort expects positive k")
    xyz1 = dev_tensor(xyz1, torch.float32, "xyz1")
    xyz2 = dev_tensor(xyz2, torch.float32, "xyz2")
    require(xyz1.dim() == 3 and xyz2.dim() == 3 and xyz1.shape[0] == xyz2.shape[0] and xyz1.shape[2] == xyz2.shape[2],
            "knn_point_old expects (batch_size, ndataset, c) xyz1 and (batch_size, npoint, c) xyz2")
    require(k <= xyz1.shape[1], "knn_point_old: input must have at least k columns")
    outi, out = select_top_k(k, expanded_sqdist(xyz1, xyz2))
    return out[:, :, :k].contiguous(), outi[:, :, :k].contiguous()


def group_point_concat(xyz, points, idx, new_xyz, xyz_last=False):
    """The tail of sample_and_group (hf/core/feature_extractors/pointnet_util.py:49-60) as one kernel:

        grouped_xyz = group_point(xyz, idx) - new_xyz[:, :, None, :]
        new_points  = concat([grouped_xyz, group_point(points, idx)], -1)     (points None: grouped_xyz)

    xyz (b,n,3), points (b,n,c) or None, idx (b,m,nsample) int32, new_xyz (b,m,3) -> (b,m,nsample,3+c).
    xyz_last=True gives the column order of pointnet_sa_module_msg (:262-264),
    concat([group_point(points, idx), grouped_xyz], -1).
    """
    xyz = dev_tensor(xyz, torch.float32, "xyz")
    idx = dev_tensor(idx, torch.int32, "idx")
    new_xyz = dev_tensor(new_xyz, torch.float32, "new_xyz")
    require(xyz.dim() == 3 and xyz.shape[2] == 3, "GroupPoint expects (batch_size, num_points, 3) xyz shape")
    b, n = xyz.shape[0], xyz.shape[1]
    require(idx.dim() == 3 and idx.shape[0] == b, "GroupPoint expects (batch_size, npoints, nsample) idx shape")
    m, nsample = idx.shape[1], idx.shape[2]
    require(nsample >= 1, "GroupPoint expects (batch_size, npoints, nsample) idx shape")
    require(tuple(new_xyz.shape) == (b, m, 3), "GroupPoint expects (batch_size, npoints, 3) new_xyz shape")
    c = 0
    if points is not None:
        points = dev_tensor(points, torch.float32, "points")
        require(points.dim() == 3 and points.shape[0] == b and points.shape[1] == n,
                "GroupPoint expects (batch_size, num_points, channel) points shape")
        c = points.shape[2]
    out = torch.empty((b, m, nsample, 3 + c), dtype=torch.float32, device=xyz.device)
    with torch.cuda.device(xyz.device):
        fn = _lib.load().hf_group_point_concat_msg if xyz_last else _lib.load().hf_group_point_concat
        rc = fn(b, n, c, m, nsample, xyz.data_ptr(), points.data_ptr() if c else None, idx.data_ptr(), new_xyz.data_ptr(),
                out.data_ptr(), stream_ptr(xyz.device))
    _lib.check(rc, "group_point_concat")
    return out
