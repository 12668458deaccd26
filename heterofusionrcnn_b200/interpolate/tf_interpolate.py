"""three_nn / three_interpolate — mirror of the reference's interpolate/tf_interpolate.py.

Native code: csrc/interpolate.cu via include/hfops.h.
"""
import torch

from .. import _lib
from .._tensors import dev_tensor, require, scatter_workspace, stream_ptr


def three_nn(xyz1, xyz2, index1=None, index2=None):
    """
    Input:
        xyz1: (b,n,3) float32 array, unknown points
        xyz2: (b,m,3) float32 array, known points
    Output:
        dist: (b,n,3) float32 array, distances to known points
        idx: (b,n,3) int32 array, indices to known points

    Reference: interpolate/tf_interpolate.py:11-20, op checks interpolate/tf_interpolate.cpp:63-96.
    As in the reference kernel the "distances" are SQUARED distances, ascending.
    index1 / index2 (optional, not in the reference): spatial.CloudIndex of xyz1 / xyz2, to share
    one sort between several ops on the same clouds; results do not depend on them.
    """
    xyz1 = dev_tensor(xyz1, torch.float32, "xyz1")
    xyz2 = dev_tensor(xyz2, torch.float32, "xyz2")
    require(xyz1.dim() == 3 and xyz1.shape[2] == 3, "ThreeNN expects (b,n,3) xyz1 shape.")
    require(xyz2.dim() == 3 and xyz2.shape[2] == 3, "ThreeNN expects (b,m,3) xyz2 shape.")
    require(xyz2.shape[0] == xyz1.shape[0], "ThreeNN expects xyz1 and xyz2 to share b")
    b, n, m = xyz1.shape[0], xyz1.shape[1], xyz2.shape[1]
    dist = torch.empty((b, n, 3), dtype=torch.float32, device=xyz1.device)
    idx = torch.empty((b, n, 3), dtype=torch.int32, device=xyz1.device)
    lib = _lib.load()
    p1 = index1.ptr_for(xyz1) if index1 is not None else None
    p2 = index2.ptr_for(xyz2) if index2 is not None else None
    with torch.cuda.device(xyz1.device):
        if p1 is not None and p2 is not None:
            rc = lib.hf_three_nn_indexed(b, n, m, xyz1.data_ptr(), xyz2.data_ptr(), dist.data_ptr(), idx.data_ptr(),
                                         p1, p2, stream_ptr(xyz1.device))
        else:
            # scratch for the spatially pruned kernel (csrc/spatial.cu); 0 bytes = every-pair kernel
            ws_bytes = int(lib.hf_spatial_workspace_bytes(b, n, m)) if lib.hf_spatial_applies(n, m) else 0
            ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=xyz1.device) if ws_bytes else None
            rc = lib.hf_three_nn_ws(b, n, m, xyz1.data_ptr(), xyz2.data_ptr(), dist.data_ptr(), idx.data_ptr(),
                                    ws.data_ptr() if ws is not None else None, ws_bytes, stream_ptr(xyz1.device))
    _lib.check(rc, "three_nn")
    return dist, idx


def three_nn_weights(xyz1, xyz2, index1=None, index2=None):
    """three_nn fused with the interpolation weights its caller computes from it
    (hf/core/feature_extractors/pointnet_util.py:303-308):

        dist, idx = three_nn(xyz1, xyz2); dist = max(dist, 1e-10)
        weight = (1/dist) / sum(1/dist, axis=2)

    Returns (dist (b,n,3) raw squared distances, idx (b,n,3) int32, weight (b,n,3)).
    """
    xyz1 = dev_tensor(xyz1, torch.float32, "xyz1")
    xyz2 = dev_tensor(xyz2, torch.float32, "xyz2")
    require(xyz1.dim() == 3 and xyz1.shape[2] == 3, "ThreeNN expects (b,n,3) xyz1 shape.")
    require(xyz2.dim() == 3 and xyz2.shape[2] == 3 and xyz2.shape[0] == xyz1.shape[0], "ThreeNN expects (b,m,3) xyz2 shape.")
    b, n = xyz1.shape[0], xyz1.shape[1]
    m = xyz2.shape[1]
    out = torch.empty((3, b, n, 3), dtype=torch.float32, device=xyz1.device)  # dist | idx bits | weight: one allocation
    dist, idx, weight = out[0], out[1].view(torch.int32), out[2]
    p1 = index1.ptr_for(xyz1) if index1 is not None else None
    p2 = index2.ptr_for(xyz2) if index2 is not None else None
    with torch.cuda.device(xyz1.device):
        rc = _lib.load().hf_three_nn_weights(b, n, m, xyz1.data_ptr(), xyz2.data_ptr(), dist.data_ptr(), idx.data_ptr(),
                                             weight.data_ptr(), p1, p2, stream_ptr(xyz1.device))
    _lib.check(rc, "three_nn")
    return dist, idx, weight


def three_interpolate(points, idx, weight):
    """
    Input:
        points: (b,m,c) float32 array, known points
        idx: (b,n,3) int32 array, indices to known points
        weight: (b,n,3) float32 array, weights on known points
    Output:
        out: (b,n,c) float32 array, interpolated point values

    Reference: interpolate/tf_interpolate.py:26-37 (transpose -> channel-first op -> transpose);
    here the kernel works channel-last directly, so no transposes are materialised.
    """
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    weight = dev_tensor(weight, torch.float32, "weight")
    require(points.dim() == 3, "ThreeInterpolate expects (b,m,c) points shape")
    b, m, c = points.shape
    require(idx.dim() == 3 and idx.shape[0] == b and idx.shape[2] == 3, "ThreeInterpolate expects (b,n,3) idx shape")
    n = idx.shape[1]
    require(weight.dim() == 3 and weight.shape[0] == b and weight.shape[1] == n and weight.shape[2] == 3,
            "ThreeInterpolate expects (b,n,3) weight shape")
    out = torch.empty((b, n, c), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        rc = _lib.load().hf_three_interpolate_nlc(b, m, c, n, points.data_ptr(), idx.data_ptr(), weight.data_ptr(),
                                                  out.data_ptr(), stream_ptr(points.device))
    _lib.check(rc, "three_interpolate")
    return out


def three_interpolate_channel_first(points, idx, weight):
    """The native op's own layout (interpolate/tf_interpolate.cpp:102-139): points (b,c,m) -> (b,c,n)."""
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    weight = dev_tensor(weight, torch.float32, "weight")
    require(points.dim() == 3, "ThreeInterpolate expects (b,c,m) points shape")
    b, c, m = points.shape
    require(idx.dim() == 3 and idx.shape[0] == b and idx.shape[2] == 3, "ThreeInterpolate expects (b,n,3) idx shape")
    n = idx.shape[1]
    require(weight.dim() == 3 and weight.shape[0] == b and weight.shape[1] == n and weight.shape[2] == 3,
            "ThreeInterpolate expects (b,n,3) weight shape")
    out = torch.empty((b, c, n), dtype=torch.float32, device=points.device)
    with torch.cuda.device(points.device):
        rc = _lib.load().hf_three_interpolate(b, c, m, n, points.data_ptr(), idx.data_ptr(), weight.data_ptr(),
                                              out.data_ptr(), stream_ptr(points.device))
    _lib.check(rc, "three_interpolate")
    return out


def _interp_grad_args(points, idx, weight, grad_out, channel_first):
    points = dev_tensor(points, torch.float32, "points")
    idx = dev_tensor(idx, torch.int32, "idx")
    weight = dev_tensor(weight, torch.float32, "weight")
    grad_out = dev_tensor(grad_out, torch.float32, "grad_out")
    lay = "(b,c,m)" if channel_first else "(b,m,c)"
    require(points.dim() == 3, "ThreeInterpolateGrad expects %s points shape" % lay)
    b = points.shape[0]
    c, m = (points.shape[1], points.shape[2]) if channel_first else (points.shape[2], points.shape[1])
    require(idx.dim() == 3 and idx.shape[0] == b and idx.shape[2] == 3, "ThreeInterpolateGrad expects (b,n,3) idx shape")
    n = idx.shape[1]
    require(weight.dim() == 3 and tuple(weight.shape) == (b, n, 3), "ThreeInterpolateGrad expects (b,n,3) weight shape")
    want = (b, c, n) if channel_first else (b, n, c)
    require(tuple(grad_out.shape) == want,
            "ThreeInterpolateGrad expects %s grad_out shape" % ("(b,c,n)" if channel_first else "(b,n,c)"))
    return points, idx, weight, grad_out, b, c, m, n


def three_interpolate_grad(points, idx, weight, grad_out):
    """Gradient of three_interpolate w.r.t. points in the Python API's layout
    (interpolate/tf_interpolate.py:40-49 with the transposes of :35-37 folded in):
    points (b,m,c) [shape only], idx / weight (b,n,3), grad_out (b,n,c) -> grad_points (b,m,c).
    Deterministic: contributions are added in ascending (point, neighbour) order.
    """
    points, idx, weight, grad_out, b, c, m, n = _interp_grad_args(points, idx, weight, grad_out, False)
    grad_points = torch.empty((b, m, c), dtype=torch.float32, device=points.device)
    lib = _lib.load()
    ws, ws_bytes = scatter_workspace(lib, b, m, 3 * n, points.device)
    with torch.cuda.device(points.device):
        rc = lib.hf_three_interpolate_grad_nlc(b, n, c, m, grad_out.data_ptr(), idx.data_ptr(), weight.data_ptr(),
                                               grad_points.data_ptr(), ws.data_ptr(), ws_bytes, stream_ptr(points.device))
    _lib.check(rc, "three_interpolate_grad")
    return grad_points


def three_interpolate_grad_channel_first(points, idx, weight, grad_out):
    """The native ThreeInterpolateGrad op's own layout (interpolate/tf_interpolate.cpp:141-185):
    points (b,c,m), grad_out (b,c,n) -> grad_points (b,c,m)."""
    points, idx, weight, grad_out, b, c, m, n = _interp_grad_args(points, idx, weight, grad_out, True)
    grad_points = torch.empty((b, c, m), dtype=torch.float32, device=points.device)
    lib = _lib.load()
    ws, ws_bytes = scatter_workspace(lib, b, m, 3 * n, points.device)
    with torch.cuda.device(points.device):
        rc = lib.hf_three_interpolate_grad(b, c, n, m, grad_out.data_ptr(), idx.data_ptr(), weight.data_ptr(),
                                           grad_points.data_ptr(), ws.data_ptr(), ws_bytes, stream_ptr(points.device))
    _lib.check(rc, "three_interpolate_grad")
    return grad_points
