"""Farthest point sampling / gather_point — mirror of the reference's sampling/tf_sampling.py.

Same names, argument order and shapes; tensors are torch CUDA tensors (used as device
buffers) instead of lazy TF graph nodes.  Native code: csrc/sampling.cu via include/hfops.h.
"""
import torch

from .. import _lib
from .._tensors import dev_tensor, require, scatter_workspace, stream_ptr

# FPS keeps clouds of up to this many points entirely on chip; beyond it the kernel needs
# the reference's (32, n) float scratch (sampling/tf_sampling.cpp:115).
_ON_CHIP_POINTS = 65536


def gather_point(inp, idx):
    """
input:
    batch_size * ndataset * 3   float32
    batch_size * npoints        int32
returns:
    batch_size * npoints * 3    float32

    Reference: sampling/tf_sampling.py:38-46, op checks sampling/tf_sampling.cpp:126-148.
    """
    inp = dev_tensor(inp, torch.float32, "inp")
    idx = dev_tensor(idx, torch.int32, "idx")
    require(inp.dim() == 3 and inp.shape[2] == 3, "GatherPoint expects (batch_size,num_points,3) inp shape")
    b, n = inp.shape[0], inp.shape[1]
    require(idx.dim() == 2 and idx.shape[0] == b, "GatherPoint expects (batch_size,num_result) idx shape")
    m = idx.shape[1]
    out = torch.empty((b, m, 3), dtype=torch.float32, device=inp.device)
    with torch.cuda.device(inp.device):
        rc = _lib.load().hf_gather_point(b, n, m, inp.data_ptr(), idx.data_ptr(), out.data_ptr(), stream_ptr(inp.device))
    _lib.check(rc, "gather_point")
    return out


def farthest_point_sample(npoint, inp):
    """
input:
    int32
    batch_size * ndataset * 3   float32
returns:
    batch_size * npoint         int32

    Reference: sampling/tf_sampling.py:61-69, op checks sampling/tf_sampling.cpp:95-123.
    Starts at index 0; ties resolve exactly as the reference kernel does
    (sampling/tf_sampling_g.cu:146-163).
    """
    require(isinstance(npoint, int) and npoint > 0, "FarthestPointSample expects positive npoint")
    inp = dev_tensor(inp, torch.float32, "inp")
    require(inp.dim() == 3 and inp.shape[2] == 3, "FarthestPointSample expects (batch_size,num_points,3) inp shape")
    b, n = inp.shape[0], inp.shape[1]
    out = torch.empty((b, npoint), dtype=torch.int32, device=inp.device)
    temp = None
    if n > _ON_CHIP_POINTS:
        temp = torch.empty((32, n), dtype=torch.float32, device=inp.device)
    with torch.cuda.device(inp.device):
        rc = _lib.load().hf_farthest_point_sample(
            b, n, npoint, inp.data_ptr(), temp.data_ptr() if temp is not None else None, out.data_ptr(),
            stream_ptr(inp.device))
    _lib.check(rc, "farthest_point_sample")
    return out


def farthest_point_sample_with_xyz(npoint, inp, index=None):
    """farthest_point_sample + gather_point of one set-abstraction level in one launch
    (hf/core/feature_extractors/pointnet_util.py:42-44: `new_xyz = gather_point(xyz, farthest_point_sample(npoint, xyz))`).

    index: optional spatial.CloudIndex of `inp` (the one ball query / three_nn / the crop of the level use anyway): the
    kernel then skips, warp by warp, the accepted centres that cannot reach the warp's region of the cloud.
    Returns (idx (b,npoint) int32, new_xyz (b,npoint,3) float32); both identical to the two separate ops.
    """
    require(isinstance(npoint, int) and npoint > 0, "FarthestPointSample expects positive npoint")
    inp = dev_tensor(inp, torch.float32, "inp")
    require(inp.dim() == 3 and inp.shape[2] == 3, "FarthestPointSample expects (batch_size,num_points,3) inp shape")
    b, n = inp.shape[0], inp.shape[1]
    out = torch.empty((b, npoint), dtype=torch.int32, device=inp.device)
    new_xyz = torch.empty((b, npoint, 3), dtype=torch.float32, device=inp.device)
    ix = index.ptr_for(inp) if index is not None else None
    if ix is not None and n <= _ON_CHIP_POINTS:
        with torch.cuda.device(inp.device):
            rc = _lib.load().hf_farthest_point_sample_indexed(b, n, npoint, inp.data_ptr(), ix, out.data_ptr(),
                                                              new_xyz.data_ptr(), stream_ptr(inp.device))
        _lib.check(rc, "farthest_point_sample_indexed")
        return out, new_xyz
    temp = torch.empty((32, n), dtype=torch.float32, device=inp.device) if n > _ON_CHIP_POINTS else None
    with torch.cuda.device(inp.device):
        rc = _lib.load().hf_farthest_point_sample_xyz(
            b, n, npoint, inp.data_ptr(), temp.data_ptr() if temp is not None else None, out.data_ptr(),
            new_xyz.data_ptr(), stream_ptr(inp.device))
    _lib.check(rc, "farthest_point_sample_xyz")
    return out, new_xyz


PYRAMID_MAX_LEVELS = 3


def farthest_point_sample_pyramid(npoints, inp):
    """Consecutive set-abstraction sampling levels in one launch: level j draws npoints[j] points from the picks of
    level j-1 (level 0 from inp), i.e. the `farthest_point_sample` + `gather_point` pairs of successive
    pointnet_sa_module calls (pointnet_util.py:42-44) with nothing else in between.

    npoints: list of 1..3 ints.  Returns ([idx_j (b,npoints[j]) int32], [new_xyz_j (b,npoints[j],3) float32]),
    identical to calling farthest_point_sample / gather_point level by level.
    """
    import ctypes

    npoints = [int(v) for v in npoints]
    require(1 <= len(npoints) <= PYRAMID_MAX_LEVELS, "fps pyramid takes 1..%d levels" % PYRAMID_MAX_LEVELS)
    require(all(v > 0 for v in npoints), "FarthestPointSample expects positive npoint")
    inp = dev_tensor(inp, torch.float32, "inp")
    require(inp.dim() == 3 and inp.shape[2] == 3, "FarthestPointSample expects (batch_size,num_points,3) inp shape")
    b, n = inp.shape[0], inp.shape[1]
    require(n <= _ON_CHIP_POINTS, "fps pyramid: clouds beyond %d points are sampled level by level" % _ON_CHIP_POINTS)
    idx = [torch.empty((b, m), dtype=torch.int32, device=inp.device) for m in npoints]
    xyz = [torch.empty((b, m, 3), dtype=torch.float32, device=inp.device) for m in npoints]
    L = len(npoints)
    m_arr = (ctypes.c_int * L)(*npoints)
    idx_arr = (ctypes.c_void_p * L)(*[t.data_ptr() for t in idx])
    xyz_arr = (ctypes.c_void_p * L)(*[t.data_ptr() for t in xyz])
    with torch.cuda.device(inp.device):
        rc = _lib.load().hf_fps_pyramid(b, L, n, ctypes.cast(m_arr, ctypes.c_void_p), inp.data_ptr(),
                                        ctypes.cast(idx_arr, ctypes.c_void_p), ctypes.cast(xyz_arr, ctypes.c_void_p),
                                        stream_ptr(inp.device))
    _lib.check(rc, "fps_pyramid")
    return idx, xyz


def gather_point_grad(inp, idx, out_g):
    """Gradient of gather_point w.r.t. inp (the reference's GatherPointGrad op,
    sampling/tf_sampling.py:54-58, op checks sampling/tf_sampling.cpp:150-179).

    inp (b,n,3) [shape only], idx (b,m) int32, out_g (b,m,3) -> inp_g (b,n,3).
    Deterministic: contributions of a repeated index are added in ascending position.
    """
    inp = dev_tensor(inp, torch.float32, "inp")
    idx = dev_tensor(idx, torch.int32, "idx")
    out_g = dev_tensor(out_g, torch.float32, "out_g")
    require(inp.dim() == 3 and inp.shape[2] == 3, "GatherPointGradGpuOp expects (batch_size,num_points,3) inp")
    b, n = inp.shape[0], inp.shape[1]
    require(idx.dim() == 2 and idx.shape[0] == b, "GatherPointGradGpuOp expects (batch_size,num_result) idx shape")
    m = idx.shape[1]
    require(out_g.dim() == 3 and tuple(out_g.shape) == (b, m, 3),
            "GatherPointGradGpuOp expects (batch_size,num_result,3) out_g shape")
    inp_g = torch.empty((b, n, 3), dtype=torch.float32, device=inp.device)
    lib = _lib.load()
    ws, ws_bytes = scatter_workspace(lib, b, n, m, inp.device)
    with torch.cuda.device(inp.device):
        rc = lib.hf_gather_point_grad(b, n, m, out_g.data_ptr(), idx.data_ptr(), inp_g.data_ptr(), ws.data_ptr(), ws_bytes,
                                      stream_ptr(inp.device))
    _lib.check(rc, "gather_point_grad")
    return inp_g


def prob_sample(inp, inpr):
    """
input:
    batch_size * ncategory float32
    batch_size * npoints   float32
returns:
    batch_size * npoints   int32

    Reference: sampling/tf_sampling.py:18-27, op checks sampling/tf_sampling.cpp:68-93.  inp holds
    (unnormalised) category weights, inpr uniform random numbers in [0, 1); the result is the index
    the reference's binary search finds in its float32 prefix sum (same rounding, bit for bit).
    """
    inp = dev_tensor(inp, torch.float32, "inp")
    inpr = dev_tensor(inpr, torch.float32, "inpr")
    require(inp.dim() == 2 and inp.shape[1] >= 1, "ProbSample expects (batch_size,num_choices) inp shape")
    b, n = inp.shape
    require(inpr.dim() == 2 and inpr.shape[0] == b, "ProbSample expects (batch_size,num_points) inpr shape")
    m = inpr.shape[1]
    out = torch.empty((b, m), dtype=torch.int32, device=inp.device)
    temp = torch.empty((b, n), dtype=torch.float32, device=inp.device)
    with torch.cuda.device(inp.device):
        rc = _lib.load().hf_prob_sample(b, n, m, inp.data_ptr(), inpr.data_ptr(), temp.data_ptr(), out.data_ptr(),
                                        stream_ptr(inp.device))
    _lib.check(rc, "prob_sample")
    return out
