"""Spatial index of a point cloud (csrc/spatial.cu), shared between the ops that search it.

A PointNet++ level runs query_ball_point and three_nn on the same two clouds; both accept the
indices of their clouds as optional keyword arguments so that each cloud is sorted once:

    ix0, ix1 = build_cloud_index(xyz), build_cloud_index(new_xyz)
    idx, cnt = query_ball_point(r, ns, xyz, new_xyz, index1=ix0, index2=ix1)
    dist, nn = three_nn(xyz, new_xyz, index1=ix0, index2=ix1)

Without them the ops keep the reference's signatures and sort internally (or run the every-pair
kernels when a call is too small to profit).  The index is only ever an accelerator: results are
identical with and without it.
"""
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


class CloudIndex:
    """Hilbert-sorted copy of a (b, n, 3) cloud + bucket bounding boxes, in device memory."""

    def __init__(self, xyz, buf):
        self.xyz, self.buf = xyz, buf  # the cloud stays referenced: the index is only valid for its contents
        self.b, self.n = xyz.shape[0], xyz.shape[1]
        self.version = xyz._version  # bumped by every in-place write torch knows about (copy_, mul_, ...)

    def matches(self, xyz):
        """True if the index was built from exactly this tensor AND the tensor has not been written in place
        since (a stale index would silently prune against the old point positions).  Writes torch cannot see
        (a raw pointer handed to another library) are the caller's responsibility: call refresh()."""
        return (xyz.data_ptr() == self.xyz.data_ptr() and tuple(xyz.shape) == (self.b, self.n, 3)
                and xyz._version == self.version)

    def refresh(self):
        """Rebuild the index in place from the current contents of its cloud (current stream)."""
        lib = _lib.load()
        with torch.cuda.device(self.xyz.device):
            rc = lib.hf_cloud_index_build(self.b, self.n, self.xyz.data_ptr(), self.buf.data_ptr(), self.buf.numel(),
                                          stream_ptr(self.xyz.device))
        _lib.check(rc, "build_cloud_index")
        self.version = self.xyz._version
        return self

    def ptr_for(self, xyz):
        """Device pointer of the index if it belongs to `xyz`, else None (the op then works without it)."""
        return self.buf.data_ptr() if self.matches(xyz) else None


def applies(nq, nc):
    """True if a call with nq query points and nc candidate points takes the pruned kernels."""
    return bool(_lib.load().hf_spatial_applies(int(nq), int(nc)))


def build_cloud_index(xyz):
    """Index of `xyz` (b, n, 3) float32 on the current stream, or None if the cloud is too large to index."""
    xyz = dev_tensor(xyz, torch.float32, "xyz")
    require(xyz.dim() == 3 and xyz.shape[2] == 3, "CloudIndex expects (batch_size, num_points, 3) xyz shape")
    lib = _lib.load()
    b, n = xyz.shape[0], xyz.shape[1]
    nbytes = int(lib.hf_cloud_index_bytes(b, n))
    if nbytes == 0:
        return None
    buf = torch.empty((nbytes,), dtype=torch.uint8, device=xyz.device)
    with torch.cuda.device(xyz.device):
        rc = lib.hf_cloud_index_build(b, n, xyz.data_ptr(), buf.data_ptr(), nbytes, stream_ptr(xyz.device))
    _lib.check(rc, "build_cloud_index")
    return CloudIndex(xyz, buf)
