"""RoI point crop — mirror of the reference's cropping/tf_cropping.py.

Native code: csrc/cropping.cu via include/hfops.h.
"""
import torch

from .. import _lib
from .._tensors import dev_tensor, require, scatter_workspace, stream_ptr


def pc_crop_and_sample(pts, fts, intensities, mask, boxes, box_ind, resize, index=None):
    """
    input:
        pts:    (B, P, 3), [x, y, z],   float32
        fts:    (B, P, C), [c1,c2,...]  float32
        intensities: (B, P, 1), [i] float32
        mask:   (B, P) point mask label, bool
        boxes:  (N, 3, 8), 8 corners (x,y,z), float32
        box_ind:(N)        [b]  which batch that box belongs to, int32
        resize = R, int32
    output:
        crop_pts:   (N, R, 3)   float32
        crop_fts:   (N, R, C)   float32
        crop_intensity: (N, R, 1) float32
        crop_mask:  (N, R)      bool
        crop_ind:   (N, R)      int32,  which point to crop, complementary to box_ind
        non_empty_box_mask:  (N)         bool

    Reference: cropping/tf_cropping.py:13-33, op checks cropping/tf_cropping.cpp:60-190.
    Slot order is deterministic here (ascending point index, first R kept); in the
    reference it is decided by an atomicInc race (cropping/tf_cropping_g.cu:85).
    """
    require(int(resize) > 0, "PcCropAndSample expects positive resize")
    pts = dev_tensor(pts, torch.float32, "pts")
    fts = dev_tensor(fts, torch.float32, "fts")
    intensities = dev_tensor(intensities, torch.float32, "intensities")
    mask = dev_tensor(mask, torch.bool, "mask")
    boxes = dev_tensor(boxes, torch.float32, "boxes")
    box_ind = dev_tensor(box_ind, torch.int32, "box_ind")
    require(pts.dim() == 3 and pts.shape[1] > 0 and pts.shape[2] == 3, "PcCropAndSample expects (B, P, 3) pts shape")
    require(fts.dim() == 3 and fts.shape[1] > 0 and fts.shape[2] > 0, "PcCropAndSample expects (B, P, C) fts shape")
    require(intensities.dim() == 3 and intensities.shape[1] > 0 and intensities.shape[2] == 1,
            "PcCropAndSample expects (B, P, 1) intensities shape")
    require(pts.shape[0] == fts.shape[0] and pts.shape[1] == fts.shape[1],
            "PcCropAndSample expects pts & fts has same (B, P, ...) shape")
    require(intensities.shape[0] == pts.shape[0] and intensities.shape[1] == fts.shape[1],
            "PcCropAndSample expects intensities & fts has same (B, P, ...) shape")
    require(mask.dim() == 2 and mask.shape[0] == pts.shape[0] and mask.shape[1] == pts.shape[1],
            "PcCropAndSample expects (B, P) mask shape")
    require(boxes.dim() == 3 and boxes.shape[1] == 3 and boxes.shape[2] == 8, "boxes must be (N, 3, 8)")
    require(box_ind.dim() == 1 and box_ind.shape[0] == boxes.shape[0], "box_index has incompatible shape")
    B, P, C = fts.shape
    IC = intensities.shape[2]
    N, R = boxes.shape[0], int(resize)
    dev = pts.device
    crop_pts = torch.empty((N, R, 3), dtype=torch.float32, device=dev)
    crop_fts = torch.empty((N, R, C), dtype=torch.float32, device=dev)
    crop_int = torch.empty((N, R, IC), dtype=torch.float32, device=dev)
    crop_mask = torch.empty((N, R), dtype=torch.bool, device=dev)
    crop_ind = torch.empty((N, R), dtype=torch.int32, device=dev)
    non_empty = torch.empty((N,), dtype=torch.bool, device=dev)
    ix = index.ptr_for(pts) if index is not None else None  # spatial.CloudIndex of pts (optional accelerator)
    with torch.cuda.device(dev):
        rc = _lib.load().hf_pc_crop_and_sample_indexed(
            pts.data_ptr(), fts.data_ptr(), intensities.data_ptr(), mask.data_ptr(), boxes.data_ptr(),
            box_ind.data_ptr(), N, B, P, R, C, IC, crop_pts.data_ptr(), crop_fts.data_ptr(), crop_int.data_ptr(),
            crop_mask.data_ptr(), crop_ind.data_ptr(), non_empty.data_ptr(), ix, stream_ptr(dev))
    _lib.check(rc, "pc_crop_and_sample")
    return crop_pts, crop_fts, crop_int, crop_mask, crop_ind, non_empty


def pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, grad_crop_fts):
    """Gradient of pc_crop_and_sample w.r.t. fts (the reference's PcCropAndSampleGradFts op,
    cropping/tf_cropping.py:36-46, op checks cropping/tf_cropping.cpp:200-235).

    fts (B,P,C) [shape only], box_ind (N) int32, crop_ind (N,R) int32, grad_crop_fts (N,R,C) -> grad_fts (B,P,C).
    Deterministic: a point's contributions are added in ascending (box, slot) order.
    """
    fts = dev_tensor(fts, torch.float32, "fts")
    box_ind = dev_tensor(box_ind, torch.int32, "box_ind")
    crop_ind = dev_tensor(crop_ind, torch.int32, "crop_ind")
    grad_crop_fts = dev_tensor(grad_crop_fts, torch.float32, "grad_crop_fts")
    require(fts.dim() == 3 and grad_crop_fts.dim() == 3 and fts.shape[2] == grad_crop_fts.shape[2],
            "PcCropAndSampleGradFts expects fts and grad_crop_fts has same shape(2)")
    require(box_ind.dim() == 1 and crop_ind.dim() == 2 and box_ind.shape[0] == crop_ind.shape[0] == grad_crop_fts.shape[0],
            "PcCropAndSampleGradFts expects box_ind, crop_ind and grad_crop_fts has same shape(0)")
    require(crop_ind.shape[1] == grad_crop_fts.shape[1],
            "PcCropAndSampleGradFts expects crop_ind and grad_crop_fts has same shape(1)")
    batch, npts, c = fts.shape
    nb, r = crop_ind.shape
    require(npts > 0 and c > 0 and batch > 0, "PcCropAndSampleGradFts expects (B, P, C) fts shape")
    grad_fts = torch.empty((batch, npts, c), dtype=torch.float32, device=fts.device)
    if nb == 0 or r == 0:
        return grad_fts.zero_()
    lib = _lib.load()
    ws, ws_bytes = scatter_workspace(lib, 1, batch * npts, nb * r, fts.device)
    with torch.cuda.device(fts.device):
        rc = lib.hf_pc_crop_and_sample_grad_fts(box_ind.data_ptr(), crop_ind.data_ptr(), grad_crop_fts.data_ptr(), nb, batch,
                                                npts, r, c, grad_fts.data_ptr(), ws.data_ptr(), ws_bytes,
                                                stream_ptr(fts.device))
    _lib.check(rc, "pc_crop_and_sample_grad_fts")
    return grad_fts
