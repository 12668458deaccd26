"""Rotated BEV IoU and oriented NMS — mirror of the reference's bev_iou/bev_iou.py.

Native code: csrc/bev_iou.cu via include/hfops.h.
"""
import torch

from .. import _lib
from .._tensors import dev_tensor, require, stream_ptr


def compute_bev_iou(proposals, gt_bboxes):
    """
    input:
        proposals: (N, 5), [x1, y1, x2, y2, ry], float32
        gt_bboxes: (M, 5), [x1, y1, x2, y2, ry]  float32
    output:
        overlap_area:   (N, M)   float32
        bev_iou:   (N, M)   float32

    Reference: bev_iou/bev_iou.py:13-24, op checks bev_iou/bev_iou.cpp:147-179.
    """
    proposals = dev_tensor(proposals, torch.float32, "proposals")
    gt_bboxes = dev_tensor(gt_bboxes, torch.float32, "gt_bboxes")
    require(proposals.dim() == 2 and proposals.shape[0] > 0 and proposals.shape[1] == 5,
            "ComputeIOU3D expects (N, 5) proposals shape")
    require(gt_bboxes.dim() == 2 and gt_bboxes.shape[0] > 0 and gt_bboxes.shape[1] == 5,
            "ComputeIOU3D expects (M, 5) gt_bboxes shape")
    n, m = proposals.shape[0], gt_bboxes.shape[0]
    overlap = torch.empty((n, m), dtype=torch.float32, device=proposals.device)
    iou = torch.empty((n, m), dtype=torch.float32, device=proposals.device)
    with torch.cuda.device(proposals.device):
        rc = _lib.load().hf_compute_bev_iou(n, proposals.data_ptr(), m, gt_bboxes.data_ptr(), overlap.data_ptr(),
                                            iou.data_ptr(), stream_ptr(proposals.device))
    _lib.check(rc, "compute_bev_iou")
    return overlap, iou


def oriented_nms_batch(boxes, thresh, return_num_kept=False, max_keep=None):
    """oriented_nms over S independent, score-sorted box sets in one call.

    boxes: (S, N, 5) -> keep_idx (S, N) int32 [, num_kept (S,) int32].
    The reference runs scenes one by one through tf.map_fn (hf/core/models/rpn_model.py:683).
    max_keep: the caller only uses keep_idx[:, :max_keep] (sb_nms_fn slices to nms_size, model_util.py:126): the
    result is that slice, (S, max_keep), num_kept = min(kept, max_keep), and the greedy sweep stops there.
    """
    require(float(thresh) >= 0, "Need nms_threshold >= 0, got %r" % (thresh,))
    boxes = dev_tensor(boxes, torch.float32, "boxes")
    require(boxes.dim() == 3 and boxes.shape[1] > 0 and boxes.shape[2] == 5, "OrientendNMS expects (N, 5) boxes shape")
    s, n = boxes.shape[0], boxes.shape[1]
    lib = _lib.load()
    dev = boxes.device
    require(max_keep is None or (isinstance(max_keep, int) and max_keep >= 1), "max_keep must be a positive int")
    width = n if max_keep is None else min(max_keep, n)
    keep = torch.empty((s, width), dtype=torch.int32, device=dev)
    num_kept = torch.empty((s,), dtype=torch.int32, device=dev)
    ws_bytes = lib.hf_oriented_nms_workspace_bytes(s, n)
    ws = torch.empty((max(ws_bytes, 8) // 8,), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = lib.hf_oriented_nms_topk(boxes.data_ptr(), s, n, float(thresh), width, keep.data_ptr(), num_kept.data_ptr(),
                                      ws.data_ptr(), ws.numel() * 8, stream_ptr(dev))
    _lib.check(rc, "oriented_nms")
    return (keep, num_kept) if return_num_kept else keep


def oriented_nms(boxes, thresh):
    """
    input:
        boxes: (N, 5), [x1, y1, x2, y2, ry], float32
    output:
        keep_idx: (N), int32

    Reference: bev_iou/bev_iou.py:30-39, op bev_iou/bev_iou.cpp:54-118.  Boxes must already
    be sorted by descending score; kept indices come first (ascending), the tail is padded
    with keep_idx[0].  Mask, greedy sweep and padding all run on the device.
    """
    require(float(thresh) >= 0, "Need nms_threshold >= 0, got %r" % (thresh,))  # attr check, bev_iou.cpp:43-47
    boxes = dev_tensor(boxes, torch.float32, "boxes")
    require(boxes.dim() == 2 and boxes.shape[0] > 0 and boxes.shape[1] == 5, "OrientendNMS expects (N, 5) boxes shape")
    return oriented_nms_batch(boxes.unsqueeze(0), thresh)[0]


def oriented_nms_mask(boxes, thresh, upper_only=False):
    """The native launcher's own output (bev_iou/bev_iou_g.cu:317): (N, ceil(N/64)) uint64 mask as int64."""
    require(float(thresh) >= 0, "Need nms_threshold >= 0, got %r" % (thresh,))
    boxes = dev_tensor(boxes, torch.float32, "boxes")
    require(boxes.dim() == 2 and boxes.shape[0] > 0 and boxes.shape[1] == 5, "OrientendNMS expects (N, 5) boxes shape")
    n = boxes.shape[0]
    cb = (n + 63) // 64
    mask = torch.empty((n, cb), dtype=torch.int64, device=boxes.device)
    with torch.cuda.device(boxes.device):
        rc = _lib.load().hf_oriented_nms_mask(boxes.data_ptr(), mask.data_ptr(), n, float(thresh),
                                              1 if upper_only else 0, stream_ptr(boxes.device))
    _lib.check(rc, "oriented_nms_mask")
    return mask
