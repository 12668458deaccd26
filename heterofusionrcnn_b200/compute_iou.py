"""Mirror of hf/core/compute_iou.py: the glue between box tensors and the BEV IoU / NMS ops.

Same function names and argument order as the reference (boxes3d_to_bev_tf, box3d_iou_tf,
oriented_nms_tf); eager on torch CUDA tensors, each one C-ABI call (csrc/glue.cu).  oriented_nms_tf also
accepts a batch (S,N,7) / (S,N), replacing the tf.map_fn over scenes of hf/core/models/rpn_model.py:683.
"""
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


def boxes3d_to_bev_tf(boxes3d):
    """
    Input:
        boxes3d: (N, 7) [x, y, z, l, w, h, ry]
    Output:
        boxes_bev: (N, 5) [x1, y1, x2, y2, ry]

    Reference: hf/core/compute_iou.py:7-20.
    """
    boxes3d = dev_tensor(boxes3d, torch.float32, "boxes3d")
    require(boxes3d.dim() == 2 and boxes3d.shape[1] == 7, "boxes3d must be (N, 7)")
    n = boxes3d.shape[0]
    bev = torch.empty((n, 5), dtype=torch.float32, device=boxes3d.device)
    with torch.cuda.device(boxes3d.device):
        rc = _lib.load().hf_boxes3d_to_bev(n, boxes3d.data_ptr(), bev.data_ptr(), stream_ptr(boxes3d.device))
    _lib.check(rc, "boxes3d_to_bev")
    return bev


def box3d_iou_tf(boxes_a, boxes_b):
    """Compute 3D bounding box IoU for Oriented BBox.

    Input:
        boxes_a: (N, 7) [x, y, z, h, w, l, ry]
        boxes_b: (M, 7) [x, y, z, h, w, l, ry]
    Output:
        iou_3d: (N, M) 3D bounding box IoU
        iou_2d: (N, M) bird's eye view 2D bounding box IoU

    Reference: hf/core/compute_iou.py:23-64.
    """
    boxes_a = dev_tensor(boxes_a, torch.float32, "boxes_a")
    boxes_b = dev_tensor(boxes_b, torch.float32, "boxes_b")
    require(boxes_a.dim() == 2 and boxes_a.shape[1] == 7 and boxes_b.dim() == 2 and boxes_b.shape[1] == 7,
            "boxes must be (N, 7) and (M, 7)")
    n, m = boxes_a.shape[0], boxes_b.shape[0]
    iou_3d = torch.zeros((n, m), dtype=torch.float32, device=boxes_a.device)
    iou_2d = torch.zeros((n, m), dtype=torch.float32, device=boxes_a.device)
    if n == 0 or m == 0:
        return iou_3d, iou_2d
    lib = _lib.load()
    ws_bytes = int(lib.hf_box3d_iou_workspace_bytes(n, m))
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=boxes_a.device)
    with torch.cuda.device(boxes_a.device):
        rc = lib.hf_box3d_iou(n, boxes_a.data_ptr(), m, boxes_b.data_ptr(), iou_3d.data_ptr(), iou_2d.data_ptr(),
                              ws.data_ptr(), ws_bytes, stream_ptr(boxes_a.device))
    _lib.check(rc, "box3d_iou")
    return iou_3d, iou_2d


def oriented_nms_tf(boxes, scores, thresh, return_num_kept=False):
    """
    Inputs:
        boxes: (N, 7) [x, y, z, h, w, l, ry]      (or (S, N, 7): S independent scenes)
        scores: (N)                                (or (S, N))
        thresh: scalar. float
    Outputs:
        keep_idx: (N)                              (or (S, N)) indices into `boxes`, best score first;
                  the tail behind the kept boxes repeats the first kept index (bev_iou.cpp:104-112)

    Reference: hf/core/compute_iou.py:67-80.
    """
    require(float(thresh) >= 0, "Need nms_threshold >= 0, got %r" % (thresh,))
    boxes = dev_tensor(boxes, torch.float32, "boxes")
    scores = dev_tensor(scores, torch.float32, "scores")
    single = boxes.dim() == 2
    if single:
        boxes, scores = boxes.unsqueeze(0), scores.unsqueeze(0)
    require(boxes.dim() == 3 and boxes.shape[2] == 7, "boxes must be (N, 7) or (S, N, 7)")
    s, n = boxes.shape[0], boxes.shape[1]
    require(tuple(scores.shape) == (s, n), "scores must be (N) or (S, N)")
    keep = torch.zeros((s, n), dtype=torch.int32, device=boxes.device)
    num = torch.zeros((s,), dtype=torch.int32, device=boxes.device)
    if s and n:
        lib = _lib.load()
        ws_bytes = int(lib.hf_oriented_nms_scored_workspace_bytes(s, n))
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=boxes.device)
        with torch.cuda.device(boxes.device):
            rc = lib.hf_oriented_nms_scored(boxes.data_ptr(), scores.data_ptr(), s, n, float(thresh), keep.data_ptr(),
                                            num.data_ptr(), ws.data_ptr(), ws_bytes, stream_ptr(boxes.device))
        _lib.check(rc, "oriented_nms")
    if single:
        keep, num = keep[0], num[0]
    return (keep, num) if return_num_kept else keep
