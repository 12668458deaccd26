"""The operator part of hf/core/models/model_util.py: sb_nms_fn (:101-142), the per-scene proposal selection that
rpn_model.py:683 maps over the batch with tf.map_fn.  Here one call takes the whole batch.

Native code: csrc/glue.cu (score sort, BEV conversion, selection) + csrc/bev_iou.cu (NMS) via include/hfops.h.
"""
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


def sb_nms_fn(boxes_and_scores, nms_iou_thresh, nms_size, fixed_num_proposal_nms, bev_extents=None):
    """NMS on single batch (or on S scenes at once).
    Input:
        boxes_and_scores: A tuple with following elements:
            sb_boxes: (P, 7) [x, y, z, h, w, l, ry]          (or (S, P, 7))
            sb_scores: (P). Scores for each 3d box             (or (S, P))
        nms_iou_thresh: float. The IoU threshold used for NMS
        nms_size: int. The number of output box indices.
        fixed_num_proposal_nms: bool.  True: the output holds min(nms_size, P) indices, possibly with duplicated
            boxes (the keep list's padding repeats its first entry).  False: duplicates are removed (tf.unique) and the
            output is padded with -1.
        bev_extents: unused, as in the reference.
    Output:
        sb_nms_indices_padded: (nms_size) int32 (or (S, nms_size)).  -1 = no selection, padding.
        num_selected_boxes_before_padding: int32 scalar tensor (or (S)).

    Reference: hf/core/models/model_util.py:101-142.
    """
    sb_boxes, sb_scores = boxes_and_scores
    require(isinstance(nms_size, int) and nms_size >= 0, "sb_nms_fn expects a non-negative int nms_size")
    require(float(nms_iou_thresh) >= 0, "Need nms_threshold >= 0, got %r" % (nms_iou_thresh,))
    boxes = dev_tensor(sb_boxes, torch.float32, "sb_boxes")
    scores = dev_tensor(sb_scores, torch.float32, "sb_scores")
    single = boxes.dim() == 2
    if single:
        boxes, scores = boxes.unsqueeze(0), scores.unsqueeze(0)
    require(boxes.dim() == 3 and boxes.shape[2] == 7, "sb_boxes must be (P, 7) or (S, P, 7)")
    s, n = boxes.shape[0], boxes.shape[1]
    require(tuple(scores.shape) == (s, n), "sb_scores must be (P) or (S, P)")
    out = torch.empty((s, nms_size), dtype=torch.int32, device=boxes.device)
    num = torch.empty((s,), dtype=torch.int32, device=boxes.device)
    if s:
        lib = _lib.load()
        ws_bytes = int(lib.hf_sb_nms_workspace_bytes(s, n))
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=boxes.device)
        with torch.cuda.device(boxes.device):
            rc = lib.hf_sb_nms(boxes.data_ptr(), scores.data_ptr(), s, n, float(nms_iou_thresh), nms_size,
                               1 if fixed_num_proposal_nms else 0, out.data_ptr(), num.data_ptr(), ws.data_ptr(), ws_bytes,
                               stream_ptr(boxes.device))
        _lib.check(rc, "sb_nms_fn")
    if single:
        return out[0], num[0]
    return out, num
