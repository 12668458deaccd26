"""Mirror of the part of hf/core/box_8c_encoder.py the operator path uses: the box -> 8-corner conversion
that feeds pc_crop_and_sample (hf/core/models/rcnn_model.py:478-489)."""
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


def tf_box_3d_to_box_8co(boxes_3d, expand_length=0.0):
    """Computes the 3D bounding box corner positions from Box3D format.

    Args:
        boxes_3d: N x 7 tensor of box_3d in the format [x, y, z, l, w, h, ry]
        expand_length: (not in the reference's signature) the pooling context its caller adds first,
            hf/core/models/rcnn_model.py:462-476: y += e and l, w, h += 2 e
    Returns:
        corners_3d: a tensor of shape (N x 3 x 8): [[[x1,...,x8],[y1...,y8],[z1,...,z8]]]

    Reference: hf/core/box_8c_encoder.py:101-184 (same corner order P1..P8 as np_box_3d_to_box_8co, :8-98).
    """
    boxes_3d = dev_tensor(boxes_3d, torch.float32, "boxes_3d")
    require(boxes_3d.dim() == 2 and boxes_3d.shape[1] == 7, "boxes_3d must be (N, 7)")
    n = boxes_3d.shape[0]
    corners = torch.empty((n, 3, 8), dtype=torch.float32, device=boxes_3d.device)
    with torch.cuda.device(boxes_3d.device):
        rc = _lib.load().hf_boxes3d_to_corners(n, boxes_3d.data_ptr(), float(expand_length), corners.data_ptr(),
                                               stream_ptr(boxes_3d.device))
    _lib.check(rc, "tf_box_3d_to_box_8co")
    return corners
