"""hfops-b200: B200-native (sm_100a) implementation of HeteroFusionRCNN's point-cloud operator path.

Drop-in modules (same function names, argument order and shapes as the reference's
tf_sampling / tf_grouping / tf_interpolate / tf_cropping / bev_iou), eager on torch CUDA
tensors, calling hand-written CUDA through the C-ABI of include/hfops.h:

    from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point
    from heterofusionrcnn_b200.grouping.tf_grouping import query_ball_point, group_point
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn, three_interpolate
    from heterofusionrcnn_b200.cropping.tf_cropping import pc_crop_and_sample
    from heterofusionrcnn_b200.bev_iou.bev_iou import compute_bev_iou, oriented_nms
"""
import os as _os

# The chain spreads one pass over 13 CUDA streams; with the default 8 hardware queues several of
# them alias and pick up false dependencies (a three_nn queued behind the NMS sweep).  Must be set
# before the CUDA context exists, so it is done at import; an explicit setting wins.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from ._lib import HfOpsError, InvalidArgumentError, launch_count, load  # noqa: F401

__version__ = "0.2.0"


def version():
    """The native library's version string, e.g. 'hfops-b200 0.2.0 (sm_100a) build 1a2b3c4d5e6f'."""
    return load().hf_version().decode()


def build_id():
    """The 12-hex-digit hash of the sources libhfops.so was built from (heterofusionrcnn_b200/build.py:source_id)."""
    return version().rsplit(" ", 1)[-1]
