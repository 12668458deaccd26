"""ctypes binding of libhfops.so (C-ABI declared in include/hfops.h).

This is the only place the product touches native code.  There is NO fallback: if the
library is missing or a call fails, an exception is raised.
"""
import ctypes
import os

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "libhfops.so")

HF_OK = 0
HF_ERR_INVALID_ARG = -1
HF_ERR_CUDA = -2
HF_ERR_WORKSPACE = -3
HF_ERR_UNSUPPORTED = -4

_c_int = ctypes.c_int
_c_float = ctypes.c_float
_ptr = ctypes.c_void_p
_size_t = ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/hfops.h one to one.
SIGNATURES = {
    "hf_version": (ctypes.c_char_p, []),
    "hf_last_error": (ctypes.c_char_p, []),
    "hf_launch_count": (ctypes.c_ulonglong, []),
    "hf_fp32_probe_out_floats": (_size_t, [_c_int]),
    "hf_fp32_probe_flops": (ctypes.c_double, [_c_int, _c_int, _c_int]),
    "hf_fp32_probe": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr]),
    "hf_farthest_point_sample": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_farthest_point_sample_xyz": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_farthest_point_sample_indexed": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_fps_pyramid": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_fps_set_cluster": (None, [_c_int]),
    "hf_fps_set_threads": (None, [_c_int]),
    "hf_fps_set_variant": (None, [_c_int, _c_int, _c_int]),
    "hf_fps_get_variant": (None, [_ptr, _ptr, _ptr]),
    "hf_gather_point": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_gather_set_mode": (None, [_c_int]),
    "hf_group_point_pair": (_c_int, [_c_int, _c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_query_ball_point": (_c_int, [_c_int, _c_int, _c_int, _c_float, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_spatial_workspace_bytes": (_size_t, [_c_int, _c_int, _c_int]),
    "hf_spatial_set_mode": (None, [_c_int]),
    "hf_spatial_set_pair_counter": (None, [_ptr]),
    "hf_query_ball_point_ws": (_c_int, [_c_int, _c_int, _c_int, _c_float, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t,
                                        _ptr]),
    "hf_three_nn_ws": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_cloud_index_bytes": (_size_t, [_c_int, _c_int]),
    "hf_spatial_applies": (_c_int, [_c_int, _c_int]),
    "hf_cloud_index_build": (_c_int, [_c_int, _c_int, _ptr, _ptr, _size_t, _ptr]),
    "hf_query_ball_point_indexed": (_c_int, [_c_int, _c_int, _c_int, _c_float, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr,
                                             _ptr]),
    "hf_three_nn_indexed": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_scatter_workspace_bytes": (_size_t, [_c_int, _c_int, _c_int]),
    "hf_gather_point_grad": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_group_point_grad": (_c_int, [_c_int, _c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_three_interpolate_grad": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_three_interpolate_grad_nlc": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t,
                                               _ptr]),
    "hf_pc_crop_and_sample_grad_fts": (_c_int, [_ptr, _ptr, _ptr, _c_int, _c_int, _c_int, _c_int, _c_int, _ptr, _ptr,
                                                _size_t, _ptr]),
    "hf_three_nn_weights": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_knn_workspace_bytes": (_size_t, [_c_int, _c_int, _c_int]),
    "hf_knn_point": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_knn_indices_general": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_select_top_k": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_boxes3d_to_bev": (_c_int, [ctypes.c_longlong, _ptr, _ptr, _ptr]),
    "hf_boxes3d_to_corners": (_c_int, [ctypes.c_longlong, _ptr, _c_float, _ptr, _ptr]),
    "hf_box3d_iou_workspace_bytes": (_size_t, [_c_int, _c_int]),
    "hf_box3d_iou": (_c_int, [_c_int, _ptr, _c_int, _ptr, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_box3d_iou_static": (_c_int, [_c_int, _ptr, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_oriented_nms_scored_workspace_bytes": (_size_t, [_c_int, _c_int]),
    "hf_oriented_nms_scored": (_c_int, [_ptr, _ptr, _c_int, _c_int, _c_float, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_sb_nms_workspace_bytes": (_size_t, [_c_int, _c_int]),
    "hf_sb_nms": (_c_int, [_ptr, _ptr, _c_int, _c_int, _c_float, _c_int, _c_int, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_lidar_fov_filter_workspace_bytes": (_size_t, [_c_int, _c_int]),
    "hf_lidar_fov_filter": (_c_int, [_c_int, _c_int, _ptr, _ptr, _c_int, _c_int, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_pc_sample_points": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr, _c_float, _c_float, _ptr, _ptr, _ptr,
                                     _ptr]),
    "hf_prob_sample": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_group_point_concat": (_c_int, [_c_int] * 5 + [_ptr] * 6),
    "hf_group_point_concat_msg": (_c_int, [_c_int] * 5 + [_ptr] * 6),
    "hf_group_point": (_c_int, [_c_int, _c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_three_nn": (_c_int, [_c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_three_interpolate": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_three_interpolate_nlc": (_c_int, [_c_int, _c_int, _c_int, _c_int, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "hf_pc_crop_and_sample": (_c_int, [_ptr] * 6 + [_c_int] * 6 + [_ptr] * 6 + [_ptr]),
    "hf_pc_crop_and_sample_indexed": (_c_int, [_ptr] * 6 + [_c_int] * 6 + [_ptr] * 6 + [_ptr, _ptr]),
    "hf_compute_bev_iou": (_c_int, [_c_int, _ptr, _c_int, _ptr, _ptr, _ptr, _ptr]),
    "hf_oriented_nms_mask": (_c_int, [_ptr, _ptr, _c_int, _c_float, _c_int, _ptr]),
    "hf_oriented_nms_workspace_bytes": (_size_t, [_c_int, _c_int]),
    "hf_nms_set_spatial_sort": (None, [_c_int]),
    "hf_oriented_nms_last_pair_count": (ctypes.c_longlong, [_c_int, _c_int, _ptr, _ptr]),
    "hf_oriented_nms": (_c_int, [_ptr, _c_int, _c_int, _c_float, _ptr, _ptr, _ptr, _size_t, _ptr]),
    "hf_oriented_nms_topk": (_c_int, [_ptr, _c_int, _c_int, _c_float, _c_int, _ptr, _ptr, _ptr, _size_t, _ptr]),
}

_lib = None


class HfOpsError(RuntimeError):
    """A libhfops call returned a non-zero status."""


class InvalidArgumentError(HfOpsError, ValueError):
    """Counterpart of the reference's OP_REQUIRES -> tf.errors.InvalidArgumentError."""


def load():
    """Load libhfops.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HfOpsError(
            "libhfops.so not found at %s - build it with `python -m heterofusionrcnn_b200.build` "
            "(there is no CPU or PyTorch fallback for these ops)" % LIB_PATH
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc, what):
    if rc == HF_OK:
        return
    msg = load().hf_last_error().decode("utf-8", "replace")
    if rc == HF_ERR_INVALID_ARG:
        raise InvalidArgumentError("%s: %s" % (what, msg))
    raise HfOpsError("%s failed (status %d): %s" % (what, rc, msg))


def launch_count():
    return int(load().hf_launch_count())
