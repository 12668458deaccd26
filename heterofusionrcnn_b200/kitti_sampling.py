"""Fixed-size sampling of a lidar scan — the operator part of KittiDataset.load_samples
(hf/datasets/kitti/kitti_dataset.py:341-371) on the device, behind hf_lidar_fov_filter and in front of FPS.

The reference draws with numpy's global RNG; here the randomness is drawn by the caller (draw_randomness) without any
knowledge of the scans and handed in, and the near/far split, the draws, the shuffle, the gather and the intensity
shift run in one kernel (csrc/pc_sample.cu via include/hfops.h).
"""
import numpy as np
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr

DEPTH_SPLIT = 40.0  # kitti_dataset.py:343
INTENSITY_SHIFT = 0.5  # kitti_dataset.py:365
MAX_POINTS = 131072


def draw_randomness(rng, scans, n_cap, pc_sample_pts):
    """Host-side draws standing in for the reference's np.random calls: (keys, slot_of, extra) numpy arrays
    (uint32 (S,n_cap), int32 (S,P) permutations, uint32 (S,P)); see include/hfops.h:hf_pc_sample_points."""
    keys = rng.integers(0, 2 ** 32, (scans, n_cap), dtype=np.uint32)
    slot_of = np.stack([rng.permutation(pc_sample_pts) for _ in range(scans)]).astype(np.int32) if scans else \
        np.zeros((0, pc_sample_pts), np.int32)
    extra = rng.integers(0, 2 ** 32, (scans, pc_sample_pts), dtype=np.uint32)
    return keys, slot_of, extra


def sample_point_cloud(pts, counts, pc_sample_pts, keys, slot_of, extra=None, return_choice=False):
    """pts (S,n_cap,4) float32 [x,y,z,intensity] and counts (S) int32 (as hf_lidar_fov_filter writes them, on the
    device), keys (S,n_cap) / slot_of (S,P) / extra (S,P) from draw_randomness -> sampled_pc (S,P,4) float32 with the
    intensity shifted to [-0.5, 0.5], status (S) int32 (0 where the reference would raise), and with return_choice the
    source row of every output row.  Reference: hf/datasets/kitti/kitti_dataset.py:341-371."""
    require(isinstance(pc_sample_pts, int) and pc_sample_pts >= 0, "pc_sample_pts must be a non-negative int")
    pts = dev_tensor(pts, torch.float32, "pts")
    require(pts.dim() == 3 and pts.shape[2] == 4, "pts must be (S, n, 4)")
    s, n_cap = pts.shape[0], pts.shape[1]
    require(1 <= n_cap <= MAX_POINTS, "at most %d points per scan" % MAX_POINTS)
    counts = dev_tensor(counts, torch.int32, "counts")
    require(tuple(counts.shape) == (s,), "counts must be (S)")
    # the key / permutation arrays are bit patterns: int32 tensors of the same bytes are accepted too
    keys = dev_tensor(keys.view(torch.int32) if keys.dtype == torch.uint32 else keys, torch.int32, "keys")
    require(tuple(keys.shape) == (s, n_cap), "keys must be (S, n)")
    slot_of = dev_tensor(slot_of, torch.int32, "slot_of")
    require(tuple(slot_of.shape) == (s, pc_sample_pts), "slot_of must be (S, pc_sample_pts)")
    if extra is not None:
        extra = dev_tensor(extra.view(torch.int32) if extra.dtype == torch.uint32 else extra, torch.int32, "extra")
        require(tuple(extra.shape) == (s, pc_sample_pts), "extra must be (S, pc_sample_pts)")
    out = torch.empty((s, pc_sample_pts, 4), dtype=torch.float32, device=pts.device)
    choice = torch.empty((s, pc_sample_pts), dtype=torch.int32, device=pts.device) if return_choice else None
    status = torch.empty((s,), dtype=torch.int32, device=pts.device)
    with torch.cuda.device(pts.device):
        rc = _lib.load().hf_pc_sample_points(
            s, n_cap, pc_sample_pts, pts.data_ptr(), counts.data_ptr(), keys.data_ptr(), slot_of.data_ptr(),
            extra.data_ptr() if extra is not None else None, DEPTH_SPLIT, INTENSITY_SHIFT, out.data_ptr(),
            choice.data_ptr() if choice is not None else None, status.data_ptr(), stream_ptr(pts.device))
    _lib.check(rc, "pc_sample_points")
    return (out, status, choice) if return_choice else (out, status)
