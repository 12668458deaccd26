"""Input side of the point path on the GPU: hf/core/obj_utils.py:221-277 (get_lidar_point_cloud) with
hf/core/calib_utils.py:370-407 (lidar_to_cam_frame) and :280-296 (project_to_image).

The reference reads the velodyne `.bin` with numpy, moves it to the rectified camera frame and keeps the points in
front of the camera that project into the image, all in float64 on the host.  Here the raw float32 scan goes to the
device once and `hf_lidar_fov_filter` (csrc/lidar_input.cu) does the transform (float64) and the order-preserving
filter; the two small calibration products stay on the host in float64 exactly as the reference forms them.
"""
import numpy as np
import torch

from . import _lib
from ._tensors import dev_tensor, require, stream_ptr


def rectified_velo_to_cam(r0_rect, tr_velodyne_to_cam):
    """r0_rect (3,3) and tr_velodyne_to_cam (3,4) -> the first three rows of the padded 4x4 product
    (calib_utils.py:388-403), float64 (3,4)."""
    r0 = np.zeros((4, 4), dtype=np.float64)
    r0[:3, :3] = np.asarray(r0_rect, dtype=np.float64)
    r0[3, 3] = 1.0
    tf = np.zeros((4, 4), dtype=np.float64)
    tf[:3, :4] = np.asarray(tr_velodyne_to_cam, dtype=np.float64)
    tf[3, 3] = 1.0
    return np.dot(r0, tf)[:3]


def pack_calib(r0_rect, tr_velodyne_to_cam, p2):
    """One scan's 24 float64 calibration words of hf_lidar_fov_filter."""
    p2 = np.asarray(p2, dtype=np.float64)
    require(p2.shape == (3, 4), "p2 must be (3, 4)")
    return np.concatenate([rectified_velo_to_cam(r0_rect, tr_velodyne_to_cam).reshape(-1), p2.reshape(-1)])


def lidar_fov_filter(points, calib, im_size):
    """
    Input:
        points: (S, n, 4) float32 CUDA tensor, raw velodyne [x, y, z, i] (S scans padded to the same n; pad with
                points behind the sensor, e.g. x = -1, which the filter drops)
        calib: (S, 24) float64 (numpy or CUDA tensor), `pack_calib` per scan
        im_size: [w, h]
    Output:
        out: (S, n, 4) float32 [x, y, z, i] in the rectified camera frame; rows [0, count[s]) are the kept points of
             scan s in scan order, the rest is unspecified
        count: (S,) int32
    """
    points = dev_tensor(points, torch.float32, "points")
    require(points.dim() == 3 and points.shape[2] == 4, "points must be (S, n, 4)")
    s, n = int(points.shape[0]), int(points.shape[1])
    if not isinstance(calib, torch.Tensor):
        calib = torch.from_numpy(np.ascontiguousarray(calib, dtype=np.float64)).to(points.device)
    calib = dev_tensor(calib, torch.float64, "calib")
    require(tuple(calib.shape) == (s, 24), "calib must be (S, 24)")
    require(len(im_size) == 2 and int(im_size[0]) > 0 and int(im_size[1]) > 0, "im_size must be [w, h] > 0")
    lib = _lib.load()
    out = torch.empty((s, n, 4), dtype=torch.float32, device=points.device)
    count = torch.empty((s,), dtype=torch.int32, device=points.device)
    nbytes = int(lib.hf_lidar_fov_filter_workspace_bytes(s, n))
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=points.device)
    with torch.cuda.device(points.device):
        rc = lib.hf_lidar_fov_filter(
            s, n, points.data_ptr(), calib.data_ptr(), int(im_size[0]), int(im_size[1]), out.data_ptr(),
            count.data_ptr(), ws.data_ptr(), nbytes, stream_ptr(points.device),
        )
    _lib.check(rc, "lidar_fov_filter")
    return out, count


def read_velodyne(path):
    """The float32 (n, 4) [x, y, z, i] records of a KITTI velodyne file (calib_utils.py:352-366)."""
    return np.fromfile(path, dtype=np.float32).reshape(-1, 4)


def read_calibration(path):
    """P2, R0_rect and Tr_velo_to_cam of a KITTI calibration text file (`key: v v v ...` lines), float64."""
    rows = {}
    with open(path, "r") as f:
        for line in f:
            key, sep, rest = line.partition(":")
            if sep and rest.strip():
                rows[key.strip()] = np.array(rest.split(), dtype=np.float64)
    for key in ("P2", "R0_rect", "Tr_velo_to_cam"):
        require(key in rows, "calibration file %s has no %s row" % (path, key))
    return rows["P2"].reshape(3, 4), rows["R0_rect"].reshape(3, 3), rows["Tr_velo_to_cam"].reshape(3, 4)


def get_lidar_point_cloud(img_idx, calib_dir, velo_dir, im_size, device="cuda"):
    """
    The reference's call with `im_size` given: the scan's points in the image field of view.

    Input:
        img_idx: sample index; calib_dir, velo_dir: KITTI directories; im_size: [w, h]
    Output:
        (N, 4) float32 CUDA tensor [x, y, z, i], rectified camera frame, scan order
    """
    p2, r0, tr = read_calibration("%s/%06d.txt" % (calib_dir, img_idx))
    scan = torch.from_numpy(read_velodyne("%s/%06d.bin" % (velo_dir, img_idx))).pin_memory()
    points = scan.to(device, non_blocking=True)[None]
    out, count = lidar_fov_filter(points, pack_calib(r0, tr, p2)[None], im_size)
    return out[0, : int(count.item())]
