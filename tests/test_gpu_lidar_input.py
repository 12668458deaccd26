"""GPU: the input side (csrc/lidar_input.cu) through the Python mirror -> ctypes -> C-ABI against the reference's own
output on its fixture scan (tests/golden/lidar_fov_000000.npz) and against the oracle on synthetic scans.
Bar: the kept set and its order bit-exact; coordinates equal to the float32 rounding of the reference's float64
result (a 1-ulp float32 slack covers a last-bit float64 difference that lands on a float32 rounding boundary)."""
import os

import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import obj_utils
from heterofusionrcnn_b200._lib import InvalidArgumentError

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lidar_fov_000000.npz")


def close_f32(got, want64):
    want = want64.astype(np.float32)
    return np.all(np.abs(got - want) <= np.spacing(np.abs(want)))


def test_fixture_scan_equals_the_reference_output(cuda):
    g = np.load(GOLD)
    calib = obj_utils.pack_calib(g["r0_rect"], g["tr_velodyne_to_cam"], g["p2"])
    out, count = obj_utils.lidar_fov_filter(torch.from_numpy(g["raw"]).to(cuda)[None], calib[None], g["image_size"])
    n = int(count.item())
    assert n == len(g["expect"]) == 12045
    got = out[0, :n].cpu().numpy()
    assert np.array_equal(got[:, 3], g["expect"][:, 3].astype(np.float32))  # same points, same order
    assert close_f32(got[:, :3], g["expect"][:, :3])
    assert (got[:, :3] == g["expect"][:, :3].astype(np.float32)).mean() > 0.999


def synthetic_scans(seed, scans, n):
    rng = np.random.default_rng(seed)
    pts = np.empty((scans, n, 4), dtype=np.float32)
    pts[..., 0] = rng.uniform(-40, 80, (scans, n))
    pts[..., 1] = rng.uniform(-40, 40, (scans, n))
    pts[..., 2] = rng.uniform(-3, 3, (scans, n))
    pts[..., 3] = rng.uniform(0, 1, (scans, n))
    g = np.load(GOLD)
    calibs = []
    for s in range(scans):
        tr = g["tr_velodyne_to_cam"].copy()
        tr[:, 3] += rng.normal(0, 0.05, 3)
        p2 = g["p2"].copy()
        p2[0, 2] += rng.normal(0, 5.0)
        calibs.append((g["r0_rect"], tr, p2))
    return pts, calibs, g["image_size"]


@pytest.mark.parametrize("scans,n", [(1, 1), (1, 1023), (1, 1024), (3, 1025), (2, 20000), (4, 131072)])
def test_synthetic_scans_equal_the_oracle(cuda, oracle, scans, n):
    pts, calibs, im = synthetic_scans(n + scans, scans, n)
    packed = np.stack([obj_utils.pack_calib(*c) for c in calibs])
    out, count = obj_utils.lidar_fov_filter(torch.from_numpy(pts).to(cuda), packed, im)
    out, count = out.cpu().numpy(), count.cpu().numpy()
    for s in range(scans):
        want, kept = oracle.lidar_fov_filter(pts[s], *calibs[s], im)
        assert count[s] == len(kept)
        got = out[s, : count[s]]
        assert np.array_equal(got[:, 3], pts[s, kept, 3])
        assert close_f32(got[:, :3], want[:, :3])
    if n >= 20000:
        assert 0 < count.min() and count.max() < n


def test_empty_and_all_dropped_scans(cuda):
    g = np.load(GOLD)
    calib = obj_utils.pack_calib(g["r0_rect"], g["tr_velodyne_to_cam"], g["p2"])[None]
    out, count = obj_utils.lidar_fov_filter(torch.zeros((1, 0, 4), device=cuda), calib, g["image_size"])
    assert out.shape == (1, 0, 4) and int(count.item()) == 0
    behind = torch.zeros((1, 5000, 4), device=cuda)
    behind[..., 0] = -10.0
    assert int(obj_utils.lidar_fov_filter(behind, calib, g["image_size"])[1].item()) == 0
    with pytest.raises(InvalidArgumentError):
        obj_utils.lidar_fov_filter(behind, calib, [0, 375])
    with pytest.raises(InvalidArgumentError):
        obj_utils.lidar_fov_filter(behind[0], calib, g["image_size"])


def test_get_lidar_point_cloud_reads_kitti_files(cuda, tmp_path):
    g = np.load(GOLD)
    (tmp_path / "velodyne").mkdir()
    (tmp_path / "calib").mkdir()
    g["raw"].tofile(str(tmp_path / "velodyne" / "000007.bin"))
    rows = {"P0": np.zeros(12), "P2": g["p2"].reshape(-1), "R0_rect": g["r0_rect"].reshape(-1),
            "Tr_velo_to_cam": g["tr_velodyne_to_cam"].reshape(-1)}
    with open(str(tmp_path / "calib" / "000007.txt"), "w") as f:
        for k, v in rows.items():
            f.write("%s: %s\n" % (k, " ".join(repr(float(x)) for x in v)))
        f.write("\n")
    pc = obj_utils.get_lidar_point_cloud(7, str(tmp_path / "calib"), str(tmp_path / "velodyne"), g["image_size"], cuda)
    assert pc.shape == (12045, 4) and close_f32(pc.cpu().numpy(), g["expect"])
