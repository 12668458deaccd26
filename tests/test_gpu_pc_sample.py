"""GPU: hf_pc_sample_points (csrc/pc_sample.cu) — the fixed-size sampling of the KITTI loader
(hf/datasets/kitti/kitti_dataset.py:341-371) — against tests/golden/pc_sample_ref.npz (the reference's own lines
executed with np.random bound to the host-input definitions, tests/golden/make_golden_pc_sample.py) and against the
oracle restatement on seeded scans.  Bar: bit-exact rows, source indices and status."""
import os

import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import kitti_sampling

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pc_sample_ref.npz")


def T(a, dev):
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).to(dev)


def N(t):
    return t.detach().cpu().numpy()


def golden_cases():
    g = np.load(GOLDEN)
    names = sorted({k.split(".")[0] for k in g.files}, key=lambda c: int(c[1:]))
    return [{k.split(".", 1)[1]: g[k] for k in g.files if k.startswith(c + ".")} for c in names]


def test_pc_sample_matches_the_reference_block(cuda):
    for ci, c in enumerate(golden_cases()):
        pts = np.concatenate([c["pts"], c["inten"]], axis=1).astype(np.float32)[None]
        P = int(c["P"])
        out, status, choice = kitti_sampling.sample_point_cloud(
            T(pts, cuda), T(np.array([pts.shape[1]], np.int32), cuda), P, T(c["keys"][None], cuda), T(c["slot_of"][None], cuda),
            T(c["extra"][None], cuda), return_choice=True)
        assert int(status[0]) == int(c["ok"]), ci
        if int(c["ok"]):
            assert np.array_equal(N(choice[0]), c["choice"]), ci
            assert np.array_equal(N(out[0]), c["sampled"].astype(np.float32)), ci
        else:
            assert not N(out).any() and (N(choice) == -1).all()


@pytest.mark.parametrize("S,n_cap,P,seed", [(4, 60000, 16384, 1), (3, 20000, 16384, 2), (2, 131072, 16384, 3),
                                            (5, 5000, 4096, 4), (2, 100, 300, 5)])
def test_pc_sample_matches_oracle_batched(cuda, oracle, S, n_cap, P, seed):
    """Ragged batches: every scan has its own count (read on the device), incl. an empty one and a full one."""
    rng = np.random.default_rng(seed)
    z = np.where(rng.random((S, n_cap)) < 0.15, rng.uniform(40, 75, (S, n_cap)), rng.uniform(0.1, 40, (S, n_cap)))
    pts = np.stack([rng.uniform(-40, 40, (S, n_cap)), rng.uniform(-3, 3, (S, n_cap)), z, rng.random((S, n_cap))],
                   axis=2).astype(np.float32)
    counts = rng.integers(1, n_cap + 1, S).astype(np.int32)
    counts[0] = n_cap
    if S > 2:
        counts[1] = 0
        counts[2] = max(1, P // 3)  # draw with replacement
    keys, slot_of, extra = kitti_sampling.draw_randomness(rng, S, n_cap, P)
    if seed == 4:
        keys %= 16  # many equal keys: the lower index wins
    out, status, choice = kitti_sampling.sample_point_cloud(T(pts, cuda), T(counts, cuda), P, T(keys, cuda), T(slot_of, cuda),
                                                            T(extra, cuda), return_choice=True)
    for s in range(S):
        want, wchoice, ok = oracle.pc_sample_points(pts[s], counts[s], P, keys[s], slot_of[s], extra[s])
        assert int(status[s]) == ok, s
        assert np.array_equal(N(choice[s]), wchoice), s
        assert np.array_equal(N(out[s]), want), s
        if ok and counts[s] > P:  # every far point is kept, no point twice
            far = np.nonzero(~(pts[s, :counts[s], 2] < 40.0))[0]
            assert np.isin(far, wchoice).all() and len(set(wchoice.tolist())) == P


def test_pc_sample_behind_the_fov_filter(cuda, oracle):
    """The two input-side ops chained on the device: raw velodyne records -> hf_lidar_fov_filter -> hf_pc_sample_points,
    the counts never leave the GPU; the result is the oracle's for the filter's output."""
    from heterofusionrcnn_b200 import obj_utils

    g = np.load(os.path.join(os.path.dirname(GOLDEN), "lidar_fov_000000.npz"))
    raw = torch.from_numpy(g["raw"][None]).to(cuda)
    calib = obj_utils.pack_calib(g["r0_rect"], g["tr_velodyne_to_cam"], g["p2"])[None]
    kept, counts = obj_utils.lidar_fov_filter(raw, calib, [int(v) for v in g["image_size"]])
    rng = np.random.default_rng(9)
    P = 8192
    keys, slot_of, extra = kitti_sampling.draw_randomness(rng, 1, kept.shape[1], P)
    out, status = kitti_sampling.sample_point_cloud(kept, counts, P, T(keys, cuda), T(slot_of, cuda), T(extra, cuda))
    want, _, ok = oracle.pc_sample_points(N(kept[0]), int(counts[0]), P, keys[0], slot_of[0], extra[0])
    assert ok == 1 and int(status[0]) == 1 and int(counts[0]) > P
    assert np.array_equal(N(out[0]), want)


def test_pc_sample_argument_errors(cuda):
    from heterofusionrcnn_b200._lib import InvalidArgumentError

    pts = torch.zeros((1, 8, 4), device=cuda)
    cnt = torch.full((1,), 8, dtype=torch.int32, device=cuda)
    keys = torch.zeros((1, 8), dtype=torch.int32, device=cuda)
    slot = torch.arange(4, dtype=torch.int32, device=cuda)[None]
    with pytest.raises(InvalidArgumentError):
        kitti_sampling.sample_point_cloud(pts[:, :, :3], cnt, 4, keys, slot)
    with pytest.raises(InvalidArgumentError):
        kitti_sampling.sample_point_cloud(pts, cnt, 4, keys[:, :5], slot)
    with pytest.raises(InvalidArgumentError):
        kitti_sampling.sample_point_cloud(pts, cnt, 5, keys, slot)


@pytest.mark.parametrize("n,P", [(600, 256), (100, 256)])
def test_pc_sample_reports_a_slot_outside_the_output(cuda, n, P):
    """slot_of must be a permutation of range(P).  An entry outside [0, P) cannot be honoured: the scan is reported with
    status 0 and zero rows (include/hfops.h), the other scans of the batch are unaffected."""
    rng = np.random.default_rng(n)
    pts = rng.uniform(0.1, 39, (2, n, 4)).astype(np.float32)
    counts = np.array([n, n], np.int32)
    keys, slot_of, extra = kitti_sampling.draw_randomness(rng, 2, n, P)
    good, gstatus = kitti_sampling.sample_point_cloud(T(pts, cuda), T(counts, cuda), P, T(keys, cuda), T(slot_of, cuda),
                                                      T(extra, cuda))
    assert N(gstatus).tolist() == [1, 1]
    bad = slot_of.copy()
    bad[1, 7] = P  # one past the end
    bad[1, 100] = -3
    out, status, choice = kitti_sampling.sample_point_cloud(T(pts, cuda), T(counts, cuda), P, T(keys, cuda), T(bad, cuda),
                                                            T(extra, cuda), return_choice=True)
    assert N(status).tolist() == [1, 0]
    assert torch.equal(out[0], good[0])
    assert not N(out[1]).any() and (N(choice[1]) == -1).all()
