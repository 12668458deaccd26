"""GPU: the fused sampling entry points — farthest_point_sample + gather_point in one launch
(hf_farthest_point_sample_xyz) and several set-abstraction sampling levels in one launch (hf_fps_pyramid) —
against the oracle (sampling/tf_sampling_g.cu:105-181 restated), the separate ops, and the unmodified reference
kernels.  Reference call order: hf/core/feature_extractors/pointnet_util.py:42-44, once per SA module."""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import _lib, synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import (farthest_point_sample, farthest_point_sample_pyramid,
                                                        farthest_point_sample_with_xyz, gather_point)

import _refcuda

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not _refcuda.available(), reason="oracle/_ref/libref_cuda.so not built")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def N(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("b,n,m,dup,cluster", [(2, 16384, 1024, 0, 0), (3, 4096, 1024, 500, 0), (2, 1000, 256, 0, 0),
                                               (4, 256, 64, 0, 0), (2, 5000, 300, 0, 8), (1, 3, 7, 0, 0),
                                               (2, 700, 700, 0, 2), (2, 16384, 600, 4000, 4), (17, 2048, 64, 0, 0),
                                               (2, 40000, 300, 3000, 0), (1, 70000, 64, 0, 0)])
def test_fps_with_xyz_matches_oracle(cuda, oracle, b, n, m, dup, cluster):
    """indices == oracle FPS, points == oracle gather of them (bit-exact copies), every kernel family behind the
    entry (multi-pick writes the points itself; > 65536 points goes through the global-memory kernel + a gather)."""
    xyz, _ = synthetic.make_scans(101 + n + m, b, n, duplicate_pad=dup)
    want = oracle.farthest_point_sample(m, xyz)
    _lib.load().hf_fps_set_cluster(cluster)
    try:
        idx, pts = farthest_point_sample_with_xyz(m, T(xyz, cuda))
    finally:
        _lib.load().hf_fps_set_cluster(0)
    assert idx.dtype == torch.int32 and tuple(idx.shape) == (b, m) and tuple(pts.shape) == (b, m, 3)
    assert np.array_equal(N(idx), want)
    assert np.array_equal(N(pts), oracle.gather_point(xyz, want))


@pytest.mark.parametrize("algo", [1, 2])
def test_fps_with_xyz_other_kernels(cuda, algo):
    """The one-pick and the bucket-pruned kernels behind the same entry: a gather launch follows them."""
    xyz, _ = synthetic.make_scans(7, 2, 6000, duplicate_pad=700)
    x = T(xyz, cuda)
    lib = _lib.load()
    want = farthest_point_sample(300, x)
    lib.hf_fps_set_variant(algo, 0, 0)
    try:
        idx, pts = farthest_point_sample_with_xyz(300, x)
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    assert torch.equal(idx, want) and torch.equal(pts, gather_point(x, want))


def test_fps_cluster_of_six(cuda, oracle):
    """6 CTAs x 8 warps (hf_fps_set_cluster(6)): 11 / 3 points per thread, the tail of the last 512-block padded."""
    lib = _lib.load()
    for n, m, dup in [(16384, 600, 3000), (16000, 300, 0), (4608, 256, 100), (4000, 128, 0), (300, 64, 0)]:
        xyz, _ = synthetic.make_scans(61 + n, 2, n, duplicate_pad=dup)
        want = oracle.farthest_point_sample(m, xyz)
        lib.hf_fps_set_cluster(6)
        try:
            got = farthest_point_sample(m, T(xyz, cuda))
        finally:
            lib.hf_fps_set_cluster(0)
        assert np.array_equal(N(got), want), (n, m)


@pytest.mark.parametrize("b,n,m,dup", [(2, 65536, 700, 9000), (1, 20000, 300, 0), (2, 40000, 512, 3000), (1, 16385, 64, 0)])
def test_fps_shared_memory_shape(cuda, oracle, b, n, m, dup):
    """Clouds of 16385..65536 points with the coordinates in shared memory (4 CTAs x 16 warps, the shape batches of more
    than 15 scans take): forced here for small batches, against the oracle."""
    xyz, _ = synthetic.make_scans(401 + n + m, b, n, duplicate_pad=dup)
    want = oracle.farthest_point_sample(m, xyz)
    lib = _lib.load()
    lib.hf_fps_set_variant(3, 16, 0)
    try:
        idx, pts = farthest_point_sample_with_xyz(m, T(xyz, cuda))
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    assert np.array_equal(N(idx), want)
    assert np.array_equal(N(pts), oracle.gather_point(xyz, want))


@pytest.mark.parametrize("b,n,m,dup,warps", [(2, 16384, 1024, 4000, 0), (16, 16384, 200, 0, 0), (3, 5000, 300, 100, 0),
                                             (2, 4097, 200, 0, 0), (2, 16000, 500, 0, 4), (2, 16384, 300, 2000, 8), (2, 16384, 400, 3000, 6), (3, 9000, 256, 0, 6),
                                             (2, 40000, 300, 3000, 0), (1, 65536, 256, 8000, 0), (2, 3000, 64, 0, 0),
                                             (1, 33000, 2000, 0, 0)])
def test_fps_over_a_spatial_index(cuda, oracle, b, n, m, dup, warps):
    """index= : the cloud is read in curve order and threads (or warps) skip the centres that cannot reach their part
    of it.  Same picks
    (the tie-break runs on the original indices): against the oracle, duplicate-padded clouds included, every shape
    (8 x 4 and 4 x 8 warps at <= 16384 points, 8 x 8 beyond, clouds indexed in parts, the fallback below 4097 points)."""
    from heterofusionrcnn_b200 import spatial

    xyz, _ = synthetic.make_scans(501 + n + m, b, n, duplicate_pad=dup)
    x = T(xyz, cuda)
    ix = spatial.build_cloud_index(x)
    want = oracle.farthest_point_sample(m, xyz)
    lib = _lib.load()
    for mode in (0, 1, 2):  # default (= compact threads) / compact warps / compact threads
        if warps or mode:
            lib.hf_fps_set_variant(3, warps, mode)
        try:
            idx, pts = farthest_point_sample_with_xyz(m, x, index=ix)
        finally:
            lib.hf_fps_set_variant(0, 0, 0)
        assert np.array_equal(N(idx), want), mode
        assert np.array_equal(N(pts), oracle.gather_point(xyz, want)), mode


@pytest.mark.parametrize("b,n,m", [(16, 16384, 4096), (4, 65536, 16384), (15, 16384, 4096)])
def test_fps_over_a_spatial_index_full_size(cuda, b, n, m):
    """BASELINE config 2 / config 5 sizes: identical to the kernel without the index (itself pinned on the reference
    kernel at these sizes, test_gpu_bench_config.py)."""
    from heterofusionrcnn_b200 import spatial

    xyz, _ = synthetic.make_scans(79, b, n, duplicate_pad=n // 8)
    x = T(xyz, cuda)
    i0, p0 = farthest_point_sample_with_xyz(m, x)
    i1, p1 = farthest_point_sample_with_xyz(m, x, index=spatial.build_cloud_index(x))
    assert torch.equal(i0, i1) and torch.equal(p0, p1)


def _level_by_level(x, ms):
    idxs, pts = [], []
    for m in ms:
        i = farthest_point_sample(m, x)
        x = gather_point(x, i)
        idxs.append(i)
        pts.append(x)
    return idxs, pts


@pytest.mark.parametrize("b,n,ms,dup", [
    (16, 4096, [1024, 256, 64], 0),      # the RPN chain's levels 2-4 at the benchmark batch
    (2, 4096, [1024, 256, 64], 600),     # duplicate points: exact distance ties at every level
    (3, 4000, [1000, 250, 60], 0),       # ragged
    (2, 2048, [512, 128, 32], 0), (2, 1024, [256, 64], 0), (2, 512, [128, 32, 8], 0), (1, 300, [100], 0),
    (2, 100, [50, 20, 5], 0), (1, 1, [1, 1, 1], 0),
    (2, 16384, [4096, 1024, 256], 2000),  # config 5's levels 2-4, cluster 8 x 4 warps
    (16, 16384, [4096, 1024, 256], 0),    # the same at 16 scans: cluster 4 x 8 warps
    (2, 4096, [3000, 100], 0),            # no single-launch shape: issued level by level inside the entry
    (2, 20000, [2000, 500], 0),           # cloud beyond the pyramid kernel: level by level
    (2, 4096, [4096, 4096, 4096], 0),     # m == n at every level
])
def test_fps_pyramid_equals_level_by_level(cuda, b, n, ms, dup):
    xyz, _ = synthetic.make_scans(301 + n + sum(ms), b, n, duplicate_pad=dup)
    x = T(xyz, cuda)
    want_i, want_p = _level_by_level(x, ms)
    got_i, got_p = farthest_point_sample_pyramid(ms, x)
    for j in range(len(ms)):
        assert torch.equal(got_i[j], want_i[j]), "indices of level %d" % j
        assert torch.equal(got_p[j], want_p[j]), "points of level %d" % j


def test_fps_pyramid_matches_oracle(cuda, oracle):
    xyz, _ = synthetic.make_scans(17, 2, 4096, duplicate_pad=300)
    got_i, got_p = farthest_point_sample_pyramid([1024, 256, 64], T(xyz, cuda))
    cloud = xyz
    for j, m in enumerate([1024, 256, 64]):
        want = oracle.farthest_point_sample(m, cloud)
        assert np.array_equal(N(got_i[j]), want), j
        cloud = oracle.gather_point(cloud, want)
        assert np.array_equal(N(got_p[j]), cloud), j


@needs_ref
def test_fps_spine_matches_reference_kernels_at_bench_shape(cuda):
    """16 x (16384 -> 4096 -> 1024 -> 256 -> 64) exactly as chain.run_chain issues it (one launch for level 1, one for
    levels 2-4), against the unmodified reference FPS + gather kernels level by level."""
    xyz, _ = synthetic.make_scans(2, 16, 16384)
    x = T(xyz, cuda)
    i1, p1 = farthest_point_sample_with_xyz(4096, x)
    idxs, pts = farthest_point_sample_pyramid([1024, 256, 64], p1)
    cloud = x
    for got_i, got_p, m in zip([i1] + idxs, [p1] + pts, [4096, 1024, 256, 64]):
        ri = _refcuda.fps(m, cloud)
        assert torch.equal(got_i, ri), m
        cloud = _refcuda.gather_point(cloud, ri)
        assert torch.equal(got_p, cloud), m


def test_fps_pyramid_argument_errors(cuda):
    from heterofusionrcnn_b200._tensors import InvalidArgumentError

    x = torch.zeros((1, 64, 3), device=cuda)
    with pytest.raises(InvalidArgumentError):
        farthest_point_sample_pyramid([], x)
    with pytest.raises(InvalidArgumentError):
        farthest_point_sample_pyramid([8, 4, 2, 1], x)
    with pytest.raises(InvalidArgumentError):
        farthest_point_sample_pyramid([8, 0], x)
