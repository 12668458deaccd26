"""GPU: the CUDA path (through the Python mirror -> ctypes -> C-ABI) against the CPU oracle
and, where oracle/_ref/libref_cuda.so travelled with the tree, against the UNMODIFIED
reference CUDA kernels running on the same B200.

Bar: bit-exact for every integer / index / bool output; floats within 1e-5 relative
(BASELINE.json north_star), and in practice bit-exact where the arithmetic order is pinned.
"""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.bev_iou.bev_iou import compute_bev_iou, oriented_nms, oriented_nms_batch, oriented_nms_mask
from heterofusionrcnn_b200.cropping.tf_cropping import pc_crop_and_sample
from heterofusionrcnn_b200.grouping.tf_grouping import group_point, query_ball_point
from heterofusionrcnn_b200.interpolate.tf_interpolate import (three_interpolate, three_interpolate_channel_first,
                                                              three_nn)
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point

import _refcuda

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # north_star float tolerance


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def N(t):
    return t.detach().cpu().numpy()


def close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = np.maximum(np.abs(a), np.abs(b))
    return np.all(np.abs(a - b) <= rtol * scale + 1e-12)


needs_ref = pytest.mark.skipif(not _refcuda.available(), reason="oracle/_ref/libref_cuda.so not built")


# ---------------------------------------------------------------- FPS / gather
@pytest.mark.parametrize("b,n,m,dup", [(2, 16384, 1024, 0), (3, 4096, 1024, 0), (2, 1024, 256, 0), (4, 256, 64, 0),
                                       (2, 5000, 300, 0), (2, 700, 700, 0), (1, 3, 7, 0), (2, 16384, 600, 4000),
                                       (1, 8192, 64, 8000)])
def test_fps_matches_oracle(cuda, oracle, b, n, m, dup):
    xyz, _ = synthetic.make_scans(11 + n + m, b, n, duplicate_pad=dup)
    want = oracle.farthest_point_sample(m, xyz)
    got = N(farthest_point_sample(m, T(xyz, cuda)))
    assert got.dtype == np.int32 and got.shape == (b, m)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("cluster", [1, 2, 4, 8])
def test_fps_every_cluster_shape(cuda, oracle, cluster):
    from heterofusionrcnn_b200 import _lib

    xyz, _ = synthetic.make_scans(5, 2, 6000, duplicate_pad=1500)
    want = oracle.farthest_point_sample(512, xyz)
    _lib.load().hf_fps_set_cluster(cluster)
    try:
        got = N(farthest_point_sample(512, T(xyz, cuda)))
    finally:
        _lib.load().hf_fps_set_cluster(0)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n,m,dup,w,j", [(16384, 700, 3000, 16, 32), (16000, 300, 0, 32, 16), (8192, 300, 0, 8, 32),
                                         (8000, 300, 500, 16, 16), (7000, 200, 0, 32, 8), (4096, 512, 0, 4, 32),
                                         (4096, 512, 1000, 8, 16), (4000, 256, 0, 16, 8), (3000, 256, 0, 32, 4),
                                         (2048, 256, 0, 4, 16), (2000, 256, 0, 8, 8), (1500, 200, 100, 16, 4),
                                         (2048, 128, 0, 32, 2), (1024, 256, 0, 4, 8), (1024, 256, 0, 8, 4),
                                         (1000, 128, 0, 16, 2), (512, 128, 0, 4, 4), (500, 64, 0, 8, 2),
                                         (300, 64, 0, 16, 1), (256, 64, 0, 2, 4), (200, 300, 0, 4, 2), (37, 16, 0, 8, 1),
                                         (1, 4, 0, 2, 4), (20000, 300, 0, 16, 32), (40000, 300, 3000, 16, 32),
                                         (30000, 256, 0, 32, 16), (60000, 256, 0, 32, 16), (100000, 256, 0, 16, 32),
                                         (131072, 200, 9000, 32, 16), (12000, 256, 0, 16, 16), (30000, 256, 0, 16, 16),
                                         (65536, 256, 8000, 16, 16), (33000, 200, 0, 0, 0)])
def test_fps_bucket_kernel_shapes(cuda, oracle, n, m, dup, w, j):
    """Every instantiated shape of the bucket-pruned kernel (fps_bucket.cu), incl. clusters, ragged
    sizes, duplicate-padded clouds (exact ties) and m > n, against the oracle."""
    from heterofusionrcnn_b200 import _lib

    xyz, _ = synthetic.make_scans(21 + n + m, 2, n, duplicate_pad=dup)
    want = oracle.farthest_point_sample(m, xyz)
    _lib.load().hf_fps_set_variant(2, w, j)
    try:
        got = N(farthest_point_sample(m, T(xyz, cuda)))
    finally:
        _lib.load().hf_fps_set_variant(0, 0, 0)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n,m", [(16384, 4096), (65536, 16384)])
def test_fps_bucket_equals_plain_kernel_full_size(cuda, n, m):
    """BASELINE config 2 / config 5 sizes: the pruned kernel and the plain every-point kernel
    (itself pinned against the oracle and the reference kernel) give identical indices."""
    from heterofusionrcnn_b200 import _lib

    xyz, _ = synthetic.make_scans(77, 2, n, duplicate_pad=n // 8)
    x = T(xyz, cuda)
    lib = _lib.load()
    try:
        lib.hf_fps_set_variant(1, 0, 0)
        plain = farthest_point_sample(m, x)
        lib.hf_fps_set_variant(2, 0, 0)
        bucket = farthest_point_sample(m, x)
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    assert torch.equal(plain, bucket)


@pytest.mark.parametrize("n,m,dup,cluster", [(16384, 700, 3000, 0), (16384, 300, 0, 4), (16000, 300, 0, 2), (8192, 300, 0, 1),
                                             (8000, 300, 500, 8), (7000, 200, 0, 2), (4096, 512, 0, 0), (4096, 512, 1000, 1),
                                             (4000, 256, 0, 2), (3000, 256, 0, 4), (2048, 256, 0, 0), (2000, 256, 0, 1),
                                             (1500, 200, 100, 8), (1024, 256, 0, 0), (1000, 128, 0, 1), (512, 128, 0, 0),
                                             (500, 64, 0, 8), (300, 64, 0, 2), (256, 64, 0, 0), (200, 300, 0, 0),
                                             (37, 16, 0, 4), (1, 4, 0, 0), (3, 7, 0, 2), (20000, 300, 0, 0),
                                             (40000, 300, 3000, 0), (30000, 256, 0, 4), (60000, 256, 0, 0),
                                             (65536, 256, 8000, 8), (33000, 200, 0, 0), (700, 700, 0, 0), (5000, 1, 0, 0)])
def test_fps_multi_pick_kernel_shapes(cuda, oracle, n, m, dup, cluster):
    """Every instantiated shape of the multi-pick kernel (fps_multi.cu), incl. every cluster size,
    ragged sizes, duplicate-padded clouds (exact ties) and m > n, against the oracle."""
    from heterofusionrcnn_b200 import _lib

    xyz, _ = synthetic.make_scans(31 + n + m, 2, n, duplicate_pad=dup)
    want = oracle.farthest_point_sample(m, xyz)
    _lib.load().hf_fps_set_variant(3, 0, 0)
    _lib.load().hf_fps_set_cluster(cluster)
    try:
        got = N(farthest_point_sample(m, T(xyz, cuda)))
    finally:
        _lib.load().hf_fps_set_cluster(0)
        _lib.load().hf_fps_set_variant(0, 0, 0)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n,m,b", [(16384, 4096, 16), (65536, 16384, 3), (4096, 4096, 20)])
def test_fps_multi_pick_equals_plain_kernel_full_size(cuda, n, m, b):
    """BASELINE config 2 / config 5 sizes: the multi-pick kernel and the one-pick-per-exchange
    kernel (itself pinned against the oracle and the reference kernel) give identical indices."""
    from heterofusionrcnn_b200 import _lib

    xyz, _ = synthetic.make_scans(78, b, n, duplicate_pad=n // 8)
    x = T(xyz, cuda)
    lib = _lib.load()
    try:
        lib.hf_fps_set_variant(1, 0, 0)
        plain = farthest_point_sample(m, x)
        lib.hf_fps_set_variant(3, 0, 0)
        multi = farthest_point_sample(m, x)
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    assert torch.equal(plain, multi)


@needs_ref
def test_fps_matches_reference_kernel_full_size(cuda):
    """BASELINE config 2 size, against the reference's own kernel on this GPU (ties included)."""
    xyz, _ = synthetic.make_scans(2, 4, 16384, duplicate_pad=2000)
    x = T(xyz, cuda)
    assert torch.equal(farthest_point_sample(4096, x), _refcuda.fps(4096, x))


def test_fps_65536_properties(cuda):
    """BASELINE config 5 size (too slow for the serial oracle): size-independent properties."""
    xyz, _ = synthetic.make_scans(9, 2, 65536)
    x = T(xyz, cuda)
    idx = farthest_point_sample(2048, x)
    i = N(idx)
    assert i[:, 0].tolist() == [0, 0] and i.min() >= 0 and i.max() < 65536
    for s in range(2):
        assert len(set(i[s].tolist())) == 2048  # distinct points => no repeats
    # greedy property: the min-distance to the chosen set is non-increasing along the sequence
    p = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3))
    for s in range(2):
        d = torch.cdist(p[s, :512].double(), p[s, :512].double())
        prev = torch.full((512,), float("inf"), dtype=torch.float64, device=cuda)
        mins = []
        for j in range(1, 512):
            mins.append(d[j, :j].min().item())
        assert all(mins[k] >= mins[k + 1] - 1e-6 for k in range(len(mins) - 1))


def test_gather_point(cuda, oracle):
    rng = np.random.default_rng(3)
    xyz = rng.standard_normal((3, 777, 3)).astype(np.float32)
    idx = rng.integers(0, 777, (3, 129)).astype(np.int32)
    assert np.array_equal(N(gather_point(T(xyz, cuda), T(idx, cuda))), oracle.gather_point(xyz, idx))


# ---------------------------------------------------------------- ball query / group
@pytest.mark.parametrize("b,n,m,r,ns", [(2, 16384, 512, 0.8, 32), (2, 4096, 1024, 0.2, 32), (2, 1024, 256, 0.4, 32),
                                        (3, 1001, 130, 2.0, 16), (1, 50, 7, 100.0, 64), (2, 4096, 100, 0.05, 8),
                                        (1, 2048, 64, 4.0, 400)])
def test_query_ball_point_matches_oracle(cuda, oracle, b, n, m, r, ns):
    xyz, _ = synthetic.make_scans(n + m, b, n)
    q = xyz[:, :m] if m % 2 else oracle.gather_point(xyz, oracle.farthest_point_sample(m, xyz))
    q = np.ascontiguousarray(q)
    widx, wcnt = oracle.query_ball_point(r, ns, xyz, q)
    gidx, gcnt = query_ball_point(r, ns, T(xyz, cuda), T(q, cuda))
    assert np.array_equal(N(gcnt), wcnt)
    assert np.array_equal(N(gidx), widx)  # rows without a hit are 0 on both sides


def test_query_ball_point_radius_edge_is_bit_exact(cuda, oracle):
    """Points sitting exactly on / one ulp around the radius: the sqrt-free test must agree."""
    rng = np.random.default_rng(0)
    r = np.float32(0.3)
    n = 4096
    dirs = rng.standard_normal((n, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    scale = r * (1 + rng.integers(-4, 5, n) * 2.0 ** -23)
    xyz = (dirs * scale[:, None]).astype(np.float32)[None]
    q = np.zeros((1, 1, 3), np.float32)
    widx, wcnt = oracle.query_ball_point(float(r), n, xyz, q)
    gidx, gcnt = query_ball_point(float(r), n, T(xyz, cuda), T(q, cuda))
    assert 0 < wcnt[0, 0] < n
    assert np.array_equal(N(gcnt), wcnt) and np.array_equal(N(gidx), widx)


@needs_ref
def test_query_ball_point_matches_reference_kernel(cuda):
    xyz, _ = synthetic.make_scans(21, 2, 16384)
    x = T(xyz, cuda)
    q = gather_point(x, farthest_point_sample(1024, x))
    for r in (0.1, 0.8, 4.0):
        gidx, gcnt = query_ball_point(r, 32, x, q)
        ridx, rcnt = _refcuda.query_ball_point(r, 32, x, q)
        assert torch.equal(gcnt, rcnt)
        hit = rcnt > 0  # the reference leaves rows without a hit uninitialised
        assert torch.equal(gidx[hit], ridx[hit])


# ---------------------------------------------------------------- spatially pruned ball query / three_nn
@pytest.fixture
def pruned(cuda):
    """Force the pruned kernels of csrc/spatial.cu on every shape they have (default: large calls only)."""
    from heterofusionrcnn_b200 import _lib

    _lib.load().hf_spatial_set_mode(2)
    yield
    _lib.load().hf_spatial_set_mode(0)


@pytest.mark.parametrize("b,n,m,r,ns,dup", [(2, 4096, 1024, 0.2, 32, 0), (2, 5000, 777, 0.4, 16, 600), (1, 16384, 300, 0.1, 32, 0),
                                            (2, 3000, 3000, 2.0, 32, 0), (1, 2048, 512, 50.0, 32, 500), (2, 100, 33, 1.0, 8, 0),
                                            (1, 31, 1, 0.5, 4, 0), (1, 1, 5, 0.5, 3, 0), (2, 1500, 200, 4.0, 64, 0),
                                            (1, 6000, 128, 8.0, 128, 1000), (1, 700, 64, 1e-6, 32, 0),
                                            (1, 20000, 900, 0.5, 32, 3000), (1, 65536, 4096, 0.4, 32, 0),
                                            (1, 40000, 20000, 1.0, 16, 0)])
def test_query_ball_point_pruned_matches_oracle(cuda, oracle, pruned, b, n, m, r, ns, dup):
    """Ragged sizes, duplicate points, rows with no hit, rows with far more hits than nsample
    (the nsample smallest indices must survive), every query inside every ball (r = 50)."""
    xyz, _ = synthetic.make_scans(n + m + 3, b, n, duplicate_pad=dup)
    q = (xyz[:, ::-1][:, :m] if m <= n else np.resize(xyz, (b, m, 3))).copy()
    widx, wcnt = oracle.query_ball_point(r, ns, xyz, q)
    gidx, gcnt = query_ball_point(r, ns, T(xyz, cuda), T(q, cuda))
    assert np.array_equal(N(gcnt), wcnt)
    assert np.array_equal(N(gidx), widx)


@pytest.mark.parametrize("b,n,m,dup", [(2, 4096, 1024, 0), (2, 5000, 777, 600), (1, 16384, 300, 2000), (2, 3000, 3000, 0),
                                       (2, 100, 33, 10), (1, 31, 1, 0), (1, 5, 2, 0), (1, 1, 700, 0), (2, 1500, 3, 0),
                                       (1, 16384, 16384, 4000), (1, 20000, 900, 3000), (1, 65536, 4096, 0),
                                       (1, 33000, 40000, 5000)])
def test_three_nn_pruned_matches_oracle(cuda, oracle, pruned, b, n, m, dup):
    """Incl. exact distance ties (duplicate points: the lower index must win) and m < 3 (+inf, index 0)."""
    xyz, _ = synthetic.make_scans(n * 3 + m + 1, b, max(n, m), duplicate_pad=min(dup, max(n, m) - 1))
    unknown = np.ascontiguousarray(xyz[:, :n])
    known = xyz[:, ::-1][:, :m].copy()
    wd, wi = oracle.three_nn(unknown, known)
    gd, gi = three_nn(T(unknown, cuda), T(known, cuda))
    assert np.array_equal(N(gi), wi)
    assert np.array_equal(N(gd), wd)


def test_pruned_kernels_equal_every_pair_kernels_on_the_chain_shapes(cuda):
    """BASELINE config 2 shapes, both radius sets: pruned == every-pair, bit for bit."""
    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()
    xyz, _ = synthetic.make_scans(91, 3, 16384, duplicate_pad=1500)
    x0 = T(xyz, cuda)
    x1 = gather_point(x0, farthest_point_sample(4096, x0))
    x2 = gather_point(x1, farthest_point_sample(1024, x1))
    try:
        for (a, q, radii) in [(x0, x1, (0.1, 0.5, 4.0)), (x1, x2, (0.2, 1.0))]:
            lib.hf_spatial_set_mode(1)
            want_nn = three_nn(a, q)
            want_bq = [query_ball_point(r, 32, a, q) for r in radii]
            lib.hf_spatial_set_mode(2)
            got_nn = three_nn(a, q)
            got_bq = [query_ball_point(r, 32, a, q) for r in radii]
            assert torch.equal(got_nn[0], want_nn[0]) and torch.equal(got_nn[1], want_nn[1])
            for g, w in zip(got_bq, want_bq):
                assert torch.equal(g[0], w[0]) and torch.equal(g[1], w[1])
    finally:
        lib.hf_spatial_set_mode(0)


def test_pruned_kernels_equal_every_pair_kernels_at_config5_level1(cuda):
    """BASELINE config 5, level 1 (65536 -> 16384): clouds indexed in four parts; pruned == every-pair."""
    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()
    xyz, _ = synthetic.make_scans(17, 2, 65536, duplicate_pad=3000)
    x0 = T(xyz, cuda)
    x1 = gather_point(x0, farthest_point_sample(16384, x0))
    try:
        lib.hf_spatial_set_mode(1)
        want_nn, want_bq = three_nn(x0, x1), [query_ball_point(r, 32, x0, x1) for r in (0.1, 0.5)]
        lib.hf_spatial_set_mode(0)  # auto: these shapes take the pruned kernels
        got_nn, got_bq = three_nn(x0, x1), [query_ball_point(r, 32, x0, x1) for r in (0.1, 0.5)]
    finally:
        lib.hf_spatial_set_mode(0)
    assert torch.equal(got_nn[0], want_nn[0]) and torch.equal(got_nn[1], want_nn[1])
    for g, w in zip(got_bq, want_bq):
        assert torch.equal(g[0], w[0]) and torch.equal(g[1], w[1])


@pytest.mark.parametrize("c", [1, 3, 4, 96, 130, 256])
def test_group_point(cuda, oracle, c):
    rng = np.random.default_rng(c)
    pts = rng.standard_normal((2, 500, c)).astype(np.float32)
    idx = rng.integers(0, 500, (2, 37, 9)).astype(np.int32)
    got = group_point(T(pts, cuda), T(idx, cuda))
    assert got.shape == (2, 37, 9, c)
    assert np.array_equal(N(got), oracle.group_point(pts, idx))


@pytest.mark.parametrize("c", [1, 2, 3, 4, 7, 96, 130, 256, 512])
@pytest.mark.parametrize("b,n,m,ns", [(2, 500, 37, 9), (17, 300, 64, 32), (1, 5, 3, 1)])
def test_group_point_pair(cuda, oracle, c, b, n, m, ns):
    """The coordinates and the features of the same neighbourhoods in one launch (pointnet_util.py:49,56): every kernel
    behind the entry — 128-bit rows with the coordinates on 12 lanes, thread-per-row for c <= 4, two launches otherwise."""
    from heterofusionrcnn_b200.grouping.tf_grouping import group_point_pair

    rng = np.random.default_rng(c + n)
    xyz = rng.standard_normal((b, n, 3)).astype(np.float32)
    pts = rng.standard_normal((b, n, c)).astype(np.float32)
    idx = rng.integers(0, n, (b, m, ns)).astype(np.int32)
    gx, gp = group_point_pair(T(xyz, cuda), T(pts, cuda), T(idx, cuda))
    assert tuple(gx.shape) == (b, m, ns, 3) and tuple(gp.shape) == (b, m, ns, c)
    assert np.array_equal(N(gx), oracle.group_point(xyz, idx))
    assert np.array_equal(N(gp), oracle.group_point(pts, idx))


@pytest.mark.parametrize("c", [4, 96, 256, 512, 1024, 4096, 8192])
def test_group_point_tma_row_mover(cuda, oracle, c):
    """hf_gather_set_mode(1): rows moved by cp.async.bulk (global -> shared per gathered row, one bulk copy shared ->
    global per stage).  Opt-in (slower than the register kernel for these row sizes); rows beyond a stage (16 KB) take
    the register kernel."""
    from heterofusionrcnn_b200 import _lib

    rng = np.random.default_rng(c)
    pts = rng.standard_normal((3, 700, c)).astype(np.float32)
    idx = rng.integers(0, 700, (3, 41, 13)).astype(np.int32)
    _lib.load().hf_gather_set_mode(1)
    try:
        got = group_point(T(pts, cuda), T(idx, cuda))
    finally:
        _lib.load().hf_gather_set_mode(0)
    assert np.array_equal(N(got), oracle.group_point(pts, idx))


@pytest.mark.parametrize("c", [4, 8, 32])
def test_group_point_tma_row_mover_stage_reuse_stress(cuda, c):
    """Many short rows through the bulk row mover: every warp walks its ring of three shared-memory stages hundreds
    of times, so a stage refilled while the previous store still reads it (write-after-read on the stage) would show
    as rows of the wrong source.  Row r of the source holds the value r in every channel, so validity is checked on
    the device against the index itself, over several repetitions."""
    from heterofusionrcnn_b200 import _lib

    b, n, m, ns = 4, 4096, 16384, 32
    g = torch.Generator(device="cpu").manual_seed(c)
    idx = torch.randint(0, n, (b, m, ns), generator=g, dtype=torch.int32).to(cuda)
    pts = (torch.arange(n, dtype=torch.float32, device=cuda)[None, :, None] +
           n * torch.arange(b, dtype=torch.float32, device=cuda)[:, None, None]).expand(b, n, c).contiguous()
    want = (idx.to(torch.float32) + n * torch.arange(b, dtype=torch.float32, device=cuda)[:, None, None])[..., None].expand(b, m, ns, c)
    _lib.load().hf_gather_set_mode(1)
    try:
        for _ in range(5):
            got = group_point(pts, idx)
            assert torch.equal(got, want)
    finally:
        _lib.load().hf_gather_set_mode(0)


# ---------------------------------------------------------------- three_nn / interpolate
@pytest.mark.parametrize("b,n,m", [(2, 16384, 4096), (2, 4096, 1024), (3, 1024, 256), (2, 256, 64), (2, 333, 1027),
                                   (1, 10, 2), (1, 5, 1)])
def test_three_nn_matches_oracle(cuda, oracle, b, n, m):
    xyz, _ = synthetic.make_scans(n * 3 + m, b, max(n, m), duplicate_pad=max(n, m) // 8)
    unknown = np.ascontiguousarray(xyz[:, :n])
    known = xyz[:, ::-1][:, :m].copy()
    wd, wi = oracle.three_nn(unknown, known)
    gd, gi = three_nn(T(unknown, cuda), T(known, cuda))
    assert np.array_equal(N(gi), wi)
    assert np.array_equal(N(gd), wd)  # includes the +inf of m < 3


@needs_ref
def test_three_nn_and_interpolate_match_reference_kernels(cuda):
    xyz, _ = synthetic.make_scans(4, 2, 16384, duplicate_pad=1000)
    x = T(xyz, cuda)
    known = gather_point(x, farthest_point_sample(4096, x))
    gd, gi = three_nn(x, known)
    rd, ri = _refcuda.three_nn(x, known)
    assert torch.equal(gi, ri) and torch.equal(gd, rd)
    w = T(synthetic.fp_weights(N(gd)), cuda)
    feats = torch.randn(2, 4096, 128, device=cuda)
    got = three_interpolate(feats, gi, w)
    ref = _refcuda.three_interpolate_channel_first(feats.transpose(1, 2).contiguous(), gi, w).transpose(1, 2)
    assert torch.equal(got, ref)  # same FMA order => bit-identical, stronger than the 1e-5 bar
    cf = three_interpolate_channel_first(feats.transpose(1, 2).contiguous(), gi, w)
    assert torch.equal(cf.transpose(1, 2), ref)


@pytest.mark.parametrize("c", [1, 3, 128, 132, 1024])
def test_three_interpolate(cuda, oracle, c):
    rng = np.random.default_rng(c)
    pts = rng.standard_normal((2, 300, c)).astype(np.float32)
    idx = rng.integers(0, 300, (2, 701, 3)).astype(np.int32)
    w = rng.random((2, 701, 3)).astype(np.float32)
    w /= w.sum(-1, keepdims=True)
    want = oracle.three_interpolate(pts, idx, w)
    got = N(three_interpolate(T(pts, cuda), T(idx, cuda), T(w, cuda)))
    assert close(got, want) and np.array_equal(got, want)
    cf = N(three_interpolate_channel_first(T(pts.transpose(0, 2, 1), cuda), T(idx, cuda), T(w, cuda)))
    assert np.array_equal(cf.transpose(0, 2, 1), want)


@pytest.mark.parametrize("b,n,m,c", [(2, 16384, 4096, 128), (3, 4096, 1024, 512), (2, 701, 300, 132), (1, 33, 7, 4),
                                     (2, 5000, 1000, 3), (2, 1000, 256, 260), (2, 300, 64, 1024), (1, 1, 5, 8),
                                     (17, 300, 100, 64)])
def test_three_interpolate_on_three_nn_geometry(cuda, oracle, b, n, m, c):
    """Real three_nn indices and weights (neighbouring rows share sources), ragged sizes, more scenes than one CTA row of
    the grid covers, and the reference kernel's own output in its channel-first layout."""
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn_weights

    xyz, _ = synthetic.make_scans(50 + n, b, n)
    unk = T(xyz, cuda)
    kn = gather_point(unk, farthest_point_sample(m, unk))
    _, nidx, w = three_nn_weights(unk, kn)
    pts = torch.randn((b, m, c), device=cuda)
    got = three_interpolate(pts, nidx, w)
    assert np.array_equal(N(got), oracle.three_interpolate(N(pts), N(nidx), N(w)))
    if _refcuda.available():
        ref = _refcuda.three_interpolate_channel_first(pts.transpose(1, 2).contiguous(), nidx, w)
        assert torch.equal(got, ref.transpose(1, 2))


# ---------------------------------------------------------------- crop
def crop_inputs(seed, B, P, C, nbox):
    xyz, inten = synthetic.make_scans(seed, B, P)
    rng = np.random.default_rng(seed)
    fts = rng.standard_normal((B, P, C)).astype(np.float32)
    mask = rng.random((B, P)) < 0.3
    b3 = [synthetic.make_boxes3d(seed * 13 + i, xyz[i], nbox) for i in range(B)]
    b3[0][:3, 3:6] *= 6.0  # a few big boxes: more than `resize` points inside
    b3[-1][0, 0] += 500.0  # and an empty one
    boxes = np.concatenate([synthetic.boxes3d_to_corners(x) for x in b3])
    box_ind = np.repeat(np.arange(B, dtype=np.int32), nbox)
    return xyz, fts, inten, mask, boxes, box_ind


@pytest.mark.parametrize("C,R", [(288, 512), (7, 64), (4, 16)])
def test_crop_matches_oracle(cuda, oracle, C, R):
    args = crop_inputs(5 + C, 2, 16384, C, 40)
    want = oracle.pc_crop_and_sample(*args, R)
    got = pc_crop_and_sample(*[T(a, cuda) for a in args], R)
    names = ["crop_pts", "crop_fts", "crop_intensities", "crop_mask", "crop_ind", "non_empty_box"]
    assert not want[5].all() and want[5].any()
    for name, g, w in zip(names, got, want):
        assert N(g).shape == w.shape, name
        assert np.array_equal(N(g), w), name


@pytest.mark.parametrize("C,R,P,nbox", [(288, 512, 16384, 40), (7, 64, 16384, 40), (4, 16, 16384, 40), (8, 128, 5000, 25),
                                        (4, 32, 40000, 12), (12, 8, 70000, 6), (8, 256, 65536, 48), (4, 64, 16416, 20),
                                        (4, 128, 33000, 30)])
def test_crop_with_cloud_index_matches_oracle(cuda, oracle, C, R, P, nbox):
    """The index only prunes the point-in-box scan: identical outputs, incl. boxes with more hits than the hit
    list holds (R = 16 / 8: falls back to the ordered scan), clouds indexed in parts with the second box level
    (16416 .. 65536 points: only the groups of 32 buckets whose super-box the RoI touches are tested) and clouds too
    large to index (70000: plain path)."""
    from heterofusionrcnn_b200 import spatial

    args = crop_inputs(9 + C, 2, P, C, nbox)
    want = oracle.pc_crop_and_sample(*args, R)
    dargs = [T(a, cuda) for a in args]
    ix = spatial.build_cloud_index(dargs[0])
    assert (ix is None) == (P > 65536)
    got = pc_crop_and_sample(*dargs, R, index=ix)
    assert want[5].any()
    for name, g, w in zip(["crop_pts", "crop_fts", "crop_intensities", "crop_mask", "crop_ind", "non_empty_box"], got, want):
        assert np.array_equal(N(g), w), name


@pytest.mark.parametrize("R,scale", [(3, 1.0), (1, 1.0), (2, 6.0), (100, 6.0), (500, 6.0), (100, 1.0), (37, 3.0)])
def test_crop_with_cloud_index_odd_resize(cuda, oracle, R, scale):
    """resize that is not a power of two / smaller than 4: the indexed select's hit list holds 8*R entries but
    is sorted over the next power of two, so boxes with hcap/2 < hits <= hcap (and any hit at R < 4) must take
    the ordered scan instead (cropping.cu wrote past the list there).  Boxes are blown up so that
    several hold more than R and more than 8*R points."""
    from heterofusionrcnn_b200 import spatial

    B, P, C, nbox = 2, 16384, 8, 24
    xyz, inten = synthetic.make_scans(300 + R, B, P)
    rng = np.random.default_rng(R)
    fts = rng.standard_normal((B, P, C)).astype(np.float32)
    mask = rng.random((B, P)) < 0.3
    b3 = [synthetic.make_boxes3d(R * 13 + i, xyz[i], nbox) for i in range(B)]
    for i in range(B):
        b3[i][:, 3:6] *= np.linspace(0.5, 2.0, nbox)[:, None] * scale  # hit counts sweep through R, 4R, 8R
    b3[-1][0, 0] += 500.0
    boxes = np.concatenate([synthetic.boxes3d_to_corners(x) for x in b3])
    box_ind = np.repeat(np.arange(B, dtype=np.int32), nbox)
    args = (xyz, fts, inten, mask, boxes, box_ind)
    want = oracle.pc_crop_and_sample(*args, R)
    dargs = [T(a, cuda) for a in args]
    ix = spatial.build_cloud_index(dargs[0])
    for index in (ix, None):
        got = pc_crop_and_sample(*dargs, R, index=index)
        for name, g, w in zip(["crop_pts", "crop_fts", "crop_intensities", "crop_mask", "crop_ind", "non_empty_box"], got, want):
            assert np.array_equal(N(g), w), (name, index is not None)
    # the sweep really crosses the list capacity: some boxes hold more than 8*R points, some fewer than R
    inside = np.array([len(np.unique(want[4][k])) for k in range(len(boxes))])
    assert (inside >= min(R, 2)).any() and want[5].any() and not want[5].all()


@needs_ref
def test_crop_agrees_with_reference_kernel_up_to_its_race(cuda):
    """The reference assigns slots with a racing atomicInc: compare what is race-free
    (which boxes are empty, and the SET of cropped points when fewer than R are inside)."""
    args = [T(a, cuda) for a in crop_inputs(77, 2, 16384, 16, 64)]
    R = 512
    got = pc_crop_and_sample(*args, R)
    ref = _refcuda.pc_crop_and_sample(*args, R)
    assert torch.equal(got[5], ref[5])
    gi, ri = N(got[4]), N(ref[4])
    checked = 0
    for b in range(gi.shape[0]):
        u = np.unique(gi[b])
        if N(got[5])[b] and len(u) < R:
            assert np.array_equal(u, np.unique(ri[b]))
            checked += 1
    assert checked > 20


# ---------------------------------------------------------------- BEV IoU / NMS
def test_bev_iou_known_answers_and_oracle(cuda, oracle):
    a = np.array([[0, 0, 1, 1, 0], [2, 2, 3, 3, 0]], np.float32)
    g = np.array([[0, 0, 1, 1, 0], [2, 2, 4, 4, 0], [5, 5, 6, 6, 0]], np.float32)
    ov, iou = compute_bev_iou(T(a, cuda), T(g, cuda))  # demo of bev_iou/bev_iou.py:47-57
    assert np.array_equal(N(ov), [[1, 0, 0], [0, 1, 0]]) and np.array_equal(N(iou), [[1, 0, 0], [0, 0.25, 0]])
    boxes = synthetic.make_nms_boxes(3, 700, clusters=6)
    wo, wi = oracle.compute_bev_iou(boxes[:300], boxes[300:])
    go, gi = compute_bev_iou(T(boxes[:300], cuda), T(boxes[300:], cuda))
    assert (wo > 0).mean() > 0.05
    # libdevice vs glibc sinf/cosf/atan2f differ in the last ulps (the oracle runs on the host): each side is within
    # north_star's 1e-5 of the reference kernel (test_bev_iou_matches_reference_kernel, test_oracle_golden.py), so the
    # two are within 2e-5 of each other; the observed maximum over these 120 000 pairs is printed (pytest -s)
    assert np.array_equal(N(go) == 0, wo == 0)
    for name, got, want in (("overlap", N(go), wo), ("iou", N(gi), wi)):
        nz = want != 0
        rel = np.abs(got[nz].astype(np.float64) - want[nz]) / np.abs(want[nz])
        print("bev_iou %s: %d non-zero pairs, max relative difference CUDA vs host oracle %.3g" % (name, nz.sum(), rel.max()))
        assert np.allclose(got, want, rtol=2e-5, atol=1e-6), (name, rel.max())


@needs_ref
def test_bev_iou_matches_reference_kernel(cuda):
    boxes = T(synthetic.make_nms_boxes(8, 2000, clusters=10), cuda)
    go, gi = compute_bev_iou(boxes[:1000], boxes[1000:])
    ro, ri = _refcuda.compute_bev_iou(boxes[:1000], boxes[1000:])
    assert torch.equal(go == 0, ro == 0)
    assert torch.allclose(go, ro, rtol=RTOL, atol=1e-6) and torch.allclose(gi, ri, rtol=RTOL, atol=1e-6)


@pytest.mark.parametrize("n,thresh,clusters", [(3, 0.5, 1), (64, 0.5, 2), (65, 0.3, 2), (1000, 0.7, 8),
                                               (2500, 0.01, 12), (2500, 0.85, 12)])
def test_oriented_nms_matches_oracle(cuda, oracle, n, thresh, clusters):
    if n == 3:
        boxes = np.array([[0, 0, 1, 1, 0], [2, 2, 3, 3, 0], [0, 0, 0.75, 0.75, 0]], np.float32)  # bev_iou.py:59-62
    else:
        boxes = synthetic.make_nms_boxes(n, n, clusters)
    for _ in range(8):  # keep-list equality is only defined away from the threshold
        want, (kept, margin) = oracle.oriented_nms(boxes, thresh, return_info=True)
        if margin > 1e-5:
            break
        thresh += 0.00137
    assert margin > 1e-5, "test input has a pair within 1e-5 of the threshold"
    keep, num = oriented_nms_batch(T(boxes, cuda)[None], thresh, return_num_kept=True)
    assert int(num[0]) == kept
    assert np.array_equal(N(keep[0]), want)
    assert np.array_equal(N(oriented_nms(T(boxes, cuda), thresh)), want)
    if n == 3:
        assert want.tolist() == [0, 1, 0]


@pytest.mark.parametrize("n,thresh", [(2500, 0.85), (9000, 0.8), (700, 0.01)])
def test_oriented_nms_spatial_sort_changes_nothing(cuda, n, thresh):
    """Enumerating the pairs in Morton order (tile culling) or in score order gives the same keep lists."""
    from heterofusionrcnn_b200 import _lib

    boxes = T(np.stack([synthetic.make_nms_boxes(400 + s, n, 9) for s in range(3)]), cuda)
    lib = _lib.load()
    try:
        lib.hf_nms_set_spatial_sort(0)
        k0, n0 = oriented_nms_batch(boxes, thresh, return_num_kept=True)
        lib.hf_nms_set_spatial_sort(1)
        k1, n1 = oriented_nms_batch(boxes, thresh, return_num_kept=True)
    finally:
        lib.hf_nms_set_spatial_sort(1)
    assert torch.equal(n0, n1) and torch.equal(k0, k1)


@pytest.mark.parametrize("n,thresh,max_keep", [(2500, 0.85, 512), (9000, 0.8, 512), (700, 0.01, 64), (700, 0.01, 1),
                                               (300, 0.5, 300), (300, 0.5, 1000), (65, 0.3, 7)])
def test_oriented_nms_max_keep_is_the_slice_of_the_full_list(cuda, n, thresh, max_keep):
    """max_keep (the sweep stops once that many boxes are kept - sb_nms_fn's slice, model_util.py:126): keep ==
    full keep list[:, :max_keep] incl. its keep[0] padding when fewer are kept, num_kept == min(kept, max_keep)."""
    boxes = T(np.stack([synthetic.make_nms_boxes(500 + s, n, 7) for s in range(3)]), cuda)
    full, num = oriented_nms_batch(boxes, thresh, return_num_kept=True)
    part, pnum = oriented_nms_batch(boxes, thresh, return_num_kept=True, max_keep=max_keep)
    w = min(max_keep, n)
    assert tuple(part.shape) == (3, w)
    assert torch.equal(part, full[:, :w]) and torch.equal(pnum, torch.clamp(num, max=w))


def test_oriented_nms_mask_matches_oracle_and_sweep(cuda, oracle):
    boxes = synthetic.make_nms_boxes(12, 900, 5)
    wmask = oracle.oriented_nms_mask(boxes, 0.6)
    full = N(oriented_nms_mask(T(boxes, cuda), 0.6)).view(np.uint64)
    assert np.array_equal(full, wmask)
    upper = N(oriented_nms_mask(T(boxes, cuda), 0.6, upper_only=True)).view(np.uint64)
    cb = (900 + 63) // 64
    for i in range(900):
        assert np.array_equal(upper[i, i // 64:], wmask[i, i // 64:]) and not upper[i, : i // 64].any()
    keep, _ = oracle.nms_sweep(upper)
    assert np.array_equal(keep, oracle.oriented_nms(boxes, 0.6))


@needs_ref
def test_nms_matches_reference_kernel_full_size(cuda, oracle):
    """BASELINE config 4: 9000 boxes; reference mask kernel + the reference's host sweep."""
    boxes = np.stack([synthetic.make_nms_boxes(100 + s, 9000) for s in range(2)])
    keep, num = oriented_nms_batch(T(boxes, cuda), 0.85, return_num_kept=True)
    for s in range(2):
        rmask = N(_refcuda.nms_mask(T(boxes[s], cuda), 0.85)).view(np.uint64)
        want, kept = oracle.nms_sweep(rmask)
        assert int(num[s]) == kept and np.array_equal(N(keep[s]), want)


def test_nms_full_size_properties(cuda):
    boxes = T(synthetic.make_nms_boxes(55, 9000), cuda)
    keep, num = oriented_nms_batch(boxes[None], 0.8, return_num_kept=True)
    k = int(num[0])
    kept = keep[0, :k].long()
    assert k > 0 and torch.all(kept[1:] > kept[:-1]) and torch.all(keep[0, k:] == keep[0, 0])
    _, iou = compute_bev_iou(boxes[kept], boxes[kept])
    iou.fill_diagonal_(0)
    assert float(iou.max()) <= 0.8  # no two survivors overlap more than the threshold
    # idempotence: NMS of the survivors keeps them all
    again, num2 = oriented_nms_batch(boxes[kept][None], 0.8, return_num_kept=True)
    assert int(num2[0]) == k


# ---------------------------------------------------------------- the chain
def test_chain_on_streams_equals_chain_in_program_order(cuda):
    """The multi-stream issue order of chain.run_chain must not change a single output bit."""
    from heterofusionrcnn_b200 import chain

    cfg = chain.ChainConfig().scaled(4096)
    cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels, cfg.nms_boxes = 32, 128, 32, 1500
    dev = chain.to_device(chain.make_host_inputs(cfg, batch=3, seed=5), cuda)

    def flat(res):
        out = []
        for k in sorted(res):
            if k.startswith("_"):  # stream bookkeeping (events), not a result
                continue
            v = res[k]
            out += [(k, t) for t in (v if isinstance(v, (list, tuple)) else [v])]
        return out

    for _ in range(3):  # repeated: a missing dependency would show up as a race
        a = flat(chain.run_chain(dev, cfg, overlap=False))
        b = flat(chain.run_chain(dev, cfg, overlap=True))
        torch.cuda.synchronize()
        assert len(a) == len(b)
        for (ka, ta), (kb, tb) in zip(a, b):
            assert ka == kb and torch.equal(ta, tb), ka


def test_chain_level1_shape_knob_restores_the_callers_fps_variant(cuda):
    """ChainConfig.fps_level1_warps is a scoped override: whatever hf_fps_set_variant the caller had chosen is back in
    place when run_chain returns (it used to be reset to the heuristic), and the results do not depend on the knob."""
    import ctypes

    from heterofusionrcnn_b200 import _lib, chain

    lib = _lib.load()

    def variant():
        v = [ctypes.c_int(-1) for _ in range(3)]
        lib.hf_fps_get_variant(*[ctypes.byref(x) for x in v])
        return [x.value for x in v]

    cfg = chain.ChainConfig().scaled(8192)
    cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels, cfg.nms_boxes = 16, 64, 16, 500
    dev = chain.to_device(chain.make_host_inputs(cfg, batch=2, seed=11), cuda)
    want = chain.run_chain(dev, cfg)["fps_idx"]
    knob = chain.ChainConfig(**{**cfg.__dict__})
    knob.fps_level1_warps = 8
    lib.hf_fps_set_variant(3, 4, 1)  # the caller's own choice (a probe under tools/)
    try:
        got = chain.run_chain(dev, knob)["fps_idx"]
        assert variant() == [3, 4, 1]
    finally:
        lib.hf_fps_set_variant(0, 0, 0)
    torch.cuda.synchronize()
    assert len(got) == len(want) and all(torch.equal(a, b) for a, b in zip(got, want))


@pytest.mark.parametrize("points,batch", [(8192, 3), (16384, 2), (20000, 2)])
def test_chain_fused_entries_equal_the_reference_op_sequence(cuda, points, batch):
    """The chain as bench.py issues it (fused sampling spine, group_point pairs, FPS over the input's index) against the
    same chain issued op by op like the reference's Python (farthest_point_sample + gather_point per level, two
    group_point calls per level, plain FPS): every output identical."""
    from heterofusionrcnn_b200 import chain

    cfg = chain.ChainConfig().scaled(points)
    cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels, cfg.nms_boxes = 32, 128, 32, 1200
    dev = chain.to_device(chain.make_host_inputs(cfg, batch=batch, seed=9), cuda)
    plain = chain.ChainConfig(**{**cfg.__dict__})
    plain.fused_sampling, plain.fused_grouping, plain.sorted_fps = False, False, False
    a = chain.run_chain(dev, plain, overlap=False)
    for pyramid_from, saved in ((4, 4 + 4), (1, 6 + 4), (2, 5 + 4)):  # a launch per level / levels 2-4 (3-4) in one launch
        fused = chain.ChainConfig(**{**cfg.__dict__})
        fused.sorted_fps, fused.pyramid_from = True, pyramid_from
        for overlap in (False, True):
            b = chain.run_chain(dev, fused, overlap=overlap)
            torch.cuda.synchronize()
            for l in range(4):
                assert torch.equal(a["fps_idx"][l], b["fps_idx"][l]), l
                assert torch.equal(a["xyz_levels"][l + 1], b["xyz_levels"][l + 1]), l
                for x, y in zip(a["sa%d" % (l + 1)], b["sa%d" % (l + 1)]):
                    assert torch.equal(x, y), l
                for x, y in zip(a["fp%d" % (l + 1)], b["fp%d" % (l + 1)]):
                    assert torch.equal(x, y), l
            for x, y in zip(a["crop"], b["crop"]):
                assert torch.equal(x, y)
            assert torch.equal(a["nms_keep"], b["nms_keep"]) and torch.equal(a["nms_num"], b["nms_num"])
        assert chain.chain_launches(plain) - chain.chain_launches(fused) == saved


# ---------------------------------------------------------------- three_nn fused with the FP weights
@pytest.mark.parametrize("b,n,m,indexed", [(2, 16384, 4096, True), (2, 16384, 4096, False), (3, 1024, 256, False),
                                           (2, 300, 2, False), (1, 64, 3, False)])
def test_three_nn_weights_matches_three_nn_plus_reference_formula(cuda, oracle, b, n, m, indexed):
    """Same dist / idx as three_nn (bit-exact); weights = pointnet_util.py:304-307 within 1e-5 relative."""
    from heterofusionrcnn_b200 import spatial
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn_weights

    xyz1, _ = synthetic.make_scans(31 + n, b, n, duplicate_pad=n // 8)  # duplicates: zero distances hit the 1e-10 clamp
    x1 = T(xyz1, cuda)
    x2 = gather_point(x1, farthest_point_sample(m, x1))
    ix = (spatial.build_cloud_index(x1), spatial.build_cloud_index(x2)) if indexed else (None, None)
    d0, i0 = three_nn(x1, x2)
    d, i, w = three_nn_weights(x1, x2, index1=ix[0], index2=ix[1])
    assert torch.equal(d, d0) and torch.equal(i, i0)
    wd, wi = oracle.three_nn(xyz1, N(x2))
    assert np.array_equal(N(i), wi) and np.array_equal(N(d), wd)
    want = synthetic.fp_weights(wd)
    assert close(N(w), want)
    assert np.allclose(N(w).sum(axis=2), 1.0, atol=1e-6)


# ---------------------------------------------------------------- the chain as a captured CUDA graph
def test_chain_graph_replay_equals_eager_issue(cuda):
    """chain.ChainGraph (one cudaGraphLaunch per step, what bench.py times) gives the results of eager issue."""
    from heterofusionrcnn_b200 import chain

    cfg = chain.ChainConfig().scaled(4096)
    cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels, cfg.nms_boxes = 32, 64, 32, 800
    dev = chain.to_device(chain.make_host_inputs(cfg, batch=3, seed=11), cuda)
    eager = chain.run_chain(dev, cfg)
    torch.cuda.synchronize()
    graph = chain.ChainGraph(dev, cfg)
    assert graph.launches == chain.chain_launches(cfg)
    for _ in range(2):
        res = graph.replay()
    torch.cuda.synchronize()
    for l in range(4):
        assert torch.equal(res["fps_idx"][l], eager["fps_idx"][l])
        for g, e in zip(res["sa%d" % (l + 1)], eager["sa%d" % (l + 1)]):
            assert torch.equal(g, e)
        for g, e in zip(res["fp%d" % (l + 1)], eager["fp%d" % (l + 1)]):
            assert torch.equal(g, e)
    for g, e in zip(res["crop"], eager["crop"]):
        assert torch.equal(g, e)
    assert torch.equal(res["nms_keep"], eager["nms_keep"]) and torch.equal(res["nms_num"], eager["nms_num"])


# ---------------------------------------------------------------- BASELINE configs[0]: a real KITTI scan
def test_config0_real_kitti_scan(cuda, oracle):
    """Fixture scan 000000 (reference loader + FOV filter + 16384 sampling, tests/golden/make_golden_kitti.py):
    the SA / FP ops of level 1 against the oracle, bit-exact indices, and the unmodified reference kernels."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kitti_000000_16384.npz"))
    xyz = g["xyz"][None]
    x = T(xyz, cuda)
    want_f = oracle.farthest_point_sample(4096, xyz)
    fidx = farthest_point_sample(4096, x)
    assert np.array_equal(N(fidx), want_f)
    known = gather_point(x, fidx)
    wk = oracle.gather_point(xyz, want_f)
    assert np.array_equal(N(known), wk)
    for r in (0.1, 0.5, 1.0):
        wi, wc = oracle.query_ball_point(r, 32, xyz, wk)
        gi, gc = query_ball_point(r, 32, x, known)
        assert np.array_equal(N(gc), wc) and np.array_equal(N(gi), wi), r
    wd, wi = oracle.three_nn(xyz, wk)
    gd, gi = three_nn(x, known)
    assert np.array_equal(N(gi), wi) and np.array_equal(N(gd), wd)
    w = synthetic.fp_weights(wd)
    feats = np.random.default_rng(0).standard_normal((1, 4096, 128)).astype(np.float32)
    got = three_interpolate(T(feats, cuda), gi, T(w, cuda))
    assert close(N(got), oracle.three_interpolate(feats, wi, w))
    if _refcuda.available():
        assert torch.equal(_refcuda.fps(4096, x), fidx)
        rd, ri = _refcuda.three_nn(x, known)
        assert torch.equal(ri, gi) and torch.equal(rd, gd)
