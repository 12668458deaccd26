"""GPU: knn_point and select_top_k (csrc/spatial.cu knn_kernel, csrc/topk.cu) through the Python mirror ->
ctypes -> C-ABI against the CPU oracle and the unmodified reference SelectionSort kernel on the same B200.
Bar: bit-exact (indices, values, and for select_top_k the whole output including the unsorted tail)."""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import _lib, synthetic
from heterofusionrcnn_b200.grouping.tf_grouping import knn_point, select_top_k
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point

import _refcuda

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not _refcuda.available(), reason="oracle/_ref/libref_cuda.so not built")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def N(t):
    return t.detach().cpu().numpy()


@pytest.mark.parametrize("b,n,m,k,dup,mode", [(2, 16384, 4096, 16, 0, 0), (2, 16384, 4096, 32, 2000, 0),
                                              (2, 16384, 4096, 16, 0, 1), (2, 4096, 1024, 32, 500, 2),
                                              (2, 4096, 1024, 32, 500, 1), (1, 1000, 77, 64, 0, 0), (2, 300, 300, 1, 0, 0),
                                              (1, 40, 5, 40, 10, 0), (2, 2000, 130, 128, 0, 2), (1, 5000, 33, 8, 0, 2)])
def test_knn_point_matches_oracle(cuda, oracle, b, n, m, k, dup, mode):
    """mode 0 = auto, 1 = every-pair kernel forced, 2 = pruned kernel forced."""
    xyz, _ = synthetic.make_scans(41 + n + k, b, n, duplicate_pad=dup)
    x1 = T(xyz, cuda)
    x2 = gather_point(x1, farthest_point_sample(m, x1)) if m <= n else x1
    want_v, want_i = oracle.knn_point(k, xyz, N(x2))
    _lib.load().hf_spatial_set_mode(mode)
    try:
        val, idx = knn_point(k, x1, x2)
    finally:
        _lib.load().hf_spatial_set_mode(0)
    assert val.shape == (b, m, k) and idx.dtype == torch.int32
    assert np.array_equal(N(idx), want_i)
    assert np.array_equal(N(val), want_v)


@pytest.mark.parametrize("b,n,m,k,mode", [(2, 16384, 1024, 16, 0), (2, 16384, 512, 32, 1), (1, 40000, 256, 32, 2),
                                          (1, 65536, 300, 16, 0), (2, 20000, 200, 64, 2)])
def test_knn_point_is_pinned_on_the_reference_formula(cuda, oracle, b, n, m, k, mode):
    """grouping/tf_grouping.py:78-93 restated in numpy float32 (oracle.knn_point_reference_form): the CUDA result has
    the reference's neighbours at every rank that formula decides beyond its own rounding error, and its values
    within that error.  Includes clouds above 16384 points (searched in passes over the parts of the index)."""
    xyz, _ = synthetic.make_scans(71 + n + k, b, n)
    x1 = T(xyz, cuda)
    x2 = gather_point(x1, farthest_point_sample(m, x1))
    _lib.load().hf_spatial_set_mode(mode)
    try:
        val, idx = knn_point(k, x1, x2)
    finally:
        _lib.load().hf_spatial_set_mode(0)
    frac = oracle.check_knn_against_reference_form(k, xyz, N(x2), N(val), N(idx))
    assert frac > 0.5, frac
    if n <= 20000:  # and bit-exact against the oracle's direct form (serial: small cases only)
        want_v, want_i = oracle.knn_point(k, xyz, N(x2))
        assert np.array_equal(N(idx), want_i) and np.array_equal(N(val), want_v)


def test_knn_large_cloud_pruned_equals_every_pair(cuda):
    """Clouds of 16385..65536 points: the pruned search (index in parts) against the every-pair kernel, bit for bit."""
    xyz, _ = synthetic.make_scans(13, 2, 50000, duplicate_pad=3000)
    x1 = T(xyz, cuda)
    x2 = gather_point(x1, farthest_point_sample(3000, x1))
    lib = _lib.load()
    try:
        lib.hf_spatial_set_mode(1)
        v1, i1 = knn_point(32, x1, x2)
        lib.hf_spatial_set_mode(2)
        v2, i2 = knn_point(32, x1, x2)
    finally:
        lib.hf_spatial_set_mode(0)
    assert torch.equal(i1, i2) and torch.equal(v1, v2)


def test_knn_indices_general_matches_knn_point(cuda, oracle):
    """hf/core/pointfly.py:185-214: (distances, (batch, point) index pairs) of the same search."""
    from heterofusionrcnn_b200.pointfly import knn_indices_general

    xyz, _ = synthetic.make_scans(5, 3, 3000, duplicate_pad=100)
    pts = T(xyz, cuda)
    qrs = gather_point(pts, farthest_point_sample(200, pts))
    for k in (8, 48):
        dist, ind = knn_indices_general(qrs, pts, k, True)
        assert tuple(dist.shape) == (3, 200, k) and tuple(ind.shape) == (3, 200, k, 2) and ind.dtype == torch.int32
        val, idx = knn_point(k, pts, qrs)
        assert torch.equal(dist, val) and torch.equal(ind[..., 1], idx)
        assert torch.equal(ind[..., 0], torch.arange(3, device=cuda, dtype=torch.int32)[:, None, None].expand(-1, 200, k))
        # tf.gather_nd(points, indices) == the neighbours
        nb = pts[ind[..., 0].long(), ind[..., 1].long()]
        d = ((nb - qrs[:, :, None, :]) ** 2).sum(-1)
        assert torch.allclose(d, dist, rtol=1e-5, atol=1e-6)
        oracle.check_knn_against_reference_form(k, xyz, N(qrs), N(dist), N(ind[..., 1]))


def test_knn_point_rejects_bad_arguments(cuda):
    x = torch.zeros((1, 8, 3), device=cuda)
    with pytest.raises(_lib.InvalidArgumentError):
        knn_point(0, x, x)
    with pytest.raises(_lib.InvalidArgumentError):
        knn_point(9, x, x)
    with pytest.raises(_lib.InvalidArgumentError):
        knn_point(2, torch.zeros((1, 8, 4), device=cuda), x)


@pytest.mark.parametrize("b,m,n,k,ties", [(2, 64, 1024, 32, False), (1, 33, 4096, 16, True), (2, 5, 17, 17, True),
                                          (1, 3, 25600, 8, False), (1, 4, 7, 20, True)])
def test_select_top_k_matches_oracle(cuda, oracle, b, m, n, k, ties):
    rng = np.random.default_rng(n + k)
    d = rng.integers(0, 50, (b, m, n)).astype(np.float32) if ties else rng.random((b, m, n)).astype(np.float32)
    wi, wv = oracle.select_top_k(k, d)
    gi, gv = select_top_k(k, T(d, cuda))
    assert np.array_equal(N(gi), wi) and np.array_equal(N(gv), wv)


@needs_ref
def test_select_top_k_matches_reference_kernel(cuda):
    rng = np.random.default_rng(12)
    d = T(rng.integers(0, 30, (2, 40, 300)).astype(np.float32), cuda)
    ri, rv = _refcuda.select_top_k(12, d)
    gi, gv = select_top_k(12, d)
    assert torch.equal(gi, ri) and torch.equal(gv, rv)


def test_select_top_k_of_the_distance_matrix_agrees_with_knn_point(cuda):
    """The two halves of the reference's knn_point_old path (tf_grouping.py:98-120): distance matrix +
    SelectionSort gives the same neighbours as the fused kernel when there are no ties."""
    xyz, _ = synthetic.make_scans(77, 1, 2048)
    x1 = T(xyz, cuda)
    x2 = x1[:, :128].contiguous()
    d = ((x2[:, :, None, :] - x1[:, None, :, :]) ** 2).sum(-1)
    oi, _ = select_top_k(8, d)
    _, ki = knn_point(8, x1, x2)
    assert (oi[:, :, :8] == ki).float().mean().item() > 0.999  # the torch matrix rounds differently from sqdist


@pytest.mark.parametrize("b,n,m,k", [(2, 2048, 128, 8), (1, 1000, 77, 32), (2, 300, 300, 1)])
def test_knn_point_old_is_the_reference_composition(cuda, oracle, b, n, m, k):
    """knn_point_old (grouping/tf_grouping.py:98-131) = expanded-form matrix -> SelectionSort -> first k columns.
    Exact against the oracle's selection sort run on the very matrix the op sorted; pinned on the reference's float32
    formula (restated in numpy) within that formula's own rounding error; and the neighbours agree with the fused
    knn_point except at near-ties."""
    from heterofusionrcnn_b200.grouping.tf_grouping import expanded_sqdist, knn_point_old

    xyz, _ = synthetic.make_scans(n + k, b, n)
    x1 = T(xyz, cuda)
    x2 = x1[:, ::max(1, n // m)][:, :m].contiguous() + 0.01
    val, idx = knn_point_old(k, x1, x2)
    assert val.shape == (b, m, k) and idx.shape == (b, m, k) and idx.dtype == torch.int32
    wi, wv = oracle.select_top_k(k, N(expanded_sqdist(x1, x2)))
    assert np.array_equal(N(idx), wi[:, :, :k]) and np.array_equal(N(val), wv[:, :, :k])
    decided = oracle.check_knn_against_reference_form(k, N(x1), N(x2), N(val), N(idx))
    assert decided > 0.5  # most ranks are decided by a margin above the formula's own rounding error
    _, ki = knn_point(k, x1, x2)
    assert (idx == ki).float().mean().item() > 0.98


# ---------------------------------------------------------------- prob_sample
@pytest.mark.parametrize("b,n,m", [(2, 40, 1000), (3, 8192, 4096), (2, 20000, 5000), (1, 8195, 300), (2, 1, 10), (1, 3, 7),
                                   (1, 100000, 20000)])
def test_prob_sample_matches_oracle_and_reference_kernel(cuda, oracle, b, n, m):
    """The sampled indices depend on the rounding of the float32 prefix sum: bit-exact against the oracle's
    restatement of the reference's scan tree and against the unmodified reference kernel on the same B200."""
    from heterofusionrcnn_b200.sampling.tf_sampling import prob_sample

    rng = np.random.default_rng(n + m)
    w = rng.random((b, n)).astype(np.float32)
    w[:, ::7] = 0.0  # zero-weight categories: runs of equal prefix sums
    r = rng.random((b, m)).astype(np.float32)
    r[:, 0] = 0.0
    want = oracle.prob_sample(w, r)
    got = prob_sample(T(w, cuda), T(r, cuda))
    assert got.dtype == torch.int32 and np.array_equal(N(got), want)
    if _refcuda.available():
        ref, ref_cs = _refcuda.prob_sample(T(w, cuda), T(r, cuda))
        assert torch.equal(got, ref)
        assert np.array_equal(N(ref_cs), oracle.cumsum(w))
