"""GPU: the deterministic scatter-add gradients (csrc/backward.cu) through the Python mirror -> ctypes
-> C-ABI, against the CPU oracle (the reference's loops run serially = ascending entry order) and,
where oracle/_ref/libref_cuda.so travelled, against the UNMODIFIED reference kernels on the same B200.

Bar.  The CUDA path promises the sum in ascending entry order, which is what the oracle computes:
BIT-EXACT against the oracle.  The reference kernels use float atomicAdd, so their own result varies
from run to run in the last bits; against them the tolerance is 1e-5 relative to the sum of the
absolute contributions of the row (the north_star's 1e-5, stated on the quantity the rounding
error of a reordered sum is proportional to).
"""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.cropping.tf_cropping import pc_crop_and_sample, pc_crop_and_sample_grad_fts
from heterofusionrcnn_b200.grouping.tf_grouping import group_point_grad, query_ball_point
from heterofusionrcnn_b200.interpolate.tf_interpolate import (three_interpolate_grad,
                                                              three_interpolate_grad_channel_first, three_nn)
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point, gather_point_grad

import _refcuda

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not _refcuda.available(), reason="oracle/_ref/libref_cuda.so not built")


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def N(t):
    return t.detach().cpu().numpy()


def reorder_close(got, ref, abs_sum, rtol=1e-5):
    return np.all(np.abs(np.asarray(got, np.float64) - np.asarray(ref, np.float64)) <= rtol * abs_sum + 1e-30)


# ---------------------------------------------------------------- gather_point_grad
@pytest.mark.parametrize("b,n,m,mode", [(2, 4096, 1024, "fps"), (3, 1000, 3000, "random"), (1, 7, 100, "random"),
                                        (2, 500, 0, "random"), (2, 300, 2000, "hot")])
def test_gather_point_grad(cuda, oracle, b, n, m, mode):
    rng = np.random.default_rng(100 + n + m)
    xyz, _ = synthetic.make_scans(3, b, n)
    if mode == "fps":
        idx = N(farthest_point_sample(m, T(xyz, cuda)))
    elif mode == "hot":  # one point collects > 1024 contributions: the windowed path
        idx = rng.integers(0, n, (b, m)).astype(np.int32)
        idx[:, : m * 3 // 4] = 5
    else:
        idx = rng.integers(0, n, (b, m)).astype(np.int32)
    out_g = rng.standard_normal((b, m, 3)).astype(np.float32)
    want = oracle.gather_point_grad(xyz, idx, out_g)
    got = N(gather_point_grad(T(xyz, cuda), T(idx, cuda), T(out_g, cuda)))
    assert got.shape == (b, n, 3) and np.array_equal(got, want)


# ---------------------------------------------------------------- group_point_grad
@pytest.mark.parametrize("b,n,m,ns,c,radius", [(2, 4096, 1024, 32, 96, 0.5), (2, 4096, 1024, 32, 3, 0.2),
                                               (1, 2048, 512, 16, 1, 1.0), (2, 1024, 256, 32, 256, 2.0),
                                               (2, 1000, 250, 7, 130, 1.0), (1, 512, 128, 64, 1028, 4.0),
                                               (2, 300, 64, 32, 64, 100.0)])
def test_group_point_grad(cuda, oracle, b, n, m, ns, c, radius):
    rng = np.random.default_rng(n + c)
    xyz, _ = synthetic.make_scans(7, b, n)
    xyz_d = T(xyz, cuda)
    new_xyz = gather_point(xyz_d, farthest_point_sample(m, xyz_d))
    idx, _ = query_ball_point(radius, ns, xyz_d, new_xyz)  # real neighbourhoods, padded with the first hit
    grad_out = rng.standard_normal((b, m, ns, c)).astype(np.float32)
    pts = np.zeros((b, n, c), np.float32)
    want = oracle.group_point_grad(pts, N(idx), grad_out)
    got = group_point_grad(T(pts, cuda), idx, T(grad_out, cuda))
    assert np.array_equal(N(got), want)
    again = group_point_grad(T(pts, cuda), idx, T(grad_out, cuda))
    assert torch.equal(got, again)  # run-to-run deterministic


def test_group_point_grad_hot_row_and_bad_index(cuda, oracle):
    """All rows point at one point (131072 entries on one destination: the windowed sort), plus
    out-of-range indices, which are skipped."""
    rng = np.random.default_rng(3)
    b, n, m, ns, c = 1, 64, 4096, 32, 8
    idx = np.full((b, m, ns), 9, np.int32)
    idx[0, ::7, 3] = rng.integers(0, n, idx[0, ::7, 3].shape)
    idx[0, 5, 0], idx[0, 6, 1] = -1, n
    grad_out = rng.standard_normal((b, m, ns, c)).astype(np.float32)
    pts = np.zeros((b, n, c), np.float32)
    want = oracle.group_point_grad(pts, idx, grad_out)
    got = N(group_point_grad(T(pts, cuda), T(idx, cuda), T(grad_out, cuda)))
    assert np.array_equal(got, want)


# ---------------------------------------------------------------- three_interpolate_grad
@pytest.mark.parametrize("b,n,m,c", [(2, 4096, 1024, 128), (2, 1024, 256, 512), (1, 256, 64, 1024), (2, 1000, 37, 6),
                                     (1, 64, 3, 1), (2, 2048, 512, 36)])
def test_three_interpolate_grad(cuda, oracle, b, n, m, c):
    rng = np.random.default_rng(n + m + c)
    xyz1, _ = synthetic.make_scans(5, b, n)
    xyz2 = xyz1[:, :m] + 0.01
    dist, idx = three_nn(T(xyz1, cuda), T(xyz2, cuda))
    w = synthetic.fp_weights(N(dist))
    grad_out = rng.standard_normal((b, n, c)).astype(np.float32)
    pts = np.zeros((b, m, c), np.float32)
    want = oracle.three_interpolate_grad(pts, N(idx), w, grad_out)
    got = N(three_interpolate_grad(T(pts, cuda), idx, T(w, cuda), T(grad_out, cuda)))
    assert got.shape == (b, m, c) and np.array_equal(got, want)
    # the native op's channel-first layout gives the same numbers
    got_cf = N(three_interpolate_grad_channel_first(T(np.transpose(pts, (0, 2, 1)), cuda), idx, T(w, cuda),
                                                    T(np.transpose(grad_out, (0, 2, 1)), cuda)))
    assert np.array_equal(np.transpose(got_cf, (0, 2, 1)), want)


# ---------------------------------------------------------------- crop grad_fts
@pytest.mark.parametrize("B,P,C,nb,R", [(2, 4096, 32, 24, 128), (3, 2000, 288, 10, 64), (1, 512, 5, 6, 700)])
def test_pc_crop_and_sample_grad_fts(cuda, oracle, B, P, C, nb, R):
    rng = np.random.default_rng(P + C)
    from test_gpu_parity import crop_inputs

    xyz, fts, inten, mask, boxes, box_ind = crop_inputs(17, B, P, C, nb)
    out = pc_crop_and_sample(T(xyz, cuda), T(fts, cuda), T(inten, cuda), T(mask, cuda), T(boxes, cuda), T(box_ind, cuda), R)
    crop_ind = out[4]
    nb = boxes.shape[0]
    g = rng.standard_normal((nb, R, C)).astype(np.float32)
    want = oracle.pc_crop_and_sample_grad_fts(fts, box_ind, N(crop_ind), g)
    got = N(pc_crop_and_sample_grad_fts(T(fts, cuda), T(box_ind, cuda), crop_ind, T(g, cuda)))
    assert got.shape == (B, P, C) and np.array_equal(got, want)


# ---------------------------------------------------------------- against the reference kernels (float atomics)
@needs_ref
def test_gradients_match_reference_kernels(cuda):
    rng = np.random.default_rng(9)
    b, n, m, ns, c = 2, 2048, 512, 16, 64
    xyz, _ = synthetic.make_scans(7, b, n)
    xyz_d = T(xyz, cuda)
    fidx = farthest_point_sample(m, xyz_d)
    new_xyz = gather_point(xyz_d, fidx)
    idx, _ = query_ball_point(1.0, ns, xyz_d, new_xyz)
    pts = torch.zeros((b, n, c), device=cuda)
    go = T(rng.standard_normal((b, m, ns, c)).astype(np.float32), cuda)
    abs_sum = N(_refcuda.group_point_grad(pts, idx, go.abs()))
    assert reorder_close(N(group_point_grad(pts, idx, go)), N(_refcuda.group_point_grad(pts, idx, go)), abs_sum)

    og = T(rng.standard_normal((b, m, 3)).astype(np.float32), cuda)
    abs_sum = N(_refcuda.gather_point_grad(xyz_d, fidx, og.abs()))
    assert reorder_close(N(gather_point_grad(xyz_d, fidx, og)), N(_refcuda.gather_point_grad(xyz_d, fidx, og)), abs_sum)

    dist, nidx = three_nn(xyz_d, new_xyz)
    w = T(synthetic.fp_weights(N(dist)), cuda)
    known = torch.zeros((b, c, m), device=cuda)
    g_cf = T(rng.standard_normal((b, c, n)).astype(np.float32), cuda)
    abs_sum = N(_refcuda.three_interpolate_grad_channel_first(known, nidx, w, g_cf.abs()))
    assert reorder_close(N(three_interpolate_grad_channel_first(known, nidx, w, g_cf)),
                         N(_refcuda.three_interpolate_grad_channel_first(known, nidx, w, g_cf)), abs_sum)


# ---------------------------------------------------------------- autograd pairing (the reference's RegisterGradient)
def test_autograd_functions_use_the_registered_gradients(cuda, oracle):
    """loss = <forward(x), g>  =>  x.grad == grad_op(g): the torch.autograd pairing of heterofusionrcnn_b200.autograd
    reproduces tf.RegisterGradient of the reference (points / fts get gradients, indices do not)."""
    from heterofusionrcnn_b200 import autograd as A
    from test_gpu_parity import crop_inputs

    rng = np.random.default_rng(8)
    b, n, m, ns, c = 2, 1024, 256, 16, 32
    xyz, _ = synthetic.make_scans(3, b, n)
    x = T(xyz, cuda)
    fidx = farthest_point_sample(m, x)
    new_xyz = gather_point(x, fidx)
    idx, _ = query_ball_point(1.0, ns, x, new_xyz)

    pts = T(rng.standard_normal((b, n, c)).astype(np.float32), cuda).requires_grad_(True)
    g = T(rng.standard_normal((b, m, ns, c)).astype(np.float32), cuda)
    (A.group_point(pts, idx) * g).sum().backward()
    assert np.array_equal(N(pts.grad), oracle.group_point_grad(N(pts), N(idx), N(g)))

    xr = x.clone().requires_grad_(True)
    g3 = T(rng.standard_normal((b, m, 3)).astype(np.float32), cuda)
    (A.gather_point(xr, fidx) * g3).sum().backward()
    assert np.array_equal(N(xr.grad), oracle.gather_point_grad(xyz, N(fidx), N(g3)))

    dist, nidx = three_nn(x, new_xyz)
    w = T(synthetic.fp_weights(N(dist)), cuda)
    known = T(rng.standard_normal((b, m, c)).astype(np.float32), cuda).requires_grad_(True)
    gi = T(rng.standard_normal((b, n, c)).astype(np.float32), cuda)
    (A.three_interpolate(known, nidx, w) * gi).sum().backward()
    assert np.array_equal(N(known.grad), oracle.three_interpolate_grad(N(known), N(nidx), N(w), N(gi)))

    cx, cf, ci, cm, cb, cbi = crop_inputs(5, 2, 2048, 8, 12)
    fts = T(cf, cuda).requires_grad_(True)
    out = A.pc_crop_and_sample(T(cx, cuda), fts, T(ci, cuda), T(cm, cuda), T(cb, cuda), T(cbi, cuda), 32)
    gc = T(rng.standard_normal(tuple(out[1].shape)).astype(np.float32), cuda)
    (out[1] * gc).sum().backward()
    assert np.array_equal(N(fts.grad), oracle.pc_crop_and_sample_grad_fts(cf, cbi, N(out[4]), N(gc)))
    assert not out[0].requires_grad and not out[4].requires_grad
