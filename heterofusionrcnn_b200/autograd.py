"""Differentiable versions of the ops the reference registers gradients for.

The reference attaches its gradient ops with tf.RegisterGradient:
    GatherPoint       -> gather_point_grad            sampling/tf_sampling.py:54-58
    GroupPoint        -> group_point_grad             grouping/tf_grouping.py:55-59
    ThreeInterpolate  -> three_interpolate_grad       interpolate/tf_interpolate.py:40-49
    PcCropAndSample   -> pc_crop_and_sample_grad_fts  cropping/tf_cropping.py:36-46  (w.r.t. fts only)
and declares the index producers non-differentiable (ops.NoGradient: FarthestPointSample, QueryBallPoint,
SelectionSort, ThreeNN, ProbSample).  With torch tensors the same pairing is a torch.autograd.Function; the
functions below take and return what the forward ops take and return.  Gradients are deterministic
(csrc/backward.cu).
"""
import torch

from .cropping.tf_cropping import pc_crop_and_sample as _crop, pc_crop_and_sample_grad_fts
from .grouping.tf_grouping import group_point as _group_point, group_point_grad
from .interpolate.tf_interpolate import three_interpolate as _three_interpolate, three_interpolate_grad
from .sampling.tf_sampling import gather_point as _gather_point, gather_point_grad


class _GatherPoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, idx):
        ctx.save_for_backward(inp, idx)
        return _gather_point(inp, idx)

    @staticmethod
    def backward(ctx, out_g):
        inp, idx = ctx.saved_tensors
        return gather_point_grad(inp, idx, out_g.contiguous()), None


class _GroupPoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, idx):
        ctx.save_for_backward(points, idx)
        return _group_point(points, idx)

    @staticmethod
    def backward(ctx, grad_out):
        points, idx = ctx.saved_tensors
        return group_point_grad(points, idx, grad_out.contiguous()), None


class _ThreeInterpolate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, idx, weight):
        ctx.save_for_backward(points, idx, weight)
        return _three_interpolate(points, idx, weight)

    @staticmethod
    def backward(ctx, grad_out):
        points, idx, weight = ctx.saved_tensors
        return three_interpolate_grad(points, idx, weight, grad_out.contiguous()), None, None  # as the reference: points only


class _PcCropAndSample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pts, fts, intensities, mask, boxes, box_ind, resize):
        out = _crop(pts, fts, intensities, mask, boxes, box_ind, resize)
        ctx.save_for_backward(fts, box_ind, out[4])
        ctx.mark_non_differentiable(out[0], out[2], out[3], out[4], out[5])
        return out

    @staticmethod
    def backward(ctx, g_pts, g_fts, g_int, g_mask, g_ind, g_ne):
        fts, box_ind, crop_ind = ctx.saved_tensors
        return None, pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, g_fts.contiguous()), None, None, None, None, None


def gather_point(inp, idx):
    """sampling/tf_sampling.py:38-58 with its registered gradient."""
    return _GatherPoint.apply(inp, idx)


def group_point(points, idx):
    """grouping/tf_grouping.py:44-59 with its registered gradient."""
    return _GroupPoint.apply(points, idx)


def three_interpolate(points, idx, weight):
    """interpolate/tf_interpolate.py:26-49 with its registered gradient (w.r.t. points)."""
    return _ThreeInterpolate.apply(points, idx, weight)


def pc_crop_and_sample(pts, fts, intensities, mask, boxes, box_ind, resize):
    """cropping/tf_cropping.py:13-46 with its registered gradient (w.r.t. fts)."""
    return _PcCropAndSample.apply(pts, fts, intensities, mask, boxes, box_ind, resize)
