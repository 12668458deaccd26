#!/usr/bin/env python
"""GPU micro-benchmark of the deterministic gradients (csrc/backward.cu) at the chain's shapes:
ms per call, algorithmic GB/s (read the incoming gradient once + write the result once + indices),
and the UNMODIFIED reference kernel (float atomicAdd + memset, oracle/_ref/libref_cuda.so) beside it.

    python tools/backward_bench.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.cropping.tf_cropping import pc_crop_and_sample_grad_fts
from heterofusionrcnn_b200.grouping.tf_grouping import group_point_grad, query_ball_point
from heterofusionrcnn_b200.interpolate.tf_interpolate import three_interpolate_grad, three_nn
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point, gather_point_grad
from fps_multi_sweep import timeit
import _refcuda


def main():
    hf.load()
    dev = torch.device("cuda:0")
    b = 16
    have_ref = _refcuda.available()
    x0 = torch.from_numpy(synthetic.make_scans(3, b, 16384)[0]).to(dev)
    levels = [(16384, 4096, 0.5, 1, 128), (4096, 1024, 1.0, 96, 512), (1024, 256, 2.0, 256, 512)]
    xyz = x0
    for n, m, r, c_sa, c_fp in levels:
        fidx = farthest_point_sample(m, xyz)
        new_xyz = gather_point(xyz, fidx)
        idx, _ = query_ball_point(r, 32, xyz, new_xyz)
        for c in sorted({c_sa, 3}):
            pts = torch.zeros((b, n, c), device=dev)
            go = torch.randn((b, m, 32, c), device=dev)
            ms = timeit(lambda: group_point_grad(pts, idx, go))
            nbytes = 4 * (go.numel() + pts.numel() + 3 * idx.numel())
            line = "group_point_grad      n=%5d m=%4d ns=32 c=%3d  %8.4f ms  %7.1f GB/s" % (n, m, c, ms, nbytes / ms / 1e6)
            if have_ref and (c <= 96 or n <= 4096):
                line += "   reference (atomics) %8.4f ms" % timeit(lambda: _refcuda.group_point_grad(pts, idx, go), reps=2)
            print(line)
        dist, nidx = three_nn(xyz, new_xyz)
        w = torch.from_numpy(synthetic.fp_weights(dist.cpu().numpy())).to(dev)
        known = torch.zeros((b, m, c_fp), device=dev)
        go = torch.randn((b, n, c_fp), device=dev)
        ms = timeit(lambda: three_interpolate_grad(known, nidx, w, go))
        nbytes = 4 * (go.numel() + known.numel() + 3 * 3 * nidx.numel() // 3 * 3)
        line = "three_interpolate_grad n=%5d m=%4d c=%4d       %8.4f ms  %7.1f GB/s" % (n, m, c_fp, ms, nbytes / ms / 1e6)
        if have_ref:
            kcf, gcf = known.transpose(1, 2).contiguous(), go.transpose(1, 2).contiguous()
            line += "   reference (atomics, channel-first) %8.4f ms" % timeit(
                lambda: _refcuda.three_interpolate_grad_channel_first(kcf, nidx, w, gcf), reps=2)
        print(line)
        og = torch.randn((b, m, 3), device=dev)
        ms = timeit(lambda: gather_point_grad(xyz, fidx, og))
        print("gather_point_grad     n=%5d m=%4d              %8.4f ms" % (n, m, ms))
        sys.stdout.flush()
        xyz = new_xyz
    # RoI crop: 16 scans x 512 boxes x 512 slots, C = 288 (BASELINE config 3)
    rng = np.random.default_rng(0)
    N_, R, C, P = 8192, 512, 288, 16384
    box_ind = torch.from_numpy(np.repeat(np.arange(16, dtype=np.int32), 512)).to(dev)
    centre = rng.integers(0, P, (N_, 1))
    crop_ind = torch.from_numpy(((centre + rng.integers(0, 400, (N_, R))) % P).astype(np.int32)).to(dev)
    fts = torch.zeros((16, P, C), device=dev)
    g = torch.randn((N_, R, C), device=dev)
    ms = timeit(lambda: pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, g), reps=3)
    nbytes = 4 * (g.numel() + fts.numel() + 3 * crop_ind.numel())
    line = "crop grad_fts         N=%d R=%d C=%d           %8.4f ms  %7.1f GB/s" % (N_, R, C, ms, nbytes / ms / 1e6)
    if have_ref:
        line += "   reference (atomics) %8.4f ms" % timeit(lambda: _refcuda.pc_crop_and_sample_grad_fts(fts, box_ind, crop_ind, g), reps=1)
    print(line)


if __name__ == "__main__":
    main()
