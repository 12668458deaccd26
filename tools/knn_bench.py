#!/usr/bin/env python
"""GPU micro-benchmark of knn_point (fused distance + selection, pruned / every-pair) and select_top_k
(with the unmodified reference SelectionSort kernel beside it).   python tools/knn_bench.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.grouping.tf_grouping import knn_point, select_top_k
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point
from fps_multi_sweep import timeit
import _refcuda


def main():
    lib = hf.load()
    dev = torch.device("cuda:0")
    b = 16
    xyz = torch.from_numpy(synthetic.make_scans(3, b, 16384)[0]).to(dev)
    for n, m in [(16384, 4096), (4096, 1024), (1024, 256)]:
        new_xyz = gather_point(xyz, farthest_point_sample(m, xyz))
        for k in (16, 32):
            for mode, name in ((1, "every-pair"), (2, "pruned")):
                lib.hf_spatial_set_mode(mode)
                ms = timeit(lambda: knn_point(k, xyz, new_xyz))
                lib.hf_spatial_set_mode(0)
                print("knn_point n=%5d m=%4d k=%2d b=%d  %-10s %8.4f ms   (%.1f G pair-evals/s equivalent)" % (
                    n, m, k, b, name, ms, b * n * m / ms / 1e6))
        xyz = new_xyz
    d = torch.rand((2, 1024, 4096), device=dev)
    ms = timeit(lambda: select_top_k(32, d))
    line = "select_top_k b=2 m=1024 n=4096 k=32  %8.4f ms" % ms
    if _refcuda.available():
        line += "   reference kernel %8.2f ms" % timeit(lambda: _refcuda.select_top_k(32, d), reps=1)
    print(line)


if __name__ == "__main__":
    main()
