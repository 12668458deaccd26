#!/usr/bin/env python
"""GPU probe: level-1 FPS (16384 -> 4096) for batches of 14..17 scans in the cluster shapes that matter for 16 scans:
8 CTAs x 4 warps (15 co-resident on a B200), 4 x 8, and 6 x 8 warps (22 co-resident, tools/gpc_probe.cu).

    python tools/fps_six_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample
from fps_multi_sweep import timeit


def main():
    lib = hf.load()
    dev = torch.device("cuda:0")
    x = torch.from_numpy(synthetic.make_scans(3, 20, 16384)[0]).to(dev)
    for b in (8, 15, 16, 20):
        xb = x[:b].contiguous()
        ref = farthest_point_sample(4096, xb)
        for cluster, warps in [(8, 4), (4, 8), (6, 0)]:
            lib.hf_fps_set_variant(3, warps, 0)
            lib.hf_fps_set_cluster(cluster)
            try:
                out = farthest_point_sample(4096, xb)
                ms = timeit(lambda: farthest_point_sample(4096, xb), reps=5)
                print("b=%2d cluster=%d warps=%d  %8.4f ms  same=%s" % (b, cluster, warps or 8, ms, bool(torch.equal(out, ref))))
            finally:
                lib.hf_fps_set_cluster(0)
                lib.hf_fps_set_variant(0, 0, 0)
            sys.stdout.flush()


if __name__ == "__main__":
    main()
