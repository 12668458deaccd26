#!/usr/bin/env python
"""GPU probe: how many FPS clusters are co-resident?  Times fps (multi-pick) against the batch size for
cluster sizes 4 and 8: a step in ms between b and b+1 marks the end of the first wave.

    python tools/fps_wave_probe.py [n]
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
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    lib = hf.load()
    dev = torch.device("cuda:0")
    xyz, _ = synthetic.make_scans(3, 20, n)
    x = torch.from_numpy(xyz).to(dev)
    for cluster in (8, 4):
        lib.hf_fps_set_variant(3, 0, 0)
        lib.hf_fps_set_cluster(cluster)
        for b in (1, 2, 4, 8, 10, 12, 13, 14, 15, 16, 17, 18, 20):
            xb = x[:b].contiguous()
            ms = timeit(lambda: farthest_point_sample(n // 4, xb), reps=3)
            print("n=%d cluster=%d b=%2d  %8.4f ms" % (n, cluster, b, ms))
            sys.stdout.flush()
    lib.hf_fps_set_cluster(0)
    lib.hf_fps_set_variant(0, 0, 0)


if __name__ == "__main__":
    main()
