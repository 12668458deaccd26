#!/usr/bin/env python
"""GPU micro-benchmark: multi-pick FPS, (cluster, warps per CTA) shapes that hold the same number of
groups on fewer SMs.  Checks identical indices against the default shape.

    python tools/fps_warps_sweep.py [batch]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point
from fps_multi_sweep import timeit


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    lib = hf.load()
    dev = torch.device("cuda:0")
    for n, m in [(16384, 4096), (4096, 1024), (1024, 256), (65536, 16384)]:
        x = torch.from_numpy(synthetic.make_scans(3, batch, max(n, 16384))[0]).to(dev)
        if n < 16384:
            x = gather_point(x, farthest_point_sample(n, x))
        ref = farthest_point_sample(m, x)
        for cluster, warps in [(0, 0), (8, 4), (4, 4), (4, 8), (2, 8), (2, 16), (1, 16), (1, 8)]:
            lib.hf_fps_set_variant(3, warps, 0)
            lib.hf_fps_set_cluster(cluster)
            try:
                out = farthest_point_sample(m, x)
                ms = timeit(lambda: farthest_point_sample(m, x))
                print("n=%6d m=%5d b=%2d cluster=%d warps=%2d  %8.4f ms  same=%s" % (
                    n, m, batch, cluster, warps, ms, bool(torch.equal(out, ref))))
            except Exception as ex:  # noqa: BLE001
                print("n=%d cluster=%d warps=%d: %s" % (n, cluster, warps, str(ex)[:90]))
            finally:
                lib.hf_fps_set_cluster(0)
                lib.hf_fps_set_variant(0, 0, 0)
            sys.stdout.flush()


if __name__ == "__main__":
    main()
