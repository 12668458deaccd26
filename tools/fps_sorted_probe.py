#!/usr/bin/env python
"""GPU probe: FPS with and without the cloud's spatial index (warps skip the centres that cannot reach them).

    python tools/fps_sorted_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import spatial, synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample_with_xyz
from fps_multi_sweep import timeit


def main():
    hf.load()
    dev = torch.device("cuda:0")
    for n, m, batches in ((16384, 4096, (8, 15, 16)), (65536, 16384, (8, 15, 16))):
        x = torch.from_numpy(synthetic.make_scans(3, 16, n)[0]).to(dev)
        for b in batches:
            xb = x[:b].contiguous()
            ix = spatial.build_cloud_index(xb)
            t_ix = timeit(lambda: spatial.build_cloud_index(xb), reps=5)
            ref = farthest_point_sample_with_xyz(m, xb)[0]
            out = farthest_point_sample_with_xyz(m, xb, index=ix)[0]
            t0 = timeit(lambda: farthest_point_sample_with_xyz(m, xb), reps=3)
            t1 = timeit(lambda: farthest_point_sample_with_xyz(m, xb, index=ix), reps=3)
            print("n=%5d m=%5d b=%2d  plain %8.3f ms   with index %8.3f ms (+ index build %6.3f ms)  same=%s" % (
                n, m, b, t0, t1, t_ix, bool(torch.equal(ref, out))))
            if n <= 16384:
                lib = hf.load()
                for w, mode, name in ((4, 1, "warps: 8 CTAs x 4 warps x 16"), (8, 1, "warps: 4 CTAs x 8 warps x 16"),
                                      (4, 2, "threads: 8 CTAs x 4 warps x 16"), (8, 2, "threads: 4 CTAs x 8 warps x 16"),
                                      (6, 2, "threads: 6 CTAs x 8 warps x 11")):
                    lib.hf_fps_set_variant(3, w, mode)
                    try:
                        o2 = farthest_point_sample_with_xyz(m, xb, index=ix)[0]
                        t2 = timeit(lambda: farthest_point_sample_with_xyz(m, xb, index=ix), reps=3)
                    finally:
                        lib.hf_fps_set_variant(0, 0, 0)
                    print("        with index, compact %-32s %8.3f ms  same=%s" % (name, t2, bool(torch.equal(ref, o2))))
            else:
                lib = hf.load()
                lib.hf_fps_set_variant(3, 0, 1)
                try:
                    o2 = farthest_point_sample_with_xyz(m, xb, index=ix)[0]
                    t2 = timeit(lambda: farthest_point_sample_with_xyz(m, xb, index=ix), reps=3)
                finally:
                    lib.hf_fps_set_variant(0, 0, 0)
                print("        with index, compact warps (8 x 8 x 32)       %8.3f ms  same=%s" % (t2, bool(torch.equal(ref, o2))))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
