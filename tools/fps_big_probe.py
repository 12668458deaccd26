#!/usr/bin/env python
"""GPU probe: FPS 65536 -> 16384 (BASELINE configs[4] cloud size), shared-memory shape (4 CTAs x 16 warps, default)
against the register shape (8 CTAs x 8 warps, hf_fps_set_variant(3, 8, 0)) for 8 / 15 / 16 scans.

    python tools/fps_big_probe.py
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
    x = torch.from_numpy(synthetic.make_scans(3, 16, 65536)[0]).to(dev)
    for b in (1, 8, 15, 16):
        xb = x[:b].contiguous()
        ref = None
        for name, warps in (("4 CTAs x 16 warps, xyz in shared memory", 16), ("8 CTAs x 8 warps, xyz in registers", 8)):
            lib.hf_fps_set_variant(3, warps, 0)
            try:
                out = farthest_point_sample(16384, xb)
                ms = timeit(lambda: farthest_point_sample(16384, xb), reps=3)
            finally:
                lib.hf_fps_set_variant(0, 0, 0)
            same = True if ref is None else bool(torch.equal(out, ref))
            ref = out if ref is None else ref
            print("n=65536 m=16384 b=%2d  %-42s %8.3f ms  same=%s" % (b, name, ms, same))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
