#!/usr/bin/env python
"""GPU probe: the overlapped step (graph replay, batch 16 x 16384) against the shape of the level-1 FPS: plain / over the
input's spatial index, 4 CTAs x 8 warps / 8 CTAs x 4 warps per scan.

    python tools/fps_shape_in_step.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import chain


def main():
    lib = hf.load()
    dev = torch.device("cuda:0")
    base = chain.ChainConfig()
    res = chain.to_device(chain.make_host_inputs(base, batch=16, seed=2), dev)
    for name, sorted_fps, warps in (("plain, 4 x 8 warps (default)", False, 0), ("plain, 8 x 4 warps", False, 4),
                                    ("index, 4 x 8 warps", True, 8), ("index, 8 x 4 warps", True, 4)):
        cfg = chain.ChainConfig(**{**base.__dict__})
        cfg.sorted_fps, cfg.fps_level1_warps = sorted_fps, warps
        for with_rest in (True, False):
            cfg.with_crop = cfg.with_nms = with_rest
            g = chain.ChainGraph(res, cfg)
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            print("%-30s %-18s %.4f ms per step" % (name, "whole chain" if with_rest else "SA/FP chain only", e0.elapsed_time(e1) / 10))
            del g


if __name__ == "__main__":
    main()
