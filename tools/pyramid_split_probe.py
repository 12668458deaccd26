#!/usr/bin/env python
"""GPU probe: the overlapped step (graph replay, batch 16 x 16384) against ChainConfig.pyramid_from (the first 0-based
level whose sampling goes into the pyramid launch) and nms_early_stop; 3 graph builds each.

    python tools/pyramid_split_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import chain


def main():
    hf.load()
    dev = torch.device("cuda:0")
    base = chain.ChainConfig()
    res = chain.to_device(chain.make_host_inputs(base, batch=16, seed=2), dev)
    for early in (True, False):
        for pf in (1, 2, 3, 4):
            out = []
            for rep in range(3):
                cfg = chain.ChainConfig(**{**base.__dict__})
                cfg.pyramid_from, cfg.nms_early_stop = pf, early
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
                out.append(e0.elapsed_time(e1) / 10)
                n = g.launches
                del g
            print("nms_early_stop=%-5s pyramid_from=%d (%d launches): %s ms per step" % (early, pf, n, " ".join("%.4f" % v for v in out)))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
