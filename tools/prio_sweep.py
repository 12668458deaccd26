#!/usr/bin/env python
"""GPU probe: the overlapped step (graph replay, batch 16 x 16384) against the stream priorities of the chain's lanes
(crit, group, nn/fp, crop, nms; -3 = highest on a B200) with the level-1 FPS plain or over the input's spatial index.

    python tools/prio_sweep.py
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
    for sorted_fps in (False, True):
        for prio in ("-3,-2,-2,0,-1", "-3,-2,-2,-1,0", "-3,-2,-2,-1,-1", "-3,-2,-2,-2,0", "-3,-1,-1,-2,0", "-3,-2,-2,0,0",
                     "-3,-3,-3,-1,0", "-3,-3,-3,-2,-1"):
            os.environ["HF_CHAIN_PRIORITIES"] = prio
            chain._Lanes._cache.clear()
            cfg = chain.ChainConfig(**{**base.__dict__})
            cfg.sorted_fps = sorted_fps
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
            print("fps %-6s priorities %-16s %.4f ms per step" % ("index" if sorted_fps else "plain", prio, e0.elapsed_time(e1) / 10))
            sys.stdout.flush()
            del g


if __name__ == "__main__":
    main()
