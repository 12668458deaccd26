#!/usr/bin/env python
"""GPU probe: what each branch of the chain costs INSIDE the overlapped step (graph replay, batch 16 x 16384): the step
with the crop / the NMS / both left out, and with the sampling spine alone.

    python tools/ablation_probe.py
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
    host = chain.make_host_inputs(base, batch=16, seed=2)
    resident = chain.to_device(host, dev)
    for name, crop, nms, pyr in (("whole chain", True, True, 4), ("whole chain, pyramid of levels 2-4", True, True, 1),
                                 ("without the NMS", True, False, 4), ("without the crop", False, True, 4),
                                 ("SA/FP chain only", False, False, 4), ("SA/FP chain only, pyramid of levels 2-4", False, False, 1)):
        cfg = chain.ChainConfig(**{**base.__dict__})
        cfg.with_crop, cfg.with_nms, cfg.pyramid_from = crop, nms, pyr
        g = chain.ChainGraph(resident, cfg)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        print("%-42s %.4f ms per step (%d launches)" % (name, e0.elapsed_time(e1) / 10, g.launches))
        del g


if __name__ == "__main__":
    main()
