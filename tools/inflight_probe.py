#!/usr/bin/env python
"""GPU probe: resident throughput of the chain with 1, 2 and 3 batches in flight, each batch a captured CUDA graph on its
own stream set (chain.ChainGraph(slot=k)) and its own input / output buffers.

    python tools/inflight_probe.py [steps]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import chain


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    hf.load()
    dev = torch.device("cuda:0")
    cfg = chain.ChainConfig()
    host = chain.make_host_inputs(cfg, batch=16, seed=2)
    for depth in (1, 2, 3):
        sets = [chain.to_device(host, dev) for _ in range(depth)]
        graphs = [chain.ChainGraph(sets[k], cfg, slot=k) for k in range(depth)]
        streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
        main_s = torch.cuda.current_stream(dev)

        def run(n):
            for st in streams:
                st.wait_stream(main_s)
            for it in range(n):
                with torch.cuda.stream(streams[it % depth]):
                    graphs[it % depth].replay()
            for st in streams:
                main_s.wait_stream(st)

        run(2 * depth)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(steps)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print("batches in flight %d: %.4f ms per batch of 16 scans = %.0f scans/s" % (depth, ms, 16 / ms * 1e3))
        del graphs, sets
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
