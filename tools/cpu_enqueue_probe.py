#!/usr/bin/env python
"""GPU: host time to enqueue one chain pass vs device time (is the step launch bound?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import chain
hf.load()
cfg = chain.ChainConfig()
dev = torch.device("cuda:0")
res = chain.to_device(chain.make_host_inputs(cfg, 16, 2), dev)
for _ in range(3):
    chain.run_chain(res, cfg)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    chain.run_chain(res, cfg)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host enqueue %.3f ms/step, total %.3f ms/step" % ((t1 - t0) * 100, (t2 - t0) * 100))
