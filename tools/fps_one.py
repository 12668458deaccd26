#!/usr/bin/env python
"""One FPS call per level shape (for ncu):  python tools/fps_one.py [n m algo w j]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample

n, m, algo, w, j = (int(a) for a in (sys.argv[1:6] + ["16384", "4096", "2", "0", "0"][len(sys.argv) - 1:]))
lib = hf.load()
xyz, _ = synthetic.make_scans(3, 16, n)
x = torch.from_numpy(xyz).cuda()
lib.hf_fps_set_variant(algo, w, j)
for _ in range(3):
    out = farthest_point_sample(m, x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
out = farthest_point_sample(m, x)
e1.record()
torch.cuda.synchronize()
print("n=%d m=%d algo=%d w=%d j=%d: %.4f ms, %.1f ns/iter" % (n, m, algo, w, j, e0.elapsed_time(e1), e0.elapsed_time(e1) * 1e6 / (m - 1)))
