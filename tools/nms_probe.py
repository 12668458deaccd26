#!/usr/bin/env python
"""GPU: time hf_oriented_nms on the BASELINE config-4 workload (16 scenes x 9000 boxes) and report
how many pairs reach the exact clip.   python tools/nms_probe.py [scenes] [boxes] [thresh]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic

S = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N = int(sys.argv[2]) if len(sys.argv) > 2 else 9000
TH = float(sys.argv[3]) if len(sys.argv) > 3 else 0.85
lib = hf.load()
boxes = torch.from_numpy(np.stack([synthetic.make_nms_boxes(34 + 17 * i, N) for i in range(S)])).cuda()
keep = torch.empty((S, N), dtype=torch.int32, device="cuda")
num = torch.empty((S,), dtype=torch.int32, device="cuda")
wsb = lib.hf_oriented_nms_workspace_bytes(S, N)
ws = torch.empty((wsb // 8 + 1,), dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for sort in (0, 1):
    lib.hf_nms_set_spatial_sort(sort)
    for _ in range(3):
        rc = lib.hf_oriented_nms(boxes.data_ptr(), S, N, TH, keep.data_ptr(), num.data_ptr(), ws.data_ptr(), ws.numel() * 8, st)
        assert rc == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        lib.hf_oriented_nms(boxes.data_ptr(), S, N, TH, keep.data_ptr(), num.data_ptr(), ws.data_ptr(), ws.numel() * 8, st)
    e1.record()
    torch.cuda.synchronize()
    pairs = lib.hf_oriented_nms_last_pair_count(S, N, ws.data_ptr(), st)
    print("spatial_sort=%d: %.3f ms per call, %d pairs clipped (%.0f per scene, %.3f%% of N(N-1)/2), kept %s" % (
        sort, e0.elapsed_time(e1) / 5, pairs, pairs / S, 100.0 * pairs / S / (N * (N - 1) / 2), num[:4].tolist()))
