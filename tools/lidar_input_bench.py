#!/usr/bin/env python
"""Times hf_lidar_fov_filter (csrc/lidar_input.cu) on synthetic scans of KITTI size against its HBM roofline and the
numpy float64 path of the oracle on the host.  python tools/lidar_input_bench.py"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from heterofusionrcnn_b200 import obj_utils  # noqa: E402
from test_gpu_lidar_input import synthetic_scans  # noqa: E402

dev = torch.device("cuda:0")
peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
for scans, n in [(1, 122880), (16, 122880), (64, 122880)]:
    pts, calibs, im = synthetic_scans(1, scans, n)
    packed = torch.from_numpy(np.stack([obj_utils.pack_calib(*c) for c in calibs])).to(dev)
    d = torch.from_numpy(pts).to(dev)
    for _ in range(5):
        out, count = obj_utils.lidar_fov_filter(d, packed, im)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 50
    e0.record()
    for _ in range(reps):
        out, count = obj_utils.lidar_fov_filter(d, packed, im)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    kept = int(count.sum().item())
    algo = scans * n * 16 + kept * 32  # the scan read once, kept points read again and written
    t0 = time.perf_counter()
    from oracle import binding  # noqa: E402

    binding.lidar_fov_filter(pts[0], *calibs[0], im)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"scans": scans, "n": n, "kept": kept, "ms": round(ms, 4), "GB/s": round(algo / ms / 1e6, 1),
                      "scans_per_s": round(scans / ms * 1e3, 1), "numpy_ms_per_scan": round(cpu_ms, 2)}))
