#!/usr/bin/env python
"""GPU micro-benchmark of the HBM-bound gathers at the chain's shapes (batch 16): group_point,
three_interpolate; 50 back-to-back launches between two CUDA events (the per-launch event pair of bench.py
adds ~7 us to a 40 us kernel), algorithmic bytes of SURVEY.md section 8d, fraction of the measured HBM peak.

    python tools/gather_bench.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200.grouping.tf_grouping import group_point
from heterofusionrcnn_b200.interpolate.tf_interpolate import three_interpolate
from fps_multi_sweep import timeit


def main():
    dev = torch.device("cuda:0")
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6553.0)
    b, ns = 16, 32
    lib = hf.load()
    for mode in (0, 1):
        lib.hf_gather_set_mode(mode)
        print("# hf_gather_set_mode(%d): group_point = %s" % (mode, "TMA bulk row mover" if mode else "register kernel"))
        run(dev, peak, b, ns)
    lib.hf_gather_set_mode(0)


def run(dev, peak, b, ns):
    big = torch.empty(512 * 1024 * 1024 // 4, device=dev)  # L2 flush between groups of launches
    for n, m, c in [(4096, 1024, 96), (1024, 256, 256), (256, 64, 512)]:
        pts = torch.randn((b, n, c), device=dev)
        idx = torch.randint(0, n, (b, m, ns), device=dev, dtype=torch.int32)
        big.zero_()
        ms = timeit(lambda: group_point(pts, idx), reps=50)
        nbytes = 4 * (b * m * ns * c + b * m * ns + b * n * c)
        print("group_point       n=%5d m=%4d c=%4d  %7.2f us  %7.1f GB/s  %.2f of peak" % (n, m, c, ms * 1e3, nbytes / ms / 1e6, nbytes / ms / 1e6 / peak))
    for n, m, c in [(16384, 4096, 128), (4096, 1024, 512), (1024, 256, 512), (256, 64, 1024)]:
        pts = torch.randn((b, m, c), device=dev)
        idx = torch.randint(0, m, (b, n, 3), device=dev, dtype=torch.int32)
        w = torch.rand((b, n, 3), device=dev)
        big.zero_()
        ms = timeit(lambda: three_interpolate(pts, idx, w), reps=50)
        nbytes = 4 * (b * n * c + b * m * c + 6 * b * n)
        print("three_interpolate n=%5d m=%4d c=%4d  %7.2f us  %7.1f GB/s  %.2f of peak" % (n, m, c, ms * 1e3, nbytes / ms / 1e6, nbytes / ms / 1e6 / peak))
    # real geometry (random indices have no locality to exploit): FP levels of the chain, memory order vs curve order
    from heterofusionrcnn_b200 import spatial, synthetic
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn_weights
    from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample_with_xyz
    xyz0 = torch.from_numpy(synthetic.make_scans(2, b, 16384)[0]).to(dev)
    _, xyz1 = farthest_point_sample_with_xyz(4096, xyz0)
    _, xyz2 = farthest_point_sample_with_xyz(1024, xyz1)
    for unk, kn, c in [(xyz0, xyz1, 128), (xyz1, xyz2, 512)]:
        n, m = unk.shape[1], kn.shape[1]
        _, nidx, w = three_nn_weights(unk, kn)
        pts = torch.randn((b, m, c), device=dev)
        nbytes = 4 * (b * n * c + b * m * c + 6 * b * n)
        for name, kw in (("three_nn indices", {}),):
            big.zero_()
            ms = timeit(lambda: three_interpolate(pts, nidx, w, **kw), reps=50)
            print("three_interpolate (chain geometry, %s) n=%5d m=%4d c=%4d  %7.2f us  %7.1f GB/s  %.2f of peak" % (
                name, n, m, c, ms * 1e3, nbytes / ms / 1e6, nbytes / ms / 1e6 / peak))


if __name__ == "__main__":
    main()
