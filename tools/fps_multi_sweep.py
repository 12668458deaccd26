#!/usr/bin/env python
"""GPU micro-benchmark: multi-pick FPS (fps_multi.cu, algo 3) against the one-pick cluster kernel (algo 1).

    python tools/fps_multi_sweep.py [batch]
Prints ms per call and ns per pick for every level of the chain and every cluster size; checks that
all variants give identical indices.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point


def timeit(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    lib = hf.load()
    dev = torch.device("cuda:0")
    for n, m, b in [(16384, 4096, batch), (4096, 1024, batch), (1024, 256, batch), (256, 64, batch),
                    (65536, 16384, 2), (65536, 16384, 16), (16384, 4096, 64)]:
        xyz, _ = synthetic.make_scans(3, b, n)
        x = torch.from_numpy(xyz).to(dev)
        if n < 16384:  # deeper levels of the chain see a cloud in FPS order: build it the way the chain does
            x0 = torch.from_numpy(synthetic.make_scans(3, b, 16384)[0]).to(dev)
            x = gather_point(x0, farthest_point_sample(n, x0))
        ref = None
        for algo, cluster in [(1, 0), (3, 0), (3, 1), (3, 2), (3, 4), (3, 8)]:
            lib.hf_fps_set_variant(algo, 0, 0)
            lib.hf_fps_set_cluster(cluster)
            try:
                out = farthest_point_sample(m, x)
                ms = timeit(lambda: farthest_point_sample(m, x))
            except Exception as ex:  # noqa: BLE001
                print("n=%d m=%d algo=%d cluster=%d: %s" % (n, m, algo, cluster, str(ex)[:80]))
                continue
            finally:
                lib.hf_fps_set_cluster(0)
                lib.hf_fps_set_variant(0, 0, 0)
            if ref is None:
                ref = out
            print("n=%6d m=%5d b=%2d algo=%d cluster=%d  %8.4f ms  %7.1f ns/pick  same=%s" % (
                n, m, b, algo, cluster, ms, ms * 1e6 / (m - 1), bool(torch.equal(out, ref))))
            sys.stdout.flush()


if __name__ == "__main__":
    main()
