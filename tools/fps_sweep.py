#!/usr/bin/env python
"""GPU micro-benchmark: FPS launch shapes (cluster size x CTA threads) per level of the chain.

    python tools/fps_sweep.py [batch]
Prints ms per call and ns per sequential iteration; verifies every shape gives identical indices.
"""
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import _lib, synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample


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
    for n, m in [(16384, 4096), (4096, 1024), (1024, 256), (256, 64), (65536, 16384)]:
        b = batch if n <= 16384 else 2
        xyz, _ = synthetic.make_scans(3, b, n)
        x = torch.from_numpy(xyz).to(dev)
        ref = None
        for variant, cluster, threads in [(v, c, th) for v in (0,) for c in (0, 1, 2, 4, 8) for th in (256, 512)]:
            if True:
                if n // max(cluster, 1) > 8192 and cluster != 0:
                    continue
                lib.hf_fps_set_variant(1, 0, 0)  # plain on-chip kernel
                lib.hf_fps_set_cluster(cluster)
                lib.hf_fps_set_threads(threads)
                try:
                    out = farthest_point_sample(m, x)
                    ms = timeit(lambda: farthest_point_sample(m, x))
                except Exception as ex:  # noqa: BLE001
                    print("n=%d m=%d cluster=%d threads=%d: %s" % (n, m, cluster, threads, ex))
                    continue
                if ref is None:
                    ref = out
                same = bool(torch.equal(out, ref))
                print("n=%6d m=%5d b=%2d variant=%d cluster=%d threads=%3d  %8.4f ms  %7.1f ns/iter  same=%s" % (
                    n, m, b, variant, cluster, threads, ms, ms * 1e6 / (m - 1), same))
        lib.hf_fps_set_cluster(0)
        lib.hf_fps_set_threads(0)
        lib.hf_fps_set_variant(0, 0, 0)
        # bucket-pruned kernel (fps_bucket.cu): every shape whose capacity covers the cloud
        per = n
        while per > 16384:
            per = (per + 1) // 2
        for w, j in [(0, 0)] + [(w, j) for w in (2, 4, 8, 16, 32) for j in (1, 2, 4, 8, 16, 32)]:
            if w and (w * 32 * j < per or w * 32 * j > 4 * max(per, 256)):
                continue
            lib.hf_fps_set_variant(2, w, j)
            try:
                out = farthest_point_sample(m, x)
                ms = timeit(lambda: farthest_point_sample(m, x))
            except Exception as ex:  # noqa: BLE001
                continue
            finally:
                lib.hf_fps_set_variant(0, 0, 0)
            print("n=%6d m=%5d b=%2d bucket w=%2d j=%2d  %8.4f ms  %7.1f ns/iter  same=%s" % (
                n, m, b, w, j, ms, ms * 1e6 / (m - 1), bool(torch.equal(out, ref))))


if __name__ == "__main__":
    main()
