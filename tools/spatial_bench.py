#!/usr/bin/env python
"""GPU micro-benchmark: pruned (csrc/spatial.cu) vs every-pair ball query and three_nn on the chain's shapes.

    python tools/spatial_bench.py [batch]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import heterofusionrcnn_b200 as hf
from heterofusionrcnn_b200 import synthetic
from heterofusionrcnn_b200.grouping.tf_grouping import query_ball_point
from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point


def timeit(fn, reps=10):
    for _ in range(3):
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
    x = [torch.from_numpy(synthetic.make_scans(2, batch, 16384)[0]).to(dev)]
    for m in (4096, 1024, 256):
        x.append(gather_point(x[-1], farthest_point_sample(m, x[-1])))
    for lvl, radii in [(0, (0.1, 0.5, 4.0)), (1, (0.2, 1.0)), (2, (0.4, 2.0))]:
        a, q = x[lvl], x[lvl + 1]
        for mode, name in [(1, "every-pair"), (2, "pruned")]:
            lib.hf_spatial_set_mode(mode)
            ms = timeit(lambda: three_nn(a, q))
            print("level %d  three_nn  n=%5d m=%5d            %-10s %8.4f ms" % (lvl + 1, a.shape[1], q.shape[1], name, ms))
            for r in radii:
                ms = timeit(lambda: query_ball_point(r, 32, a, q))
                cnt = query_ball_point(r, 32, a, q)[1].float().mean().item()
                print("level %d  ball      n=%5d m=%5d r=%4.1f     %-10s %8.4f ms  (mean cnt %.1f)" % (
                    lvl + 1, a.shape[1], q.shape[1], r, name, ms, cnt))
            sys.stdout.flush()
    lib.hf_spatial_set_mode(2)
    # prebuilt indices (what chain.run_chain does): build time per cloud, then the pruned kernels alone
    from heterofusionrcnn_b200 import spatial
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_nn_weights
    ix = [spatial.build_cloud_index(c) for c in x]
    for lvl, c in enumerate(x):
        ms = timeit(lambda: spatial.build_cloud_index(c))
        print("cloud_index level %d n=%5d  %8.4f ms" % (lvl, c.shape[1], ms))
    for lvl, r in [(0, 0.1), (1, 0.2), (2, 0.4)]:
        a, q = x[lvl], x[lvl + 1]
        ms = timeit(lambda: three_nn_weights(a, q, index1=ix[lvl], index2=ix[lvl + 1]))
        print("level %d  three_nn  n=%5d m=%5d            prebuilt   %8.4f ms" % (lvl + 1, a.shape[1], q.shape[1], ms))
        ms = timeit(lambda: query_ball_point(r, 32, a, q, index1=ix[lvl], index2=ix[lvl + 1]))
        print("level %d  ball      n=%5d m=%5d r=%4.1f     prebuilt   %8.4f ms" % (lvl + 1, a.shape[1], q.shape[1], r, ms))
    lib.hf_spatial_set_mode(0)


if __name__ == "__main__":
    main()
