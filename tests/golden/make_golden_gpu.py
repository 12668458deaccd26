"""Generate tests/golden/ref_cuda_b200.npz ON THE GPU BOX from the reference's own kernels.

    gpurun -- 'python tests/golden/make_golden_gpu.py'   (writes gpurun_out/ref_cuda_b200.npz)
    cp gpurun_out/ref_cuda_b200.npz tests/golden/

The outputs are produced by oracle/_ref/libref_cuda.so = the five reference *_g.cu files
compiled UNMODIFIED for sm_100a (oracle/Makefile) and run on a B200.  The CPU tests then
pin the oracle against these vectors (tests/test_oracle_golden.py).  Inputs are regenerated
from seeds by the test, only outputs are stored (int16/int32/float32, a few hundred KB).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import _refcuda  # noqa: E402
from golden_cases import CASES, inputs_for  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    out = {}
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    N = lambda t: t.cpu().numpy()  # noqa: E731
    for name, case in CASES.items():
        inp = inputs_for(name)
        op = case["op"]
        if op == "fps":
            out[name] = N(_refcuda.fps(case["m"], T(inp["xyz"])))
        elif op == "ball":
            idx, cnt = _refcuda.query_ball_point(case["r"], case["ns"], T(inp["xyz"]), T(inp["query"]))
            idx = N(idx)
            idx[N(cnt) == 0] = 0  # uninitialised rows in the reference
            out[name + ".idx"], out[name + ".cnt"] = idx, N(cnt)
        elif op == "three_nn":
            d, i = _refcuda.three_nn(T(inp["unknown"]), T(inp["known"]))
            out[name + ".dist"], out[name + ".idx"] = N(d), N(i)
        elif op == "interp":
            o = _refcuda.three_interpolate_channel_first(T(inp["points"].transpose(0, 2, 1)), T(inp["idx"]), T(inp["w"]))
            out[name] = N(o).transpose(0, 2, 1)
        elif op == "crop":
            r = _refcuda.pc_crop_and_sample(T(inp["xyz"]), T(inp["fts"]), T(inp["inten"]), T(inp["mask"]),
                                            T(inp["boxes"]), T(inp["box_ind"]), case["R"])
            ind = N(r[4])
            out[name + ".non_empty"] = N(r[5])
            # race-free summary: per box, the sorted unique cropped indices (padded with -1)
            uniq = np.full(ind.shape, -1, np.int32)
            for b in range(ind.shape[0]):
                u = np.unique(ind[b]) if N(r[5])[b] else np.zeros(0, np.int32)
                uniq[b, : len(u)] = u
            out[name + ".uniq"] = uniq
        elif op == "iou":
            o, i = _refcuda.compute_bev_iou(T(inp["a"]), T(inp["b"]))
            out[name + ".overlap"], out[name + ".iou"] = N(o), N(i)
        elif op == "nms":
            out[name + ".mask"] = N(_refcuda.nms_mask(T(inp["boxes"]), case["thresh"]))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "ref_cuda_b200.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
