"""Seeded inputs of the golden cases shared by tests/golden/make_golden_gpu.py (which runs the
reference CUDA kernels on a B200 and stores their outputs) and tests/test_oracle_golden.py
(which checks the CPU oracle against those stored outputs)."""
import numpy as np

from heterofusionrcnn_b200 import synthetic

CASES = {
    "fps_4096_1024": dict(op="fps", b=2, n=4096, m=1024, dup=0),
    "fps_dup_3000_400": dict(op="fps", b=2, n=3000, m=400, dup=1200),
    "fps_small_100_100": dict(op="fps", b=1, n=100, m=100, dup=40),
    "ball_r02": dict(op="ball", b=2, n=4096, m=256, r=0.2, ns=32),
    "ball_r20": dict(op="ball", b=2, n=4096, m=256, r=2.0, ns=16),
    "three_nn_a": dict(op="three_nn", b=2, n=2048, m=512),
    "interp_a": dict(op="interp", b=2, n=512, m=128, c=32),
    "crop_a": dict(op="crop", B=2, P=4096, C=8, nbox=24, R=256),
    "iou_a": dict(op="iou", n=96, m=80),
    "nms_a": dict(op="nms", n=300, thresh=0.5),
}


def _seed(name):
    return sum(ord(ch) * (i + 1) for i, ch in enumerate(name)) % 100000


def inputs_for(name):
    c = CASES[name]
    s = _seed(name)
    rng = np.random.default_rng(s)
    op = c["op"]
    if op == "fps":
        xyz, _ = synthetic.make_scans(s, c["b"], c["n"], duplicate_pad=c["dup"])
        return dict(xyz=xyz)
    if op == "ball":
        xyz, _ = synthetic.make_scans(s, c["b"], c["n"])
        q = np.ascontiguousarray(xyz[:, :: c["n"] // c["m"]][:, : c["m"]])
        return dict(xyz=xyz, query=q)
    if op == "three_nn":
        xyz, _ = synthetic.make_scans(s, c["b"], c["n"], duplicate_pad=c["n"] // 10)
        return dict(unknown=xyz, known=xyz[:, ::-1][:, : c["m"]].copy())
    if op == "interp":
        pts = rng.standard_normal((c["b"], c["m"], c["c"])).astype(np.float32)
        idx = rng.integers(0, c["m"], (c["b"], c["n"], 3)).astype(np.int32)
        w = rng.random((c["b"], c["n"], 3)).astype(np.float32)
        w /= w.sum(-1, keepdims=True)
        return dict(points=pts, idx=idx, w=w)
    if op == "crop":
        xyz, inten = synthetic.make_scans(s, c["B"], c["P"])
        fts = rng.standard_normal((c["B"], c["P"], c["C"])).astype(np.float32)
        mask = rng.random((c["B"], c["P"])) < 0.3
        b3 = [synthetic.make_boxes3d(s * 13 + i, xyz[i], c["nbox"]) for i in range(c["B"])]
        b3[0][:2, 3:6] *= 3.0
        b3[-1][0, 0] += 500.0
        boxes = np.concatenate([synthetic.boxes3d_to_corners(x) for x in b3])
        return dict(xyz=xyz, fts=fts, inten=inten, mask=mask, boxes=boxes,
                    box_ind=np.repeat(np.arange(c["B"], dtype=np.int32), c["nbox"]))
    if op == "iou":
        bx = synthetic.make_nms_boxes(s, c["n"] + c["m"], clusters=4)
        return dict(a=bx[: c["n"]], b=bx[c["n"]:])
    if op == "nms":
        return dict(boxes=synthetic.make_nms_boxes(s, c["n"], clusters=4))
    raise KeyError(op)
