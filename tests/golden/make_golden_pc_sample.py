#!/usr/bin/env python
"""Generates tests/golden/pc_sample_ref.npz by EXECUTING THE REFERENCE'S OWN LINES (run in the build container, where
/root/reference exists): the block `if pc_sample_pts < len(pts_rect): ... sampled_pc = np.hstack(...)` of
KittiDataset.load_samples (hf/datasets/kitti/kitti_dataset.py:341-371) is cut out of the source file, dedented and
exec'd on several clouds with `np` bound to a namespace whose `random.choice` / `random.shuffle` are the host-input
definitions of include/hfops.h:hf_pc_sample_points (keys / extra / slot_of) instead of numpy's global RNG; every other
name is numpy's.  Stored per case: the inputs, the three random inputs and the reference block's `sampled_pc` / `choice`.

    python tests/golden/make_golden_pc_sample.py
"""
import os
import textwrap
import types

import numpy as np

SRC = "/root/reference/hf/datasets/kitti/kitti_dataset.py"
lines = open(SRC).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.strip() == "if pc_sample_pts < len(pts_rect):")
end = next(i for i in range(start, len(lines)) if lines[i].strip().startswith("sampled_pc = np.hstack"))
block = textwrap.dedent("\n".join(lines[start:end + 1]))
assert start + 1 == 341 and end + 1 == 371, (start + 1, end + 1)  # the lines DESIGN.md cites


class StubRandom:
    """np.random for the block: draws defined by the caller's keys / extra / slot_of."""

    def __init__(self, keys, extra, slot_of):
        self.keys, self.extra, self.slot_of = keys, extra, slot_of

    def choice(self, a, size, replace=True):
        a = np.asarray(a)
        if size < 0:
            raise ValueError("negative dimensions are not allowed")
        if not replace:
            if size > len(a):
                raise ValueError("Cannot take a larger sample than population when 'replace=False'")
            order = np.lexsort((a, self.keys[a]))[:size]  # smallest (key, index) first
            return np.sort(a[order])                       # returned in ascending index order
        if len(a) == 0:
            raise ValueError("a must be non-empty")
        return a[(self.extra[:size] % np.uint32(len(a))).astype(np.int64)]

    def shuffle(self, x):
        old = x.copy()
        x[self.slot_of[: len(x)]] = old


def run_block(pts_rect, pts_intensity, pc_sample_pts, keys, extra, slot_of):
    fake_np = types.SimpleNamespace(**{k: getattr(np, k) for k in dir(np) if not k.startswith("__")})
    fake_np.random = StubRandom(keys, extra, slot_of)
    ns = {"np": fake_np, "pts_rect": pts_rect, "pts_intensity": pts_intensity, "pc_sample_pts": pc_sample_pts, "len": len}
    exec(block, ns)
    return ns["sampled_pc"], np.asarray(ns["choice"])


def make_cloud(rng, n, far_frac):
    z = np.where(rng.random(n) < far_frac, rng.uniform(40.0, 70.0, n), rng.uniform(0.5, 40.0, n))
    z[:3] = [40.0, np.nextafter(np.float32(40.0), np.float32(0)), 39.5][: min(3, n)]  # the boundary itself is far
    pts = np.stack([rng.uniform(-30, 30, n), rng.uniform(-2, 3, n), z], axis=1).astype(np.float32)
    inten = rng.random((n, 1)).astype(np.float32)
    return pts, inten


cases = {}
rng = np.random.default_rng(2024)
# (n, P, far fraction, tie-heavy keys)
specs = [(5000, 2048, 0.1, False), (3000, 2048, 0.5, True), (2048, 2048, 0.2, False), (1500, 2048, 0.3, False),
         (700, 2048, 0.3, True), (2100, 2048, 0.0, False), (40000, 16384, 0.12, False), (9000, 16384, 0.2, False),
         (33, 16, 0.4, True), (1, 4, 0.0, False), (2049, 2048, 1.0, False)]
for ci, (n, P, far, ties) in enumerate(specs):
    pts, inten = make_cloud(rng, n, far)
    keys = rng.integers(0, 8 if ties else 2 ** 32, n, dtype=np.uint32)
    extra = rng.integers(0, 2 ** 32, P, dtype=np.uint32)
    slot_of = rng.permutation(P).astype(np.int32)
    try:
        sampled, choice = run_block(pts.astype(np.float64), inten.astype(np.float64), P, keys, extra, slot_of)
        ok = 1
    except ValueError:
        sampled, choice, ok = np.zeros((P, 4)), np.full((P,), -1), 0
    cases["c%d" % ci] = dict(pts=pts, inten=inten, P=P, keys=keys, extra=extra, slot_of=slot_of, sampled=sampled,
                             choice=choice.astype(np.int32), ok=ok)
    print("case %d: n=%d P=%d far=%.2f ties=%s -> ok=%d %s %s" % (ci, n, P, far, ties, ok, sampled.shape, sampled.dtype))
flat = {"%s.%s" % (c, k): np.asarray(v) for c, d in cases.items() for k, v in d.items()}
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pc_sample_ref.npz")
np.savez_compressed(out, **flat)
print("wrote", out, os.path.getsize(out), "bytes")
