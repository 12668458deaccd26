"""The operator chain the benchmark times: the calls hf/core makes into the five op modules.

Per batch of scans (BASELINE.json configs 2-4):
  SA level l = 1..4   farthest_point_sample -> gather_point -> query_ball_point ->
                      group_point(xyz) -> group_point(features)
                      (hf/core/feature_extractors/pointnet_util.py:42-56)
  FP level l = 4..1   three_nn -> inverse-distance weights -> three_interpolate
                      (pointnet_util.py:303-308)
  RoI crop            pc_crop_and_sample on 512 proposals per scan (rcnn_model.py:478-489)
  NMS                 oriented_nms on 9000 score-sorted BEV boxes per scan, first 512 kept
                      (compute_iou.py:67-80, model_util.py:101-142)
The MLPs between the ops are dense NN layers outside this path; their outputs are
stood in for by resident synthetic feature tensors of the right shapes.
"""
from dataclasses import dataclass, field
from typing import List

import numpy as np
import torch

from . import spatial, synthetic
from .bev_iou.bev_iou import oriented_nms_batch
from .cropping.tf_cropping import pc_crop_and_sample
from .grouping.tf_grouping import group_point, group_point_pair, query_ball_point
from .interpolate.tf_interpolate import three_interpolate, three_nn_weights
from .sampling.tf_sampling import (PYRAMID_MAX_LEVELS, farthest_point_sample, farthest_point_sample_pyramid,
                                   farthest_point_sample_with_xyz, gather_point)


@dataclass
class ChainConfig:
    points: int = 16384
    npoint: List[int] = field(default_factory=lambda: [4096, 1024, 256, 64])
    radius: List[float] = field(default_factory=lambda: [0.1, 0.2, 0.4, 0.8])
    nsample: int = 32
    sa_channels: List[int] = field(default_factory=lambda: [1, 96, 256, 512])  # features grouped at SA l
    fp_channels: List[int] = field(default_factory=lambda: [1024, 512, 512, 128])  # interpolated at FP 4..1
    crop_boxes: int = 512
    crop_resize: int = 512
    crop_channels: int = 288
    nms_boxes: int = 9000
    nms_thresh: float = 0.85
    post_nms: int = 512
    with_crop: bool = True
    with_nms: bool = True
    # sampling spine: level 1 = one launch that writes the indices AND the picked points; levels 2-4 = one launch
    # (fps pyramid).  False = the reference's op sequence, farthest_point_sample + gather_point per level (8 launches).
    fused_sampling: bool = True
    # first level (0-based) whose sampling goes into ONE pyramid launch together with every deeper level (hf_fps_pyramid);
    # the levels before it get a launch each (indices + points).  4 = no pyramid launch (default): every level's points
    # exist as early as possible, and their ball query / three_nn / grouping - the tail behind the spine - start sooner.
    # Measured with the final kernels (tools/pyramid_split_probe.py): 1 (levels 2-4 fused) 1.98 ms per step, 2: 1.82,
    # 3 and 4: 1.80.  (With the round-1 FPS the fused pyramid was the better one by 0.03 ms: the spine was the bound.)
    pyramid_from: int = 4
    # grouping: the coordinates and the features of a level's neighbourhoods in one launch (group_point_pair);
    # False = two group_point launches per level (pointnet_util.py:49,56)
    fused_grouping: bool = True
    # level-1 FPS over the input's spatial index (threads skip the centres that cannot reach their bucket of the cloud):
    # None = whenever the index is built anyway.  16 x 16384 -> 4096: 1.25 -> 1.06 ms per launch (the SA/FP chain alone
    # 1.72 -> 1.59 ms per step, the whole step unchanged within noise: it is bound by the crop + NMS beside the spine);
    # 65536 -> 16384: 8.8 -> 5.6 ms (profiles/r2_fps_sorted_probe.txt, r2_fps_shape_in_step.txt)
    sorted_fps: object = None
    # the NMS sweep stops once post_nms boxes are kept (what sb_nms_fn's slice uses); False = the whole keep list
    nms_early_stop: bool = True
    # warps per CTA of the level-1 FPS clusters (0 = the library's heuristic: 8 CTAs x 4 warps up to 15 scans, 4 x 8
    # beyond; 4 forces 8 x 4 for any batch).  A tuning knob for tools/, not needed by callers.
    fps_level1_warps: int = 0

    def scaled(self, points):
        """Same chain on a different cloud size (npoint = n/4 per level)."""
        c = ChainConfig(**{**self.__dict__})
        c.points = points
        c.npoint = [points // 4, points // 16, points // 64, points // 256]
        return c


def make_host_inputs(cfg, batch, seed=2):
    """Host-side (numpy) inputs of one batch; `seed` follows SURVEY.md 8d (1000*config+scene)."""
    xyz, inten = synthetic.make_scans(seed, batch, cfg.points)
    rng = np.random.default_rng(seed * 7919 + 1)
    host = {"xyz": xyz, "intensity": inten}
    n_levels = [cfg.points] + list(cfg.npoint)
    host["sa_feats"] = [None] + [rng.standard_normal((batch, n_levels[l], cfg.sa_channels[l]), dtype=np.float32)
                                 for l in range(1, 4)]
    # FP level j (0 = coarsest): known points = level 4-j, channels fp_channels[j]
    host["fp_feats"] = [rng.standard_normal((batch, n_levels[4 - j], cfg.fp_channels[j]), dtype=np.float32)
                        for j in range(4)]
    if cfg.with_crop:
        host["crop_fts"] = rng.standard_normal((batch, cfg.points, cfg.crop_channels), dtype=np.float32)
        host["fg_mask"] = rng.random((batch, cfg.points)) < 0.3
        boxes = [synthetic.boxes3d_to_corners(synthetic.make_boxes3d(seed * 31 + i, xyz[i], cfg.crop_boxes))
                 for i in range(batch)]
        host["crop_boxes"] = np.concatenate(boxes, axis=0)
        host["box_ind"] = np.repeat(np.arange(batch, dtype=np.int32), cfg.crop_boxes)
    if cfg.with_nms:
        host["nms_boxes"] = np.stack([synthetic.make_nms_boxes(seed * 17 + i, cfg.nms_boxes) for i in range(batch)])
    return host


def _to_dev(a, device, pin):
    if a is None:
        return None
    t = torch.from_numpy(np.ascontiguousarray(a))
    if pin:
        t = t.pin_memory()
    return t.to(device, non_blocking=pin)


def to_device(host, device, pin=False):
    out = {}
    for k, v in host.items():
        out[k] = [_to_dev(x, device, pin) for x in v] if isinstance(v, list) else _to_dev(v, device, pin)
    return out


def to_pinned(host):
    """numpy inputs -> pinned host torch tensors (the host buffers a caller of the op API holds)."""
    def pin(a):
        return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return {k: [pin(x) for x in v] if isinstance(v, list) else pin(v) for k, v in host.items()}


def to_device_from_pinned(pinned, device):
    """One asynchronous host->device copy per input tensor on the current stream."""
    def up(t):
        return None if t is None else t.to(device, non_blocking=True)
    return {k: [up(x) for x in v] if isinstance(v, list) else up(v) for k, v in pinned.items()}


HOST_ORIGINATED = ("xyz", "intensity", "fg_mask", "crop_boxes", "box_ind", "nms_boxes")


class InputArena:
    """All input tensors of one step packed into ONE pinned host buffer, mirrored by device buffers of the same
    layout: a step's host->device transfer is one cudaMemcpyAsync instead of one per tensor (14 for the default
    chain).  The tensors a real caller holds on the host (scans, intensities, masks, boxes: HOST_ORIGINATED,
    ~8 MB per 16 scans) come first, the stand-ins for the MLP feature maps (which the reference produces on the
    device, hf/core/models/rcnn_model.py:478-489) after them, so either the whole arena or only the
    host-originated prefix can be sent with a single copy."""

    ALIGN = 256

    def __init__(self, host, pin=True):
        items = []
        for k, v in host.items():
            for i, a in (enumerate(v) if isinstance(v, list) else [(None, v)]):
                if a is not None:
                    items.append((k, i, np.ascontiguousarray(a)))
        items.sort(key=lambda it: 0 if it[0] in HOST_ORIGINATED else 1)  # stable: host-originated first
        self.layout, off, self.prefix = [], 0, 0
        for k, i, a in items:
            off = (off + self.ALIGN - 1) // self.ALIGN * self.ALIGN
            self.layout.append((k, i, tuple(a.shape), torch.from_numpy(a[:0].copy() if a.size else a).dtype, off, a.nbytes))
            off += a.nbytes
            if k in HOST_ORIGINATED:
                self.prefix = off
        self.nbytes = (off + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.prefix = (self.prefix + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.lists = {k: len(v) for k, v in host.items() if isinstance(v, list)}
        self.pinned = torch.empty((self.nbytes,), dtype=torch.uint8)
        if pin:  # (pin=False: layout tests on a machine without a CUDA driver)
            self.pinned = self.pinned.pin_memory()
        flat = self.pinned.numpy()
        for (k, i, shape, dt, o, nb), (_, _, a) in zip(self.layout, items):
            flat[o:o + nb] = a.reshape(-1).view(np.uint8)

    def device_buffer(self, device):
        return torch.empty((self.nbytes,), dtype=torch.uint8, device=device)

    def views(self, buf):
        """dict of tensors (same structure as make_host_inputs) viewing `buf` (a host or device arena)."""
        out = {k: [None] * n for k, n in self.lists.items()}
        for k, i, shape, dt, o, nb in self.layout:
            t = buf[o:o + nb].view(dt).view(shape)
            if i is None:
                out[k] = t
            else:
                out[k][i] = t
        return out

    def upload(self, dev_buf, host_originated_only=False):
        """ONE asynchronous host->device copy on the current stream; returns the bytes sent."""
        nb = self.prefix if host_originated_only else self.nbytes
        dev_buf[:nb].copy_(self.pinned[:nb], non_blocking=True)
        return nb


def host_results(res, cfg):
    """The results of one chain pass that leave the device in the reference pipeline: index
    outputs the host consumes (proposal keep lists, sampled indices, counts, empty-box flags).
    Grouped / interpolated / cropped feature tensors feed the next network layer on the GPU."""
    outs = list(res["fps_idx"]) + [res["sa%d" % (l + 1)][1] for l in range(4)]
    if cfg.with_nms:
        outs += [res["nms_keep"].contiguous(), res["nms_num"]]
    if cfg.with_crop:
        outs.append(res["crop"][5])
    return outs


class _NoProbe:
    """Default probe: does nothing.  bench.py passes one that brackets each op with CUDA events."""

    def __call__(self, name, **meta):
        return self

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class _Lanes:
    """The CUDA streams one chain pass is spread over (cached per device).

    The SA sampling chain fps_1 -> gather_1 -> fps_2 -> ... -> gather_4 is strictly sequential
    and latency bound (one cluster per scan); everything else only depends on one of its
    intermediate results or on nothing at all.  Issuing those on side streams lets their CTAs
    fill the SM resources the sampling CTAs leave idle:
        crit      fps / gather_point of every level                     (highest priority, issued first)
        group[l]  [cloud index of level l+1] query_ball_point + group_point x2 of level l (after gather_l)
        nn[l]     three_nn (+ interpolation weights) of level l as soon as its two clouds exist
        fp        the three_interpolate calls coarse -> fine (the order the real network needs)
        crop      pc_crop_and_sample                                      (independent)
        nms       oriented_nms                                            (independent)
    One stream per level: level l+1's ball query must not queue behind level l's group_point.
    Priorities (measured, profiles/r1_chain_schedule.md): sampling > grouping / three_nn /
    interpolation > NMS > crop — the crop's 8192 CTAs would otherwise starve the small kernels
    behind the sampling chain; HF_CHAIN_PRIORITIES="crit,group,nn,crop,nms" overrides.
    """

    _cache = {}

    def __init__(self, device):
        import os
        pr = [int(v) for v in os.environ.get("HF_CHAIN_PRIORITIES", "-3,-2,-2,0,-1").split(",")]
        mk = lambda p: torch.cuda.Stream(device=device, priority=p)  # noqa: E731
        self.crit = mk(pr[0])
        self.group = [mk(pr[1]) for _ in range(4)]
        self.nn = [mk(pr[2]) for _ in range(4)]
        self.fp = mk(pr[2])
        self.crop = mk(pr[3])
        self.nms = mk(pr[4])

    @classmethod
    def get(cls, device, slot=0):
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device(), slot)
        if key not in cls._cache:
            cls._cache[key] = cls(device)
        return cls._cache[key]

    def all(self):
        return [self.crit] + self.group + self.nn + [self.fp, self.crop, self.nms]


def run_chain(dev, cfg, probe=None, overlap=True, slot=0, join=True):
    """Run the chain on device-resident inputs `dev` (see make_host_inputs).  Returns a dict of results.

    overlap=True spreads the independent branches of the chain over CUDA streams (see _Lanes);
    the caller's current stream waits for all of them before this function returns, so the
    results are ordered on the current stream as usual.  overlap=False issues every op on the
    current stream in the reference's program order.

    Batches are independent, so a caller can keep several in flight: `slot` selects a separate
    set of streams and join=False returns without making the current stream wait — the events
    of the pass are in res["_done"] instead (see ChainPipeline).  The sampling chain of one
    batch is latency bound and leaves most of the GPU idle; the next batch's fills it.

    `probe(name, **shape)` (optional) must return a context manager; every libhfops launch runs
    inside one, named after its kernel family, with the shape figures its algorithmic
    bytes / flops are computed from.
    """
    probe = probe or _NoProbe()
    device = dev["xyz"].device
    main = torch.cuda.current_stream(device)
    if overlap:
        lanes = _Lanes.get(device, slot)
        crit, s_group, s_nn, s_fp, s_crop, s_nms = lanes.crit, lanes.group, lanes.nn, lanes.fp, lanes.crop, lanes.nms
        start = torch.cuda.Event()
        start.record(main)
        for s in lanes.all():
            s.wait_event(start)
    else:
        crit = s_fp = s_crop = s_nms = main
        s_group = s_nn = [main] * 4
    on = torch.cuda.stream

    def after(stream, event):
        if overlap and event is not None:
            stream.wait_event(event)

    def mark(stream):
        if not overlap:
            return None
        e = torch.cuda.Event()
        e.record(stream)
        return e

    res = {}
    xyz = [dev["xyz"]]
    fps_idx = []
    feats = [dev["intensity"]] + list(dev["sa_feats"][1:])
    b = dev["xyz"].shape[0]
    n_lv = [cfg.points] + list(cfg.npoint)
    ns = cfg.nsample

    def sample_level(l):
        """Sampling of SA level l (0-based) on the critical lane; returns the event after which xyz[l + 1] exists.
        With fused_sampling, level 0 is one launch (indices + picked points) and the call for level 1 samples ALL
        remaining levels in one launch (their events are the same one)."""
        n, m = n_lv[l], n_lv[l + 1]
        if len(xyz) > l + 1:  # already produced by the pyramid launch of level 1
            return have_xyz[l - 1]
        with on(crit):
            if not cfg.fused_sampling:
                with probe("fps", level=l + 1, b=b, n=n, m=m):
                    idx = farthest_point_sample(m, xyz[l])
                with probe("gather_point", level=l + 1, b=b, n=n, m=m):
                    new_xyz = gather_point(xyz[l], idx)
                fps_idx.append(idx)
                xyz.append(new_xyz)
            elif l < cfg.pyramid_from or len(n_lv) - 1 - l > PYRAMID_MAX_LEVELS:
                ix = index[0] if (l == 0 and sorted_fps) else None
                if ix is not None:
                    after(crit, index_ready[0])
                force = cfg.fps_level1_warps if l == 0 else 0
                if force:  # scoped: the variant the caller (a probe under tools/) had selected comes back afterwards
                    import ctypes

                    from . import _lib
                    prev = [ctypes.c_int(0) for _ in range(3)]
                    _lib.load().hf_fps_get_variant(*[ctypes.byref(v) for v in prev])
                    _lib.load().hf_fps_set_variant(3, force, 0)
                try:
                    with probe("fps", level=l + 1, b=b, n=n, m=m, xyz=1):
                        idx, new_xyz = farthest_point_sample_with_xyz(m, xyz[l], index=ix)
                finally:
                    if force:
                        _lib.load().hf_fps_set_variant(*[v.value for v in prev])
                fps_idx.append(idx)
                xyz.append(new_xyz)
            else:
                ms = n_lv[l + 1:]
                with probe("fps_pyramid", level=l + 1, b=b, n=n, ms=tuple(ms)):
                    idxs, pts = farthest_point_sample_pyramid(ms, xyz[l])
                fps_idx.extend(idxs)
                xyz.extend(pts)
        return mark(crit)

    # The sampling chain is the spine of the step: its first (longest) launch goes out before anything else,
    # so its clusters take their SMs while the GPU is still empty.  In program order (overlap=False) the
    # independent branches keep the reference's order instead.
    have_xyz = [None] * 4
    # spatial index of each level's cloud, built once and shared by ball query and three_nn of the
    # two levels that search it (only where the calls are large enough to take the pruned kernels).
    # The input's index needs whole SMs (1024-thread CTAs): it goes out before the crop floods the GPU.
    use_ix = [spatial.applies(n_lv[l + 1], n_lv[l]) and spatial.applies(n_lv[l], n_lv[l + 1]) for l in range(4)]
    index, index_ready = [None] * 5, [None] * 5
    sorted_fps = use_ix[0] and (cfg.sorted_fps if cfg.sorted_fps is not None else True)
    if overlap and not sorted_fps:
        have_xyz[0] = sample_level(0)
    if use_ix[0]:
        with on(s_group[0]), probe("cloud_index", level=0, b=b, n=n_lv[0]):
            index[0] = spatial.build_cloud_index(xyz[0])  # needs only the input: overlaps FPS level 1
        index_ready[0] = mark(s_group[0])
    if overlap and sorted_fps:  # the level-1 FPS reads the cloud through its index: right behind it
        have_xyz[0] = sample_level(0)
    if cfg.with_nms:
        with on(s_nms), probe("oriented_nms", b=b, nb=cfg.nms_boxes):
            # sb_nms_fn keeps the first post_nms entries (model_util.py:126): the sweep stops there
            keep, num = oriented_nms_batch(dev["nms_boxes"], cfg.nms_thresh, return_num_kept=True,
                                           max_keep=cfg.post_nms if cfg.nms_early_stop else None)
        res["nms_keep"] = keep[:, : cfg.post_nms]
        res["nms_num"] = num
    if cfg.with_crop:
        after(s_crop, index_ready[0])  # the input's index (if built) also prunes the crop's point-in-box scan
        with on(s_crop), probe("pc_crop_and_sample", b=b, p=cfg.points, c=cfg.crop_channels,
                               nb=dev["crop_boxes"].shape[0], r=cfg.crop_resize):
            res["crop"] = pc_crop_and_sample(dev["xyz"], dev["crop_fts"], dev["intensity"], dev["fg_mask"],
                                             dev["crop_boxes"], dev["box_ind"], cfg.crop_resize, index=index[0])
    # ---- set abstraction ----
    nn = [None] * 4
    nn_ready = [None] * 4
    for l in range(4):
        n, m = n_lv[l], n_lv[l + 1]
        if have_xyz[l] is None:
            have_xyz[l] = sample_level(l)
        new_xyz = xyz[l + 1]
        sg = s_group[l]
        after(sg, have_xyz[l])
        with on(sg):
            if use_ix[l] or (l + 1 < 4 and use_ix[l + 1]):
                with probe("cloud_index", level=l + 1, b=b, n=m):
                    index[l + 1] = spatial.build_cloud_index(new_xyz)
                index_ready[l + 1] = mark(sg)
            ix = (index[l], index[l + 1]) if use_ix[l] else (None, None)
            if use_ix[l]:
                after(sg, index_ready[l])
            with probe("query_ball_point", level=l + 1, b=b, n=n, m=m, ns=ns):
                gidx, cnt = query_ball_point(cfg.radius[l], ns, xyz[l], new_xyz, index1=ix[0], index2=ix[1])
            if cfg.fused_grouping:
                with probe("group_point_pair", level=l + 1, b=b, n=n, m=m, ns=ns, c=feats[l].shape[2]):
                    gxyz, gfeat = group_point_pair(xyz[l], feats[l], gidx)
            else:
                with probe("group_point", level=l + 1, b=b, n=n, m=m, ns=ns, c=3):
                    gxyz = group_point(xyz[l], gidx)
                with probe("group_point", level=l + 1, b=b, n=n, m=m, ns=ns, c=feats[l].shape[2]):
                    gfeat = group_point(feats[l], gidx)
        # three_nn of FP level l+1 (unknown = level l cloud, known = level l+1 cloud) is geometry only;
        # the interpolation weights (pointnet_util.py:304-307) come out of the same kernel
        sn = s_nn[l]
        after(sn, have_xyz[l])
        if use_ix[l]:
            after(sn, index_ready[l])
            after(sn, index_ready[l + 1])
        with on(sn):
            with probe("three_nn", level=l + 1, b=b, n=n, m=m):
                nn[l] = three_nn_weights(xyz[l], new_xyz, index1=ix[0], index2=ix[1])
        nn_ready[l] = mark(sn)
        res["sa%d" % (l + 1)] = (gidx, cnt, gxyz, gfeat)
    res["fps_idx"] = fps_idx
    res["xyz_levels"] = xyz
    res["_keep"] = index  # referenced until the pass is over: other streams read them
    # ---- feature propagation, coarse to fine (pointnet_util.py:289-330) ----
    for j in range(4):
        lvl = 3 - j  # unknown cloud = xyz[lvl], known = xyz[lvl + 1]
        dist, nidx, w = nn[lvl]
        n, m, c = n_lv[lvl], n_lv[lvl + 1], dev["fp_feats"][j].shape[2]
        after(s_fp, nn_ready[lvl])
        with on(s_fp), probe("three_interpolate", level=lvl + 1, b=b, n=n, m=m, c=c):
            out = three_interpolate(dev["fp_feats"][j], nidx, w)
        res["fp%d" % (lvl + 1)] = (dist, nidx, out)
        res.setdefault("fp_weights", [None] * 4)[lvl] = w
    if overlap:
        done = [mark(s) for s in lanes.all()]
        if join:
            for e in done:
                main.wait_event(e)
        res["_done"] = done
    else:
        res["_done"] = []
    return res


class ChainPipeline:
    """Keeps up to `depth` batches in flight on separate stream sets.

    submit() issues one pass and returns its results dict; before a slot is reused the host
    waits for the pass that used it last and only then drops its tensors, so memory handed back
    to the allocator is never still being read on another stream.  drain() makes the current
    stream wait for everything in flight.
    """

    def __init__(self, cfg, depth=2, overlap=True):
        self.cfg, self.depth, self.overlap = cfg, max(1, depth), overlap
        self.inflight = [None] * self.depth
        self.k = 0

    def next_slot(self):
        return self.k % self.depth

    def release(self, slot):
        """Host-wait for the pass that used `slot` last and drop its tensors.  Call it before overwriting input
        buffers that pass was reading (submit() calls it for its own slot)."""
        old = self.inflight[slot]
        if old is not None:
            for e in old[1]["_done"]:
                e.synchronize()
            self.inflight[slot] = None

    def submit(self, dev, probe=None):
        slot = self.k % self.depth
        self.k += 1
        self.release(slot)
        res = run_chain(dev, self.cfg, probe, overlap=self.overlap, slot=slot, join=not self.overlap or self.depth == 1)
        self.inflight[slot] = (dev, res)  # the inputs stay referenced as long as the pass may read them
        return res

    def drain(self):
        main = torch.cuda.current_stream()
        for item in self.inflight:
            if item is not None:
                for e in item[1]["_done"]:
                    main.wait_event(e)


class ChainGraph:
    """One chain pass over fixed device buffers, captured into a CUDA graph and replayed.

    The step is ~40 launches of 10-100 us kernels on 13 streams; issued from Python the host needs
    about as long to enqueue them as the GPU needs to run them, and every dependent launch adds
    its latency to the tail behind the level-1 FPS.  Captured once (all lanes fork from and join
    the capture stream), a step is one cudaGraphLaunch.  Inputs are read from `dev` and results
    are written to the same tensors on every replay: copy new inputs into `dev` (stream ordered)
    before replay() and consume `res` before the next one.
    """

    def __init__(self, dev, cfg, slot=0):
        from . import _lib

        self.dev, self.cfg = dev, cfg
        device = dev["xyz"].device
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):  # eager passes first: one-time attribute calls, allocator pools
            for _ in range(2):
                run_chain(dev, cfg, slot=slot)
        torch.cuda.current_stream(device).wait_stream(side)
        torch.cuda.synchronize(device)
        self.graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.res = run_chain(dev, cfg, slot=slot)
        self.launches = _lib.launch_count() - before  # libhfops kernels per replay

    def replay(self):
        self.graph.replay()
        return self.res


def chain_launches(cfg):
    """Kernels of libhfops launched by one run_chain call."""
    n_lv = [cfg.points] + list(cfg.npoint)
    use_ix = [spatial.applies(n_lv[l + 1], n_lv[l]) and spatial.applies(n_lv[l], n_lv[l + 1]) for l in range(4)]
    indexed = [use_ix[l] or (l > 0 and use_ix[l - 1]) for l in range(4)] + [use_ix[3]]
    # fps+xyz per level below pyramid_from and one pyramid launch for the rest / fps + gather per level
    sampling = (min(max(cfg.pyramid_from, 1), 4) + (1 if cfg.pyramid_from < 4 else 0)) if cfg.fused_sampling else 8
    grouping = 4 * (2 if cfg.fused_grouping else 3)  # ball query + group_point (pair) per level
    second_level = sum(1 for l in range(5) if indexed[l] and n_lv[l] > 16384)  # cloud_super_kernel of a large cloud
    return sampling + grouping + 4 * 2 + sum(indexed) + second_level + (1 if cfg.with_crop else 0) + (5 if cfg.with_nms else 0)
