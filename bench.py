#!/usr/bin/env python
"""bench.py — scans/sec of the point-cloud operator chain (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

One "step" = one pass of the whole op chain (SA levels 1-4: FPS -> gather_point ->
query_ball_point -> group_point x2; FP levels 4-1: three_nn -> weights -> three_interpolate;
RoI point crop of 512 proposals/scan; oriented NMS of 9000 BEV boxes/scan) over one batch
of 16 synthetic KITTI-shaped 16384-point scans per GPU (BASELINE.json configs[1..3]).
Scenes are independent, so with N GPUs every rank runs its own batch (weak scaling, no
data-path collective); torch.distributed is used only for the barrier and the max-over-ranks
of the device-timed duration.

Prints ONE JSON line (rank 0).  `value` = scans/s with inputs resident in HBM; `e2e` = the
same chain through the public Python op API with every input copied from pinned host
memory and the host-consumed results copied back, inside the timed region; `roofline` =
the dominant libhfops launch of the step against its roofline, `roofline_ops` = every
launch; `cpu_baseline` = the CPU oracle (the reference's CPU path / kernel restatement,
oracle/) timed on this box's host cores on a bounded sample.

--impl reference times that CPU path itself (all host threads) and prints the same line.
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # before the CUDA context exists (see heterofusionrcnn_b200/__init__.py)

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "scans_per_sec_16k_sa_fp_crop_nms_chain"
UNIT = "scans/s"
FP32_PEAK_DERIVED = 148 * 128 * 2 * 1.965e9 / 1e12  # SMs x lanes x FMA x max clock; the line reports the MEASURED figure


def measure_fp32_peak(torch, hf, dev):
    """FP32 non-tensor peak of this GPU: hf_fp32_probe (csrc/api.cu) = nothing but independent FFMA chains, 2048
    resident threads per SM, timed with CUDA events, best of 5.  Scalar FFMA and packed FFMA2 (fma.rn.f32x2) are
    both measured; the roofline denominator is the larger."""
    from heterofusionrcnn_b200 import _lib

    lib = _lib.load()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    blocks, iters = sms * 2 * 4, 8192
    out = torch.empty((int(lib.hf_fp32_probe_out_floats(blocks)),), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    res = {}
    for packed in (0, 1):
        best = None
        for it in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(lib.hf_fp32_probe(blocks, iters, packed, out.data_ptr(), st), "fp32 probe")
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            if it > 0 and (best is None or ms < best):
                best = ms
        res["ffma2" if packed else "ffma"] = float(lib.hf_fp32_probe_flops(blocks, iters, packed)) / (best / 1e3) / 1e12
    res["peak"] = max(res["ffma"], res["ffma2"])
    res["derived"] = FP32_PEAK_DERIVED
    return res


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16, help="scans per GPU per step")
    ap.add_argument("--points", type=int, default=16384)
    ap.add_argument("--msg-radii", action="store_true",
                    help="use the reference configs' outer MSG radii [0.5,1,2,4] instead of BASELINE's [0.1..0.8]")
    ap.add_argument("--cpu-scenes", type=int, default=0, help="CPU sample size in scans (0 = auto)")
    ap.add_argument("--no-side-legs", action="store_true",
                    help="skip the separately timed legs of rank 0 at N=1 (BEV IoU / NMS pair rates, BASELINE configs[0])")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--table", action="store_true", help="also print a per-launch table to stderr")
    ap.add_argument("--inflight", type=int, default=2,
                    help="e2e leg: batches kept in flight, so that a batch's host->device copies overlap the previous batch's kernels")
    ap.add_argument("--resident-inflight", type=int, default=1,
                    help="resident leg: batches kept in flight (measured: 2 is slower than 1 - the second batch's FPS "
                         "clusters cannot be gang-scheduled while the first batch's bulk kernels fill every SM)")
    ap.add_argument("--no-graph", action="store_true",
                    help="resident leg: issue every launch from Python each step instead of replaying the captured CUDA graph "
                         "(equal within noise on one GPU; with 8 ranks on one host the Python enqueue becomes the bottleneck: "
                         "2.45 -> 3.63 ms/step eager, profiles/r1_bench_8gpu_eager_v10.json)")
    ap.add_argument("--timeline", action="store_true",
                    help="print (stderr) when every op of the last timed step started and ended, relative to the step's start")
    ap.add_argument("--profile-range", action="store_true",
                    help="bracket the timed region with cudaProfilerStart/Stop (ncu --profile-from-start off)")
    ap.add_argument("--no-config5", action="store_true",
                    help="skip the BASELINE configs[4] leg (128 scans x 65536 points in total, strong scaling over the ranks)")
    ap.add_argument("--config5-steps", type=int, default=3, help="timed passes over the 128 scans of the configs[4] leg")
    ap.add_argument("--config5-scans", type=int, default=128, help="total scans of the configs[4] leg (multiple of 16 x ranks)")
    ap.add_argument("--serial", action="store_true",
                    help="issue the whole chain on one stream in program order (default: independent branches overlap on streams)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_cfg(args):
    from heterofusionrcnn_b200 import chain

    cfg = chain.ChainConfig()
    if args.points != cfg.points:
        cfg = cfg.scaled(args.points)
    if args.msg_radii:
        cfg.radius = [0.5, 1.0, 2.0, 4.0]
    return cfg


def config_dict(args, cfg, n_gpus):
    return {
        "workload": "configs[1]+[2]+[3]: RPN PointNet++ SA/FP op chain (FPS %d->%s, ball query r=%s ns=%d, group_point, "
                    "three_nn+three_interpolate c=%s) + RoI crop %d boxes x %d pts C=%d + BEV NMS %d->%d per scan; "
                    "batch %d scans/GPU" % (cfg.points, "->".join(map(str, cfg.npoint)), cfg.radius, cfg.nsample,
                                            cfg.fp_channels, cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels,
                                            cfg.nms_boxes, cfg.post_nms, args.batch),
        "scans_per_gpu": args.batch,
        "global_scans": args.batch * n_gpus,
        "points_per_scan": cfg.points,
        "sharding": "batch slices, one per GPU, no collective" if n_gpus > 1 else "single GPU",
        "issue": "one stream, program order" if args.serial else
                 "independent branches of the chain on CUDA streams%s, %d batch(es) in flight (e2e leg: eager, %d)" % (
                     "" if (args.no_graph or args.timeline) else ", one step = one replay of the captured CUDA graph",
                     max(1, args.resident_inflight), max(1, args.inflight)),
        "l2": "no explicit flush: one step reads 0.45 GB and writes 5.6 GB per GPU, far beyond the 126 MB L2",
    }


# --------------------------------------------------------------------------------------
# algorithmic work per launch
# --------------------------------------------------------------------------------------
def work(name, m):
    """-> (bound, algorithmic bytes, algorithmic flops) of one launch."""
    if name == "fps":
        return "fp32", 12 * m["b"] * m["n"] + 4 * m["b"] * m["m"], 8 * m["b"] * (m["m"] - 1) * m["n"]
    if name == "fps_pyramid":  # levels n -> ms[0] -> ms[1] ...: the FPS formula per level + the gathered points
        ns = [m["n"]] + list(m["ms"])
        by = sum(12 * m["b"] * ns[j] + 4 * m["b"] * ns[j + 1] + 28 * m["b"] * ns[j + 1] for j in range(len(ns) - 1))
        fl = sum(8 * m["b"] * (ns[j + 1] - 1) * ns[j] for j in range(len(ns) - 1))
        return "fp32", by, fl
    if name == "gather_point":
        return "hbm", 4 * m["b"] * m["m"] * 7, 0
    if name == "query_ball_point":
        return "fp32", 12 * m["b"] * (m["n"] + m["m"]) + 4 * m["b"] * m["m"] * (m["ns"] + 1), 9 * m["b"] * m["m"] * m["n"]
    if name == "group_point":
        return "hbm", 4 * (m["b"] * m["m"] * m["ns"] * m["c"] + m["b"] * m["m"] * m["ns"] + m["b"] * m["n"] * m["c"]), 0
    if name == "group_point_pair":  # xyz (3) and features (c) of the same rows: the group_point formula for 3 + c
        cc = m["c"] + 3
        return "hbm", 4 * (m["b"] * m["m"] * m["ns"] * cc + m["b"] * m["m"] * m["ns"] + m["b"] * m["n"] * cc), 0
    if name == "three_nn":
        return "fp32", 12 * m["b"] * (m["n"] + m["m"]) + 24 * m["b"] * m["n"], 8 * m["b"] * m["n"] * m["m"]
    if name == "three_interpolate":
        return "hbm", 4 * (m["b"] * m["n"] * m["c"] + m["b"] * m["m"] * m["c"] + 6 * m["b"] * m["n"]), 0
    if name == "pc_crop_and_sample":
        nb, r, c, b, p = m["nb"], m["r"], m["c"], m["b"], m["p"]
        return "hbm", 4 * nb * r * (3 + c + 1 + 1) + nb * r + nb + 4 * b * p * (3 + c + 1) + b * p + 96 * nb, 0
    if name == "oriented_nms":
        nb = m["nb"]
        return "fp32", m["b"] * (20 * nb + 4 * nb), 0  # pairs, not flops, are the work unit: reported separately
    return "other", 0, 0


def count_pair_evals(torch, hf, chain, resident, cfg):
    """-> {launch key: (query, candidate) distance evaluations executed by that launch}; 0 entries for launches that
    took an every-pair kernel (they execute the full count)."""
    counter = torch.zeros((1,), dtype=torch.int64, device=resident["xyz"].device)
    lib = hf.load()
    found = {}

    @contextlib.contextmanager
    def probe(name, **meta):
        if name not in ("three_nn", "query_ball_point"):
            yield
            return
        torch.cuda.synchronize()
        before = int(counter.item())
        yield
        torch.cuda.synchronize()
        key = name + "".join("/%s=%s" % (k, meta[k]) for k in ("level", "c") if k in meta)
        found[key] = int(counter.item()) - before

    lib.hf_spatial_set_pair_counter(counter.data_ptr())
    try:
        chain.run_chain(resident, cfg, probe, overlap=False)
        torch.cuda.synchronize()
    finally:
        lib.hf_spatial_set_pair_counter(None)
    return found


def back_to_back(torch, chain, resident, cfg, reps=20):
    """The HBM-bound gathers of the chain re-issued `reps` times back to back between ONE pair of CUDA events (L2 flushed
    first): the per-launch event pair of the `alone` leg costs ~5 us on launches of 15-60 us.  Same tensors, same calls
    as chain.run_chain.  -> {launch key: ms per launch}"""
    from heterofusionrcnn_b200.grouping.tf_grouping import group_point_pair
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_interpolate

    res = chain.run_chain(resident, cfg, None, overlap=False)
    torch.cuda.synchronize()
    dev = resident["xyz"].device
    flush = torch.empty(512 * 1024 * 1024 // 4, device=dev)
    feats = [resident["intensity"]] + list(resident["sa_feats"][1:])
    calls = {}
    for l in range(4):
        gidx = res["sa%d" % (l + 1)][0]
        if cfg.fused_grouping:
            calls["group_point_pair/level=%d/c=%d" % (l + 1, feats[l].shape[2])] = (
                lambda l=l, gidx=gidx: group_point_pair(res["xyz_levels"][l], feats[l], gidx))
    for j in range(4):
        lvl = 3 - j
        pts, nidx, w = resident["fp_feats"][j], res["fp%d" % (lvl + 1)][1], res["fp_weights"][lvl]
        calls["three_interpolate/level=%d/c=%d" % (lvl + 1, pts.shape[2])] = (
            lambda pts=pts, nidx=nidx, w=w: three_interpolate(pts, nidx, w))
    out = {}
    for key, fn in calls.items():
        for _ in range(3):
            fn()
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[key] = e0.elapsed_time(e1) / reps
    del flush
    return out


class EventProbe:
    """Brackets every op of chain.run_chain with CUDA events on the launching (current) stream."""

    def __init__(self, torch):
        self.torch = torch
        self.records = []  # (key, name, meta, start, end)
        self.enabled = True

    @contextlib.contextmanager
    def __call__(self, name, **meta):
        if not self.enabled:
            yield
            return
        s = self.torch.cuda.Event(enable_timing=True)
        e = self.torch.cuda.Event(enable_timing=True)
        s.record()
        yield
        e.record()
        key = name + "".join("/%s=%s" % (k, meta[k]) for k in ("level", "c") if k in meta)
        self.records.append((key, name, meta, s, e))

    def summarise(self, steps):
        agg = {}
        for key, name, meta, s, e in self.records:
            a = agg.setdefault(key, {"name": name, "meta": meta, "ms": 0.0, "n": 0})
            a["ms"] += s.elapsed_time(e)
            a["n"] += 1
        out = []
        for key, a in agg.items():
            bound, nbytes, flops = work(a["name"], a["meta"])
            avg = a["ms"] / a["n"]
            out.append({"launch": key, "bound": bound, "ms": round(avg, 5), "per_step": a["n"] // max(steps, 1),
                        "bytes": nbytes, "flops": flops, "meta": a["meta"]})
        return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------
# CPU path (oracle) — the timed baseline and the --impl reference arm
# --------------------------------------------------------------------------------------
def cpu_chain(O, host, cfg, scenes):
    """The same chain as chain.run_chain on the CPU oracle, first `scenes` scans of `host`."""
    from heterofusionrcnn_b200 import synthetic

    s = slice(0, scenes)
    xyz = [host["xyz"][s]]
    feats = [host["intensity"][s]] + [f[s] for f in host["sa_feats"][1:]]
    for l in range(4):
        idx = O.farthest_point_sample(cfg.npoint[l], xyz[l])
        new_xyz = O.gather_point(xyz[l], idx)
        gidx, _ = O.query_ball_point(cfg.radius[l], cfg.nsample, xyz[l], new_xyz)
        O.group_point(xyz[l], gidx)
        O.group_point(feats[l], gidx)
        xyz.append(new_xyz)
    for j in range(4):
        d, i = O.three_nn(xyz[3 - j], xyz[4 - j])
        O.three_interpolate(host["fp_feats"][j][s], i, synthetic.fp_weights(d))
    nb = cfg.crop_boxes * scenes
    O.pc_crop_and_sample(host["xyz"][s], host["crop_fts"][s], host["intensity"][s], host["fg_mask"][s],
                         host["crop_boxes"][:nb], host["box_ind"][:nb], cfg.crop_resize)
    keeps = [O.oriented_nms(host["nms_boxes"][k], cfg.nms_thresh)[: cfg.post_nms] for k in range(scenes)]
    return keeps


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        return os.cpu_count() or 1


def cpu_sample_size(args, cores):
    """Scans per CPU pass: one per host core, so that the scene-parallel ops (FPS is one thread per scan) keep
    every core busy like the row-parallel ones do (scene-parallel over ALL host cores)."""
    if args.cpu_scenes > 0:
        return args.cpu_scenes
    return max(1, min(cores, 64))


def time_cpu(args, cfg, steps, warmup, single_thread):
    """The CPU path on this box's host cores.  Two figures (SURVEY.md 8d / BASELINE.md section 2):
    all cores = `cores` std::threads over scenes / rows, one scan per core per pass (no algorithmic change);
    single thread = the code exactly as the reference is written, one scan."""
    from heterofusionrcnn_b200 import chain
    from oracle import binding as O

    O.build()
    cores = host_cores()
    scenes = cpu_sample_size(args, cores)
    host = chain.make_host_inputs(cfg, batch=scenes, seed=2)
    O.set_threads(cores)
    for _ in range(warmup):
        cpu_chain(O, host, cfg, scenes)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_chain(O, host, cfg, scenes)
    dt = (time.perf_counter() - t0) / steps
    cb = {"value": scenes / dt, "unit": UNIT, "cores": cores, "kind": "port",
          "sample": "%d scans (one per host core) of the bench workload, whole chain, oracle/hf_oracle.cpp (g++ -O2), %d std::threads "
                    "over scenes / rows; %.2f s per pass" % (scenes, cores, dt)}
    if single_thread:
        O.set_threads(1)
        t0 = time.perf_counter()
        cpu_chain(O, host, cfg, 1)
        dt1 = time.perf_counter() - t0
        O.set_threads(cores)
        cb["single_thread"] = {"value": 1.0 / dt1, "unit": UNIT, "cores": 1,
                               "sample": "1 scan, whole chain, one thread, the code as the reference is written; %.2f s" % dt1}
    return cb, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cfg = make_cfg(args)
    cb, dt = time_cpu(args, cfg, args.steps, args.warmup, single_thread=False)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, cfg, args.gpus), "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------
# legs timed on their own (SURVEY.md 8d work units that the chained step does not expose)
# --------------------------------------------------------------------------------------
def side_legs(torch, hf, dev, cfg, batch):
    """Rank 0, N=1, after the timed region.  Each leg: 3 warm-up calls, then `reps` calls between two CUDA events on
    the current stream, GPU otherwise idle, inputs resident.
      bev_iou   compute_bev_iou at SURVEY 8(a) a8's sizes 9000x64 and 2048x2048: pairs/s (N*M pairs, 8d) and the HBM
                fraction of its algorithmic bytes (20 B per input box, 8 B per pair written);
      nms       oriented_nms, `batch` scenes x cfg.nms_boxes, full keep list, at the thresholds of rpn_multiclass.config:
                25-30,43 (0.85 / 0.8 / 0.01): pairs/s over N(N-1)/2 pairs per scene (8d) and how many of them reached the
                exact polygon clip; plus the early-stopping form the chained step uses (first cfg.post_nms kept);
      config0   BASELINE configs[0]: three_nn + three_interpolate of fixture scan 000000 (16384 <- 4096 FPS picks,
                C = 128), GPU launches beside the reference's own CPU code for the pair: interpolate/interpolate.cpp
                compiled as it lies (oracle/_ref/libref_interpolate_cpu.so, built in the authoring container and shipped;
                the oracle port when it is absent - said in `cpu_kind`) after the oracle's three_nn (the reference's
                threenn_cpu ignores the query point, SURVEY 8c, so its three_nn is not timed)."""
    import numpy as np

    from heterofusionrcnn_b200 import synthetic
    from heterofusionrcnn_b200.bev_iou.bev_iou import compute_bev_iou, oriented_nms_batch
    from heterofusionrcnn_b200.interpolate.tf_interpolate import three_interpolate, three_nn
    from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample, gather_point

    hbm_peak, _ = peaks()
    lib = hf.load()

    def timed(fn, reps):
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

    out = {"note": "each leg timed alone with CUDA events (3 warm-up calls, then the stated repetitions), inputs resident"}
    # ---- BEV IoU matrix
    iou = []
    for na, nb in ((9000, 64), (2048, 2048)):
        a = torch.from_numpy(synthetic.make_nms_boxes(5, na)).to(dev)
        b = torch.from_numpy(synthetic.make_nms_boxes(6, nb)).to(dev)
        ms = timed(lambda: compute_bev_iou(a, b), 50)
        by = 20 * (na + nb) + 8 * na * nb
        iou.append({"num_a": na, "num_b": nb, "reps": 50, "ms": ms, "pairs_per_s": na * nb / (ms / 1e3),
                    "algorithmic_bytes": by, "hbm_frac": by / (ms / 1e3) / 1e9 / hbm_peak,
                    "note": "through the Python op, which also allocates both outputs per call"})
    out["bev_iou"] = iou
    # ---- oriented NMS
    nbx = cfg.nms_boxes
    boxes = torch.from_numpy(np.stack([synthetic.make_nms_boxes(2 * 17 + i, nbx) for i in range(batch)])).to(dev)
    ws_bytes = int(lib.hf_oriented_nms_workspace_bytes(batch, nbx))
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    keep = torch.empty((batch, nbx), dtype=torch.int32, device=dev)
    num = torch.empty((batch,), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    all_pairs = batch * nbx * (nbx - 1) // 2
    nms = []
    for th in (0.85, 0.8, 0.01):
        def full():
            rc = lib.hf_oriented_nms(boxes.data_ptr(), batch, nbx, th, keep.data_ptr(), num.data_ptr(), ws.data_ptr(), ws_bytes, st)
            assert rc == 0, lib.hf_last_error()
        ms = timed(full, 10)
        clipped = int(lib.hf_oriented_nms_last_pair_count(batch, nbx, ws.data_ptr(), st))
        ms_top = timed(lambda: oriented_nms_batch(boxes, th, return_num_kept=True, max_keep=cfg.post_nms), 10)
        nms.append({"thresh": th, "scenes": batch, "boxes": nbx, "reps": 10, "ms_full_keep_list": ms, "pairs": all_pairs,
                    "pairs_per_s": all_pairs / (ms / 1e3), "pairs_clipped": clipped, "clipped_share": clipped / all_pairs,
                    "kept_mean": float(num.float().mean().item()), "ms_first_%d_kept" % cfg.post_nms: ms_top})
    out["oriented_nms"] = nms
    # ---- BASELINE configs[0]
    gpath = os.path.join(ROOT, "tests", "golden", "kitti_000000_16384.npz")
    if os.path.exists(gpath):
        from oracle import binding as O  # the CPU side of this leg (the cpu_baseline use of the checker)

        O.build()
        xyz = np.load(gpath)["xyz"][None].astype(np.float32)
        x = torch.from_numpy(xyz).to(dev)
        known = gather_point(x, farthest_point_sample(4096, x))
        feats_h = np.random.default_rng(0).standard_normal((1, 4096, 128)).astype(np.float32)
        feats = torch.from_numpy(feats_h).to(dev)
        d, i = three_nn(x, known)
        w_h = synthetic.fp_weights(d.cpu().numpy())
        w = torch.from_numpy(w_h).to(dev)
        ms_nn = timed(lambda: three_nn(x, known), 50)
        ms_ip = timed(lambda: three_interpolate(feats, i, w), 50)
        known_h, i_h = known.cpu().numpy(), i.cpu().numpy()
        O.set_threads(1)
        t0 = time.perf_counter()
        O.three_nn(xyz, known_h)
        cpu_nn = (time.perf_counter() - t0) * 1e3
        use_ref = O.ref_available("libref_interpolate_cpu.so")
        ip = O.ref_interpolate_cpu if use_ref else O.three_interpolate
        ip(feats_h, i_h, w_h)
        t0 = time.perf_counter()
        for _ in range(5):
            ip(feats_h, i_h, w_h)
        cpu_ip = (time.perf_counter() - t0) * 1e3 / 5
        O.set_threads(host_cores())
        out["config0"] = {
            "workload": "configs[0]: fixture KITTI scan 000000, three_nn + three_interpolate, 16384 <- 4096, C=128, one scan",
            "gpu_ms": {"three_nn": ms_nn, "three_interpolate": ms_ip, "both": ms_nn + ms_ip},
            "cpu_ms": {"three_nn": cpu_nn, "three_interpolate": cpu_ip, "both": cpu_nn + cpu_ip}, "cpu_cores": 1,
            "cpu_kind": {"three_nn": "port (oracle/hf_oracle.cpp restating tf_interpolate_g.cu:22-65)",
                         "three_interpolate": "reference (interpolate/interpolate.cpp:84-104 compiled unmodified)" if use_ref
                         else "port (oracle/_ref/libref_interpolate_cpu.so not shipped)"},
            "scans_per_s_gpu": 1e3 / (ms_nn + ms_ip), "scans_per_s_cpu": 1e3 / (cpu_nn + cpu_ip)}
    return out


# --------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------
def nbytes_of(obj):
    import torch

    if obj is None:
        return 0
    if isinstance(obj, torch.Tensor):
        return obj.numel() * obj.element_size()
    if isinstance(obj, (list, tuple)):
        return sum(nbytes_of(x) for x in obj)
    if isinstance(obj, dict):
        return sum(nbytes_of(x) for x in obj.values())
    return 0


def parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_rank_to_local_cores(local, world):
    """Before any CUDA allocation: restrict this rank to its share of the cores that are local to its GPU
    (/sys/bus/pci/devices/<bdf>/local_cpulist), so that its pinned staging memory is allocated on, and its copies
    are driven from, the GPU's own NUMA node.  GPUs with the same local core list split it evenly.  Returns a
    description for the JSON line (None if nothing could be done)."""
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=index,pci.bus_id", "--format=csv,noheader"], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, text=True, timeout=20).stdout
        lists, nodes = {}, {}
        for ln in out.strip().splitlines():
            idx, bdf = [x.strip() for x in ln.split(",")]
            bdf = bdf.lower()
            if len(bdf.split(":")[0]) == 8:  # nvidia-smi prints an 8-digit domain, sysfs uses 4
                bdf = bdf[4:]
            base = "/sys/bus/pci/devices/" + bdf
            lists[int(idx)] = parse_cpulist(open(base + "/local_cpulist").read())
            try:
                nodes[int(idx)] = int(open(base + "/numa_node").read())
            except Exception:
                nodes[int(idx)] = None
        allowed = os.sched_getaffinity(0)
        mine = sorted(lists[local] & allowed) or sorted(allowed)
        peers = [g for g in range(world) if g in lists and lists[g] == lists[local]]
        share = len(mine) // max(1, len(peers))
        if share >= 2 and local in peers:
            k = peers.index(local)
            mine = mine[k * share:(k + 1) * share]
        os.sched_setaffinity(0, mine)
        return {"cpus": "%d-%d (%d)" % (mine[0], mine[-1], len(mine)), "numa_node": nodes.get(local),
                "gpus_sharing_the_core_list": len(peers)}
    except Exception as exc:  # no sysfs / nvidia-smi: leave the affinity alone
        return {"error": str(exc)[:120]}


def make_config5(torch, chain, cfg, dev, rank, world, total_scans, batch):
    """Inputs of the BASELINE configs[4] leg for this rank: its `total_scans / world` scans of 65536 points in
    batches of `batch`.  Geometry (scans, intensities, masks, proposal boxes, NMS boxes) is distinct per scan
    (seed = 5000 + global scan index) and generated on the host like every other input; the stand-ins for the
    MLP feature maps are ONE set of resident tensors shared by all batches (device-generated N(0,1); their values
    do not influence any kernel's work, and in the reference pipeline they are device-resident MLP outputs)."""
    import numpy as np

    from heterofusionrcnn_b200 import synthetic

    cfg5 = cfg.scaled(65536)
    per_rank = total_scans // world
    nb = per_rank // batch
    first = rank * per_rank
    geo = []
    for k in range(nb):
        xs, its, bx, nm, mk = [], [], [], [], []
        for i in range(batch):
            g = first + k * batch + i
            x, it = synthetic.make_scan(5000 + g, cfg5.points)
            xs.append(x)
            its.append(it)
            bx.append(synthetic.boxes3d_to_corners(synthetic.make_boxes3d(7000 + g, x, cfg5.crop_boxes)))
            nm.append(synthetic.make_nms_boxes(9000 + g, cfg5.nms_boxes))
            mk.append(np.random.default_rng(11000 + g).random(cfg5.points) < 0.3)
        geo.append({"xyz": torch.from_numpy(np.stack(xs)).to(dev), "intensity": torch.from_numpy(np.stack(its)).to(dev),
                    "fg_mask": torch.from_numpy(np.stack(mk)).to(dev), "crop_boxes": torch.from_numpy(np.concatenate(bx)).to(dev),
                    "nms_boxes": torch.from_numpy(np.stack(nm)).to(dev)})
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    n_lv = [cfg5.points] + list(cfg5.npoint)
    rnd = lambda *shape: torch.randn(shape, dtype=torch.float32, device=dev, generator=gen)  # noqa: E731
    static = {k: v.clone() for k, v in geo[0].items()}
    static["box_ind"] = torch.arange(batch, dtype=torch.int32, device=dev).repeat_interleave(cfg5.crop_boxes)
    static["sa_feats"] = [None] + [rnd(batch, n_lv[l], cfg5.sa_channels[l]) for l in range(1, 4)]
    static["fp_feats"] = [rnd(batch, n_lv[4 - j], cfg5.fp_channels[j]) for j in range(4)]
    static["crop_fts"] = rnd(batch, cfg5.points, cfg5.crop_channels)
    return cfg5, static, geo


def run_b200(args):
    import torch
    import torch.distributed as dist

    import heterofusionrcnn_b200 as hf
    from heterofusionrcnn_b200 import chain

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hfops operators have no CPU fallback "
                         "(use --impl reference for the CPU path)")
    hf.load()  # raises if libhfops.so is missing
    binding = bind_rank_to_local_cores(local, world) if world > 1 else None  # before the first CUDA allocation
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    cfg = make_cfg(args)
    host_np = chain.make_host_inputs(cfg, batch=args.batch, seed=2 + rank)
    arena = chain.InputArena(host_np)  # ONE pinned buffer holding every input tensor of a step
    resident_buf = arena.device_buffer(dev)
    arena.upload(resident_buf)
    resident = arena.views(resident_buf)
    torch.cuda.synchronize()
    fp32 = measure_fp32_peak(torch, hf, dev)
    FP32_PEAK_TFLOPS = fp32["peak"]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- resident-input throughput (value) + per-launch events (roofline) ----
    overlap = not args.serial
    use_graph = overlap and not args.no_graph and not args.timeline
    depth = max(1, args.resident_inflight) if overlap else 1
    probe = EventProbe(torch)
    probe.enabled = False
    pipe = chain.ChainPipeline(cfg, depth=depth, overlap=overlap)
    for _ in range(max(args.warmup, 3)):
        pipe.submit(resident, probe)
    pipe.drain()
    torch.cuda.synchronize()
    graph = chain.ChainGraph(resident, cfg) if use_graph else None
    if graph is not None:
        for _ in range(max(args.warmup, 3)):
            graph.replay()
        torch.cuda.synchronize()
    barrier()
    clocks = ClockSampler(local).start() if rank == 0 else None
    l0 = hf.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if args.profile_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    ev0.record()
    if graph is not None:
        for _ in range(args.steps):
            graph.replay()  # one cudaGraphLaunch = the whole step: every lane, every libhfops kernel
    else:
        probe.enabled = True
        for _ in range(args.steps):
            pipe.submit(resident, probe)
        pipe.drain()
    ev1.record()
    if args.profile_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    barrier()
    launches = (graph.launches * args.steps) if graph is not None else hf.launch_count() - l0
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = int(sum_over_ranks(launches))
    if graph is not None:
        # per-launch durations inside an overlapped step cannot be read out of a graph replay: measure them on
        # eagerly issued steps (same streams, same kernels) right after the timed region
        probe.enabled = True
        for _ in range(args.steps):
            pipe.submit(resident, probe)
        pipe.drain()
        torch.cuda.synchronize()
    ops = probe.summarise(args.steps)
    probe.enabled = False
    if args.timeline and rank == 0:
        per = len(probe.records) // max(args.steps, 1)
        last = probe.records[-per:]
        t0 = min(ev0.elapsed_time(r[3]) for r in last)
        for key, _, _, s_, e_ in sorted(last, key=lambda r: ev0.elapsed_time(r[3])):
            sys.stderr.write("timeline %-34s %8.3f -> %8.3f ms\n" % (key, ev0.elapsed_time(s_) - t0, ev0.elapsed_time(e_) - t0))
    # each launch timed ALONE (one stream, program order) — outside the timed region, for the per-kernel table
    alone = EventProbe(torch)
    for _ in range(2):  # untimed: the caching allocator has no blocks on this stream yet
        chain.run_chain(resident, cfg, None, overlap=False)
    torch.cuda.synchronize()
    for _ in range(min(args.steps, 5)):
        chain.run_chain(resident, cfg, alone, overlap=False)
    torch.cuda.synchronize()
    alone_ms = {o["launch"]: o["ms"] for o in alone.summarise(min(args.steps, 5))}
    serial_step_ms = sum(o["ms"] * o["per_step"] for o in alone.summarise(min(args.steps, 5)))
    # distance evaluations the pruned ball query / three_nn launches EXECUTE (the kernels count them when a device
    # counter is set): one more serial pass, outside every timed region
    pair_evals = count_pair_evals(torch, hf, chain, resident, cfg)
    b2b_ms = back_to_back(torch, chain, resident, cfg)

    # ---- end to end: host buffers in, host-consumed results out, through the public op API ----
    e2e = None
    if not args.no_e2e:
        depth = max(1, args.inflight) if overlap else 1
        dev_bufs = [arena.device_buffer(dev) for _ in range(depth)]
        for b_ in dev_bufs:  # every arena holds the whole step once (the host-originated-only leg keeps the features)
            arena.upload(b_)
        dev_views = [arena.views(b_) for b_ in dev_bufs]
        torch.cuda.synchronize()
        main = torch.cuda.current_stream(dev)
        out_streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]

        def e2e_leg(host_only):
            """`steps` passes, each: ONE host->device copy of the step's inputs (whole arena, or only the
            host-originated prefix) -> the chain through the public op API -> device->host copies of the
            host-consumed results.  `depth` batches in flight; CUDA events on the current stream, max over ranks."""
            pipe = chain.ChainPipeline(cfg, depth=depth, overlap=overlap)
            out_host = [None] * depth
            sent = 0
            warm = max(args.warmup, 3)
            for it in range(warm + args.steps):
                if it == warm:
                    for st in out_streams:
                        main.wait_stream(st)
                    pipe.drain()
                    barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    t0 = time.perf_counter()
                    e0.record()
                slot = pipe.next_slot()
                pipe.release(slot)  # the pass that read this arena last is over: it may be overwritten
                sent = arena.upload(dev_bufs[slot], host_originated_only=host_only)  # current stream: overlaps the previous batch's kernels
                res = pipe.submit(dev_views[slot], None)
                st = out_streams[slot]
                for e in res["_done"]:
                    st.wait_event(e)
                if not res["_done"]:
                    st.wait_stream(main)
                with torch.cuda.stream(st):  # device->host copies of this batch's host-consumed results
                    outs = chain.host_results(res, cfg)
                    if out_host[slot] is None:
                        out_host[slot] = [torch.empty(o.shape, dtype=o.dtype, pin_memory=True) for o in outs]
                    for o, h in zip(outs, out_host[slot]):
                        h.copy_(o, non_blocking=True)
            for st in out_streams:
                main.wait_stream(st)
            pipe.drain()
            e1.record()
            barrier()
            wall_ms = (time.perf_counter() - t0) * 1e3
            ms_ = max_over_ranks(max(e0.elapsed_time(e1), 0.0))
            del pipe
            return {"value": args.batch * n_gpus * args.steps / (ms_ / 1e3), "unit": UNIT, "h2d_bytes_per_step": sent,
                    "d2h_bytes_per_step": nbytes_of(out_host[0]), "ms_per_step": ms_ / args.steps,
                    "wall_ms_per_step": wall_ms / args.steps}

        def copy_only_leg():
            """The step's host->device copies alone (no kernels): what the host side of this box delivers per GPU
            when `world` ranks copy at once."""
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            for it in range(args.steps):
                arena.upload(dev_bufs[it % depth])
            c1.record()
            barrier()
            ms_ = max_over_ranks(c0.elapsed_time(c1))
            return {"gbs_per_gpu": arena.nbytes * args.steps / (ms_ / 1e3) / 1e9, "ms_per_step": ms_ / args.steps,
                    "ranks_copying_at_once": world}

        e2e = e2e_leg(host_only=False)
        e2e.update({
            "copied_in": "every input tensor of the step (scans, stand-in feature maps, masks, boxes): one pinned arena, ONE "
                         "cudaMemcpyAsync per step",
            "copied_out": "FPS indices (4 levels), ball-query counts, NMS keep lists + counts, crop non_empty flags",
            "pipelining": "%d batches in flight: a batch's host->device copy overlaps the previous batch's kernels" % depth,
            "copy_only": copy_only_leg(),
            "binding": binding,
        })
        ho = e2e_leg(host_only=True)
        ho["copied_in"] = ("only what a caller of the reference pipeline holds on the host: scans, intensities, foreground masks, "
                           "proposal boxes + box_ind, NMS boxes; the MLP feature maps stay resident (in the reference they are "
                           "produced on the device, rcnn_model.py:478-489)")
        e2e["host_originated_only"] = ho
        del dev_bufs, dev_views

    # ---- BASELINE configs[4]: 128 scans x 65536 points in total, strong scaling over the ranks ----
    config5 = None
    if not args.no_config5 and args.points == 16384 and args.config5_scans % (args.batch * world) == 0:
        torch.cuda.synchronize()
        cfg5, static5, geo5 = make_config5(torch, chain, cfg, dev, rank, world, args.config5_scans, args.batch)
        g5 = chain.ChainGraph(static5, cfg5)

        def pass5():
            for g in geo5:  # this rank's batches: geometry into the graph's input buffers (stream ordered), one replay
                for k_, v_ in g.items():
                    static5[k_].copy_(v_, non_blocking=True)
                g5.replay()

        pass5()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(args.config5_steps):
            pass5()
        f1.record()
        barrier()
        ms5 = max_over_ranks(f0.elapsed_time(f1)) / args.config5_steps
        alone5 = EventProbe(torch)
        chain.run_chain(static5, cfg5, None, overlap=False)
        torch.cuda.synchronize()
        for _ in range(2):
            chain.run_chain(static5, cfg5, alone5, overlap=False)
        torch.cuda.synchronize()
        a5 = {o["launch"]: o for o in alone5.summarise(2)}
        crop5 = a5.get("pc_crop_and_sample/c=%d" % cfg5.crop_channels)
        hbm_peak_, _ = peaks()
        config5 = {
            "workload": "configs[4]: the same chain on %d scans x %d points in total, batch-sharded: every rank runs its %d scans "
                        "in batches of %d (one graph replay per batch)" % (args.config5_scans, cfg5.points,
                                                                         args.config5_scans // world, args.batch),
            "scaling": "strong", "n_gpus": n_gpus, "total_scans": args.config5_scans, "steps": args.config5_steps,
            "scans_per_s": args.config5_scans / (ms5 / 1e3), "ms_per_pass": ms5, "ms_per_scan": ms5 / args.config5_scans,
            "ms_per_batch": ms5 / len(geo5),
            "fps_ms": a5["fps/level=1"]["ms"] if "fps/level=1" in a5 else None,
            "crop_ms": crop5["ms"] if crop5 else None,
            "crop_frac": (crop5["bytes"] / (crop5["ms"] / 1e3) / 1e9 / hbm_peak_) if crop5 else None,
            "gpu_launches_per_batch": g5.launches,
            "features": "stand-in feature maps: one resident set shared by all batches; geometry distinct per scan",
        }
        del g5, static5, geo5, alone5
        torch.cuda.empty_cache()
    clk = clocks.stop() if clocks else None

    # the only cross-rank step of the path: the final host gather of the host-consumed results (untimed)
    from heterofusionrcnn_b200 import shard

    torch.cuda.synchronize()
    fin = chain.run_chain(resident, cfg, overlap=overlap)
    torch.cuda.synchronize()
    gathered = shard.gather_to_host({"nms_keep": fin["nms_keep"].cpu().numpy(), "nms_num": fin["nms_num"].cpu().numpy()},
                                    rank, world)
    if rank == 0:
        total = shard.concat_scene_results(gathered)
        assert total["nms_keep"].shape == (args.batch * n_gpus, cfg.post_nms), total["nms_keep"].shape
    del fin

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    hbm_peak, hbm_src = peaks()
    def rate(o, ms_launch):
        sec = ms_launch / 1e3
        if o["bound"] == "hbm" and sec > 0:
            return o["bytes"] / sec / 1e9, "GB/s", hbm_peak
        if o["bound"] == "fp32" and sec > 0 and o["flops"]:
            return o["flops"] / sec / 1e12, "TFLOP/s", FP32_PEAK_TFLOPS
        return None, None, None

    for o in ops:
        o["achieved"], o["unit"], o["peak"] = rate(o, o["ms"])
        o["frac"] = (o["achieved"] / o["peak"]) if o["achieved"] else None
        o["ms_alone"] = alone_ms.get(o["launch"])
        a = rate(o, o["ms_alone"])[0] if o["ms_alone"] else None
        o["frac_alone"] = (a / o["peak"]) if a and o["peak"] else None
        # share of the serial (one-stream) step: with overlap the in-region durations add up to more than the step
        o["share_of_step"] = (o["ms_alone"] or o["ms"]) * o["per_step"] / serial_step_ms
        o["ms_back_to_back"] = b2b_ms.get(o["launch"])
        if o["ms_back_to_back"] and o["ms_alone"] and o["ms_back_to_back"] >= o["ms_alone"]:
            o["ms_back_to_back"] = None  # launches of ~15 us: the host cannot issue them back to back, nothing to report
        if o["ms_back_to_back"] and o["bound"] == "hbm":
            o["frac_back_to_back"] = o["bytes"] / (o["ms_back_to_back"] / 1e3) / 1e9 / hbm_peak
        # pruned launches: `frac` above divides the EVERY-PAIR flop count of SURVEY 8d by the pruned kernel's time (an
        # equivalent rate: it can exceed 1); frac_executed divides the evaluations the kernel really ran
        ex = pair_evals.get(o["launch"], 0)
        o["pair_evals_executed"] = ex or None
        o["pair_evals_every_pair"] = (o["meta"]["b"] * o["meta"]["n"] * o["meta"]["m"]) if ex else None
        if ex and o["flops"] and o["ms_alone"]:
            per_pair = o["flops"] / o["pair_evals_every_pair"]
            o["frac_equivalent"], o["frac_alone_equivalent"] = o["frac"], o["frac_alone"]
            o["frac"] = per_pair * ex / (o["ms"] / 1e3) / 1e12 / FP32_PEAK_TFLOPS
            o["frac_alone"] = per_pair * ex / (o["ms_alone"] / 1e3) / 1e12 / FP32_PEAK_TFLOPS
            o["achieved"] = per_pair * ex / (o["ms"] / 1e3) / 1e12
    ops.sort(key=lambda o: -(o["ms_alone"] or o["ms"]) * o["per_step"])
    own = [o for o in ops if o["bound"] != "other"]
    top_fps = max((o for o in own if o["launch"].startswith("fps/")), key=lambda o: o["ms"])
    top_hbm = max((o for o in own if o["bound"] == "hbm"), key=lambda o: o["ms"])

    step_bytes = sum(o["bytes"] * o["per_step"] for o in ops)

    def roof(o):
        default_cfg = args.batch == 16 and args.points == 16384 and not args.msg_radii
        traffic, traffic_source = ncu_traffic(o["launch"], hf.build_id()) if default_cfg else (None, "not the default configuration")
        return {"kernel": o["launch"], "bound": o["bound"], "achieved": o["achieved"], "peak": o["peak"],
                "unit": o["unit"], "frac": o["frac"], "traffic": traffic, "traffic_source": traffic_source,
                "ms_per_launch": o["ms"],
                "ms_per_launch_alone": o["ms_alone"], "frac_alone": o["frac_alone"],
                "algorithmic_bytes": o["bytes"], "algorithmic_flops": o["flops"],
                "peak_source": hbm_src if o["bound"] == "hbm" else
                               "measured in this run: hf_fp32_probe, FFMA %.1f / FFMA2 %.1f TFLOP/s (derived 148 SM x 128 lanes x 2 x "
                               "1.965 GHz = %.1f)" % (fp32["ffma"], fp32["ffma2"], fp32["derived"]),
                "share_of_step": o["share_of_step"],
                "note": "frac: this launch inside an overlapped step (issued eagerly right after the timed graph replays), where "
                        "it shares the GPU with the other 12 streams - the crop at the lowest priority; frac_alone: the same "
                        "launch with the GPU to itself"}

    line = {
        "metric": METRIC, "value": args.batch * n_gpus * args.steps / (ms / 1e3), "unit": UNIT, "n_gpus": n_gpus,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, cfg, n_gpus), "clocks": clk, "e2e": e2e, "gpu_launches": launches,
        # `roofline` = the dominant HBM-bound kernel (the RoI crop writes 4.9 of the step's 5.6 GB); the other
        # quarter of the step is the level-1 FPS, a latency chain with no HBM traffic: `roofline_fps`
        "roofline": roof(top_hbm), "roofline_fps": roof(top_fps),
        # the whole step against the HBM roofline: algorithmic bytes of every launch of the step / step time
        "roofline_step": {"bound": "hbm", "algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms / args.steps) / 1e6,
                          "peak": hbm_peak, "unit": "GB/s", "frac": step_bytes / (ms / args.steps) / 1e6 / hbm_peak},
        "roofline_ops": [{k: o.get(k) for k in ("launch", "bound", "ms", "ms_alone", "per_step", "achieved", "unit", "frac",
                                                "frac_alone", "share_of_step", "pair_evals_executed", "pair_evals_every_pair",
                                                "frac_equivalent", "frac_alone_equivalent", "ms_back_to_back",
                                                "frac_back_to_back")
                          if o.get(k) is not None or k in ("frac", "frac_alone", "achieved", "unit")} for o in ops],
        "serial_ms_per_step": serial_step_ms,
        "fp32_peak": fp32,
        "config5": config5,
        "build": hf.version(),
    }
    if not args.no_cpu_baseline and world == 1:  # rank 0 at N=1 only
        cb, _ = time_cpu(args, cfg, 1, 0, single_thread=True)
        line["cpu_baseline"] = cb
    if not args.no_side_legs and world == 1:
        line["side_legs"] = side_legs(torch, hf, dev, cfg, args.batch)
    if args.table:
        for o in ops:
            sys.stderr.write("%-34s %-5s %9.4f ms (alone %9.4f) x%d  %s %s  frac=%s  share=%.3f\n" % (
                o["launch"], o["bound"], o["ms"], o["ms_alone"] or 0.0, o["per_step"],
                ("%.1f" % o["achieved"]) if o["achieved"] else "-", o["unit"] or "",
                ("%.3f" % o["frac"]) if o["frac"] else "-", o["share_of_step"]))
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


NCU_KERNEL_OF = {  # probe name -> (kernel name fragment, grid) in the committed ncu --set full summary
    "pc_crop_and_sample/c=288": ("crop_kernel", "(8192, 1, 1)"),
    "fps/level=1": ("fps_multi", None),
}


def ncu_traffic(launch, build_id):
    """-> (dram__bytes_read.sum + dram__bytes_write.sum of one launch of `launch`, source description).

    The figure comes from one `ncu --set full` capture of this same configuration (tools/gpu_pass.sh +
    tools/ncu_summary.py -> profiles/r*_ncu_full_*_summary.csv, newest first).  A capture is only used if its first
    line names the build it was taken from (`build=<id>`, heterofusionrcnn_b200/build.py:source_id) and that is the
    library running now; otherwise traffic is None and the source says why - a stale figure is never printed next
    to freshly timed numbers."""
    import csv
    import glob
    import re

    if launch not in NCU_KERNEL_OF:
        return None, "no capture is kept for this launch"
    frag, grid = NCU_KERNEL_OF[launch]

    # every committed summary names the build it was taken from in its first line; use the one of THIS build
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_*_summary.csv")), key=os.path.getmtime)
    if not files:
        return None, "no capture under profiles/"
    path, rows, seen = None, None, []
    for f in files:
        first = open(f).readline()
        m = re.search(r"build=([0-9a-f]{12})", first)
        seen.append(m.group(1) if m else "-")
        if m and m.group(1) == build_id:
            path = f
    if path is None:
        return None, "no committed ncu --set full summary was taken from build %s (captures of builds %s): not used" % (
            build_id, ", ".join(sorted(set(seen))))
    rows = list(csv.reader(open(path)))
    name = "profiles/" + os.path.basename(path)
    hdr, units = rows[1], rows[2]
    ci = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in rows[3:]:
        if frag in r[ci["Kernel Name"]] and (grid is None or r[ci["Grid Size"]] == grid):
            tot = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(r[ci[k]]) * scale[units[ci[k]]]
            return tot, "%s (ncu --set full of build %s, kernel %s grid %s)" % (name, build_id, r[ci["Kernel Name"]], r[ci["Grid Size"]])
    return None, "%s holds no launch of %s with grid %s: not used" % (name, frag, grid)


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
