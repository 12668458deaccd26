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
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # derived: SMs x lanes x FMA x max clock (no measured figure)


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
                 "independent branches of the chain on 13 CUDA streams%s, %d batch(es) in flight (e2e leg: eager, %d)" % (
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
    if name == "gather_point":
        return "hbm", 4 * m["b"] * m["m"] * 7, 0
    if name == "query_ball_point":
        return "fp32", 12 * m["b"] * (m["n"] + m["m"]) + 4 * m["b"] * m["m"] * (m["ns"] + 1), 9 * m["b"] * m["m"] * m["n"]
    if name == "group_point":
        return "hbm", 4 * (m["b"] * m["m"] * m["ns"] * m["c"] + m["b"] * m["m"] * m["ns"] + m["b"] * m["n"] * m["c"]), 0
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
                        "bytes": nbytes, "flops": flops})
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


def cpu_sample_size(args, cores):
    if args.cpu_scenes > 0:
        return min(args.cpu_scenes, args.batch)
    return max(1, min(args.batch, cores // 8, 4))


def time_cpu(args, cfg, host, steps, warmup):
    from oracle import binding as O

    O.build()
    cores = os.cpu_count() or 1
    O.set_threads(cores)
    scenes = cpu_sample_size(args, cores)
    for _ in range(warmup):
        cpu_chain(O, host, cfg, scenes)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_chain(O, host, cfg, scenes)
    dt = (time.perf_counter() - t0) / steps
    kind = "port"
    return {"value": scenes / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%d of the %d scans of one batch, whole chain, oracle/hf_oracle.cpp (g++ -O2) with %d std::threads "
                      "over scenes/rows; %.2f s per pass" % (scenes, args.batch, cores, dt)}, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from heterofusionrcnn_b200 import chain

    cfg = make_cfg(args)
    cores = os.cpu_count() or 1
    scenes = cpu_sample_size(args, cores)
    host = chain.make_host_inputs(cfg, batch=scenes, seed=2)
    cb, dt = time_cpu(args, cfg, host, args.steps, args.warmup)
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
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    cfg = make_cfg(args)
    host_np = chain.make_host_inputs(cfg, batch=args.batch, seed=2 + rank)
    pinned = chain.to_pinned(host_np)
    resident = chain.to_device_from_pinned(pinned, dev)
    torch.cuda.synchronize()

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

    # ---- end to end: host buffers in, host-consumed results out, through the public op API ----
    e2e = None
    if not args.no_e2e:
        h2d = nbytes_of(pinned)
        depth = max(1, args.inflight) if overlap else 1
        pipe = chain.ChainPipeline(cfg, depth=depth, overlap=overlap)
        out_streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
        out_host = [None] * depth
        main = torch.cuda.current_stream(dev)
        for it in range(max(args.warmup, 3) + args.steps):
            if it == max(args.warmup, 3):
                for st in out_streams:
                    main.wait_stream(st)
                pipe.drain()
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                e0.record()
            slot = it % depth
            d = chain.to_device_from_pinned(pinned, dev)  # on the current stream: overlaps the previous batch's kernels
            res = pipe.submit(d, None)
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
        e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), 0.0))
        e2e = {"value": args.batch * n_gpus * args.steps / (e2e_ms / 1e3), "unit": UNIT,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": nbytes_of(out_host[0]), "ms_per_step": e2e_ms / args.steps,
               "wall_ms_per_step": wall_ms / args.steps,
               "copied_in": "every input tensor of the step (scans, stand-in feature maps, masks, boxes) from pinned host memory",
               "copied_out": "FPS indices (4 levels), ball-query counts, NMS keep lists + counts, crop non_empty flags",
               "pipelining": "%d batches in flight: a batch's host->device copies overlap the previous batch's kernels" % depth}
        del pipe
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
    ops.sort(key=lambda o: -(o["ms_alone"] or o["ms"]) * o["per_step"])
    own = [o for o in ops if o["bound"] != "other"]
    top_fps = max((o for o in own if o["launch"].startswith("fps/")), key=lambda o: o["ms"])
    top_hbm = max((o for o in own if o["bound"] == "hbm"), key=lambda o: o["ms"])

    step_bytes = sum(o["bytes"] * o["per_step"] for o in ops)

    def roof(o):
        default_cfg = args.batch == 16 and args.points == 16384 and not args.msg_radii
        return {"kernel": o["launch"], "bound": o["bound"], "achieved": o["achieved"], "peak": o["peak"],
                "unit": o["unit"], "frac": o["frac"], "traffic": ncu_traffic(o["launch"]) if default_cfg else None,
                "ms_per_launch": o["ms"],
                "ms_per_launch_alone": o["ms_alone"], "frac_alone": o["frac_alone"],
                "algorithmic_bytes": o["bytes"], "algorithmic_flops": o["flops"],
                "peak_source": hbm_src if o["bound"] == "hbm" else "derived 148 SM x 128 lanes x 2 x 1.965 GHz (not measured)",
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
        "roofline_ops": [{k: o[k] for k in ("launch", "bound", "ms", "ms_alone", "per_step", "achieved", "unit", "frac",
                                            "frac_alone", "share_of_step")} for o in ops],
        "serial_ms_per_step": serial_step_ms,
    }
    if not args.no_cpu_baseline:
        cb, _ = time_cpu(args, cfg, host_np, 1, 0)
        line["cpu_baseline"] = cb
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
    "fps/level=1": ("fps_multi_ker", None),
}


def ncu_traffic(launch):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `launch`, from the newest
    profiles/r*_ncu_full_*_summary.csv (one ncu --set full capture of this same configuration,
    tools/gpu_pass.sh + tools/ncu_summary.py); None if the kernel is not in it."""
    import csv
    import glob

    if launch not in NCU_KERNEL_OF:
        return None
    frag, grid = NCU_KERNEL_OF[launch]
    import re

    def version(path):  # r<round>_ncu_full_v<pass>_summary.csv -> (round, pass): numeric, not lexicographic
        m = re.search(r"r(\d+)_ncu_full_v(\d+)_summary", os.path.basename(path))
        return (int(m.group(1)), int(m.group(2))) if m else (0, 0)

    files = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r*_ncu_full_*_summary.csv")),
                   key=version)
    if not files:
        return None
    rows = list(csv.reader(open(files[-1])))
    hdr, units = rows[1], rows[2]
    ci = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in rows[3:]:
        if frag in r[ci["Kernel Name"]] and (grid is None or r[ci["Grid Size"]] == grid):
            tot = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(r[ci[k]]) * scale[units[ci[k]]]
            return tot
    return None


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
