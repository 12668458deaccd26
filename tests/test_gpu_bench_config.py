"""GPU: parity AT THE BENCHMARKED CONFIGURATION, THROUGH THE BENCHMARKED CODE PATH.

bench.py times chain.ChainGraph (one replay of the captured CUDA graph = one step) at 16 scans x 16384
points with the default ChainConfig.  This file builds exactly that object on exactly those inputs
(chain.make_host_inputs(cfg, 16, seed=2) is bench.py's rank-0 batch), replays it, and compares the tensors
the replay produced with the UNMODIFIED reference kernels (oracle/_ref/libref_cuda.so, compiled from
/root/reference/*/*_g.cu for sm_100a) run on the same GPU on the same inputs:

    FPS indices of all 4 levels, ball-query idx + pts_cnt of all 4 levels, group_point outputs, three_nn
    idx + dist of all 4 levels, three_interpolate outputs, the NMS keep list + count of all 16 scenes
    (reference mask kernel + the reference's host sweep, bev_iou.cpp:87-112) and the crop (empty flags and,
    where the reference's atomicInc slot race cannot matter, the set of cropped points).

Also here: the BASELINE configs[4] cloud size, FPS 65536 -> 16384, against the reference kernel at its real m.
"""
import numpy as np
import pytest
import torch

from heterofusionrcnn_b200 import chain, synthetic
from heterofusionrcnn_b200.sampling.tf_sampling import farthest_point_sample

import _refcuda

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not _refcuda.available(), reason="oracle/_ref/libref_cuda.so not built")


def N(t):
    return t.detach().cpu().numpy()


def close(a, b, rtol=1e-5):
    a, b = a.double(), b.double()
    return bool(torch.all((a - b).abs() <= rtol * torch.maximum(a.abs(), b.abs()) + 1e-12))


@needs_ref
def test_bench_chain_graph_matches_reference_kernels(cuda, oracle):
    cfg = chain.ChainConfig()
    batch = 16
    host = chain.make_host_inputs(cfg, batch=batch, seed=2)
    dev = chain.to_device(host, cuda)
    check_chain_graph_against_reference_kernels(cuda, oracle, cfg, dev, batch, min_race_free=1000)


@needs_ref
def test_bench_config5_chain_graph_matches_reference_kernels(cuda, oracle):
    """The object bench.py's `config5` leg replays (BASELINE configs[4]): ChainGraph over 16 scans x 65536 points, built
    by bench.make_config5 exactly as rank 0 of a 1-GPU run builds its first batch - two-wave sorted FPS 65536 -> 16384,
    the four-part cloud index, pruned search on three levels, the two-level crop select at C = 288 / R = 512 with hit
    lists four times longer than at 16384 points.  Same comparison, output by output, with the unmodified reference
    kernels (they take any n: tf_sampling_g.cu:105-170, tf_grouping_g.cu:3-36, tf_cropping_g.cu:43-132)."""
    import bench

    batch = 16
    cfg5, static5, geo5 = bench.make_config5(torch, chain, chain.ChainConfig(), cuda, 0, 1, batch, batch)
    assert cfg5.points == 65536 and len(geo5) == 1 and static5["xyz"].shape == (batch, 65536, 3)
    del geo5
    check_chain_graph_against_reference_kernels(cuda, oracle, cfg5, static5, batch, min_race_free=1000)


def check_chain_graph_against_reference_kernels(cuda, oracle, cfg, dev, batch, min_race_free):
    graph = chain.ChainGraph(dev, cfg)
    assert graph.launches == chain.chain_launches(cfg)
    for _ in range(3):
        res = graph.replay()
    torch.cuda.synchronize()

    # ---- set abstraction: every level against the reference kernels, fed with the reference's own outputs ----
    xyz = dev["xyz"]
    feats = [dev["intensity"]] + list(dev["sa_feats"][1:])
    levels = [xyz]
    for l in range(4):
        m = cfg.npoint[l]
        ridx = _refcuda.fps(m, xyz)
        assert torch.equal(res["fps_idx"][l], ridx), "fps level %d" % (l + 1)
        new_xyz = _refcuda.gather_point(xyz, ridx)
        assert torch.equal(res["xyz_levels"][l + 1], new_xyz), "gather_point level %d" % (l + 1)
        rgidx, rcnt = _refcuda.query_ball_point(cfg.radius[l], cfg.nsample, xyz, new_xyz)
        gidx, cnt, gxyz, gfeat = res["sa%d" % (l + 1)]
        assert int(rcnt.min()) >= 1  # every query is a dataset point: no uninitialised reference rows
        assert torch.equal(cnt, rcnt), "ball query count level %d" % (l + 1)
        assert torch.equal(gidx, rgidx), "ball query idx level %d" % (l + 1)
        assert torch.equal(gxyz, _refcuda.group_point(xyz, rgidx)), "group_point xyz level %d" % (l + 1)
        assert torch.equal(gfeat, _refcuda.group_point(feats[l], rgidx)), "group_point features level %d" % (l + 1)
        levels.append(new_xyz)
        xyz = new_xyz

    # ---- feature propagation ----
    for j in range(4):
        lvl = 3 - j
        dist, nidx, out = res["fp%d" % (lvl + 1)]
        rd, ri = _refcuda.three_nn(levels[lvl], levels[lvl + 1])
        assert torch.equal(nidx, ri), "three_nn idx level %d" % (lvl + 1)
        assert torch.equal(dist, rd), "three_nn dist level %d" % (lvl + 1)
        # weights: hf/core/feature_extractors/pointnet_util.py:304-307 in float32 (torch ops on the reference
        # distances); positive terms only, so 1e-5 relative per element is the north_star bar
        d = torch.clamp(rd, min=1e-10)
        w = (1.0 / d) / (1.0 / d).sum(dim=2, keepdim=True)
        gw = res["fp_weights"][lvl]
        assert close(gw, w), "interpolation weights level %d" % (lvl + 1)
        # interpolation: the reference kernel on the same indices and the same weights.  A row is a signed sum
        # of three products, so the bound is 1e-5 of the magnitude of its terms (cancellation can leave a result
        # far smaller than its rounding error); in practice the two kernels agree bit for bit.
        pts = dev["fp_feats"][j]
        rout = _refcuda.three_interpolate_channel_first(pts.transpose(1, 2).contiguous(), ri, gw.contiguous())
        rout = rout.transpose(1, 2)
        mag = torch.zeros_like(out)
        for k in range(3):
            rows = torch.gather(pts, 1, ri[:, :, k].long()[:, :, None].expand(-1, -1, pts.shape[2]))
            mag += (gw[:, :, k:k + 1] * rows).abs()
        assert bool(torch.all((out - rout).abs() <= 1e-5 * mag + 1e-12)), "three_interpolate level %d" % (lvl + 1)

    # ---- NMS: reference mask kernel + the reference's host sweep, every scene of the batch ----
    keep, num = N(res["nms_keep"]), N(res["nms_num"])
    margin = []
    for s in range(batch):
        rmask = N(_refcuda.nms_mask(dev["nms_boxes"][s], cfg.nms_thresh)).view(np.uint64)
        want, kept = oracle.nms_sweep(rmask)
        assert int(num[s]) == min(kept, cfg.post_nms), "nms count scene %d" % s  # the sweep stops at post_nms kept boxes
        assert np.array_equal(keep[s], want[: cfg.post_nms]), "nms keep scene %d" % s
        margin.append(kept)
    assert min(margin) > 0

    # ---- crop: flags exactly; the set of cropped points where the reference's slot race cannot matter ----
    args = (dev["xyz"], dev["crop_fts"], dev["intensity"], dev["fg_mask"], dev["crop_boxes"], dev["box_ind"])
    ref = _refcuda.pc_crop_and_sample(*args, cfg.crop_resize)
    got = res["crop"]
    assert torch.equal(got[5], ref[5]), "non_empty_box"
    gi, ri = N(got[4]), N(ref[4])
    gs, rs = np.sort(gi, axis=1), np.sort(ri, axis=1)
    distinct = (np.diff(gs, axis=1) != 0).sum(axis=1) + 1
    race_free = N(got[5]).astype(bool) & (distinct < cfg.crop_resize)  # fewer than R points inside: padding cycles them all
    print("crop: %d of %d boxes are free of the reference's slot race" % (race_free.sum(), len(race_free)))
    assert race_free.sum() >= min_race_free
    for k in np.nonzero(race_free)[0]:
        assert np.array_equal(np.unique(gs[k]), np.unique(rs[k])), "crop set box %d" % k
    # every output row is the gathered source row of its crop_ind (exact copies)
    bi = dev["box_ind"].long()[:, None].expand(-1, cfg.crop_resize)
    ci = got[4].long()
    ne = got[5].bool()[:, None]
    assert torch.equal(got[0], torch.where(ne[..., None], dev["xyz"][bi, ci], torch.zeros((), device=cuda)))
    assert torch.equal(got[2], torch.where(ne[..., None], dev["intensity"][bi, ci], torch.zeros((), device=cuda)))
    assert torch.equal(got[3], dev["fg_mask"][bi, ci] & ne)
    for lo in range(0, ci.shape[0], 1024):  # 4.8 GB tensor: compare in slices
        sl = slice(lo, lo + 1024)
        want = torch.where(ne[sl][..., None], dev["crop_fts"][bi[sl], ci[sl]], torch.zeros((), device=cuda))
        assert torch.equal(got[1][sl], want), "crop_fts rows %d.." % lo


@needs_ref
def test_fps_65536_matches_reference_kernel_at_full_m(cuda):
    """BASELINE configs[4] cloud size at its real m (65536 -> 16384), one plain scan and one duplicate-padded scan
    (exact distance ties), against the reference kernel (sampling/tf_sampling_g.cu:105-170; its global `temp`
    path handles any n)."""
    a, _ = synthetic.make_scans(41, 1, 65536)
    b, _ = synthetic.make_scans(42, 1, 65536, duplicate_pad=9000)
    x = torch.from_numpy(np.concatenate([a, b])).to(cuda)
    got = farthest_point_sample(16384, x)
    want = _refcuda.fps(16384, x)
    assert torch.equal(got, want)
