"""CPU, world_size 2 over gloo: the batch-sharded chain (host logic of heterofusionrcnn_b200/shard.py)
gives exactly the unsharded result.  Each rank runs the CPU oracle on its slice — the
sharding / gather logic is what is under test here, the CUDA ops have their own GPU tests."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from heterofusionrcnn_b200 import chain, shard


def test_shard_bounds_are_contiguous_and_balanced():
    assert shard.shard_bounds(16, 8) == [(2 * r, 2 * r + 2) for r in range(8)]
    assert shard.shard_bounds(5, 2) == [(0, 3), (3, 5)]
    assert shard.shard_bounds(1, 4) == [(0, 1), (1, 1), (1, 1), (1, 1)]
    with pytest.raises(ValueError):
        shard.shard_bounds(4, 0)


def _small_cfg():
    cfg = chain.ChainConfig().scaled(512)
    cfg.crop_boxes, cfg.crop_resize, cfg.crop_channels, cfg.nms_boxes, cfg.post_nms = 6, 16, 8, 150, 32
    return cfg


def _oracle_pass(host, cfg):
    """Index outputs of the chain for the scenes in `host`, on the CPU oracle."""
    from oracle import binding as O

    xyz = host["xyz"]
    out = {}
    cur = xyz
    for l in range(2):
        idx = O.farthest_point_sample(cfg.npoint[l], cur)
        new = O.gather_point(cur, idx)
        bq, cnt = O.query_ball_point(cfg.radius[l] * 8, cfg.nsample, cur, new)
        out["fps%d" % l], out["bq%d" % l], out["cnt%d" % l] = idx, bq, cnt
        cur = new
    if xyz.shape[0]:
        keep = np.stack([O.oriented_nms(host["nms_boxes"][s], cfg.nms_thresh)[: cfg.post_nms]
                         for s in range(xyz.shape[0])])
    else:
        keep = np.zeros((0, cfg.post_nms), np.int32)
    out["keep"] = keep
    crop = O.pc_crop_and_sample(host["xyz"], host["crop_fts"], host["intensity"], host["fg_mask"],
                                host["crop_boxes"], host["box_ind"], cfg.crop_resize)
    return out, crop


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, batch, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = _small_cfg()
        host = chain.make_host_inputs(cfg, batch=batch, seed=9)  # every rank builds the same global batch
        mine = shard.shard_host_inputs(host, rank, world)
        scene_out, crop = _oracle_pass(mine, cfg)
        local = {"scene": scene_out, "rows": mine["_box_rows"], "crop_ind": crop[4], "crop_pts": crop[0],
                 "non_empty": crop[5]}
        dist.barrier()
        got = shard.gather_to_host(local, rank, world)
        if rank == 0:
            merged = shard.concat_scene_results([g["scene"] for g in got])
            nb = host["crop_boxes"].shape[0]
            merged["crop_ind"] = shard.scatter_box_results([g["rows"] for g in got], [g["crop_ind"] for g in got], nb)
            merged["crop_pts"] = shard.scatter_box_results([g["rows"] for g in got], [g["crop_pts"] for g in got], nb)
            merged["non_empty"] = shard.scatter_box_results([g["rows"] for g in got], [g["non_empty"] for g in got], nb)
            q.put(merged)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("batch", [4, 3])
def test_two_rank_sharded_chain_equals_unsharded(oracle, batch):
    cfg = _small_cfg()
    host = chain.make_host_inputs(cfg, batch=batch, seed=9)
    want, crop = _oracle_pass(host, cfg)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for k, v in want.items():
        assert np.array_equal(merged[k], v), k
    assert np.array_equal(merged["crop_ind"], crop[4])
    assert np.array_equal(merged["crop_pts"], crop[0])
    assert np.array_equal(merged["non_empty"], crop[5])
