"""Batch sharding of the operator chain across GPUs.

Every op of the path is independent per scene (the reference indexes scenes with
blockIdx.x everywhere; crop boxes carry `box_ind`; NMS is mapped per scene,
hf/core/models/rpn_model.py:683), so the multi-GPU plan is the simplest possible one:
contiguous batch slices, one per rank, NO collective on the data path, and a final host
gather of the per-rank results.  This module is pure host logic (numpy / torch CPU);
`torch.distributed` is only used for the final gather, on whatever backend the caller
initialised (NCCL on the GPU box, gloo in the CPU tests).
"""
import numpy as np


def shard_bounds(total, world):
    """Contiguous, balanced [lo, hi) scene ranges: the first `total % world` ranks get one more."""
    if world < 1 or total < 0:
        raise ValueError("shard_bounds(total=%r, world=%r)" % (total, world))
    base, extra = divmod(total, world)
    bounds, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        bounds.append((lo, hi))
        lo = hi
    return bounds


def shard_host_inputs(host, rank, world):
    """The slice of `chain.make_host_inputs` output that `rank` owns.

    Per-scene tensors are sliced on the batch axis; crop boxes follow their scene and their
    `box_ind` is rebased to the local batch (the op indexes the local `pts`/`fts`).
    """
    batch = host["xyz"].shape[0]
    lo, hi = shard_bounds(batch, world)[rank]
    out = {}
    for k, v in host.items():
        if k in ("crop_boxes", "box_ind"):
            continue
        if isinstance(v, list):
            out[k] = [None if x is None else x[lo:hi] for x in v]
        else:
            out[k] = v[lo:hi]
    if "box_ind" in host:
        sel = (host["box_ind"] >= lo) & (host["box_ind"] < hi)
        out["crop_boxes"] = host["crop_boxes"][sel]
        out["box_ind"] = (host["box_ind"][sel] - lo).astype(np.int32)
        out["_box_rows"] = np.nonzero(sel)[0]  # global row of every local box, for the gather
    return out


def gather_to_host(local, rank, world, dst=0):
    """Final host gather: a list (one entry per rank, rank order) of each rank's `local`
    (any picklable object of numpy arrays) on `dst`, None elsewhere.  world == 1 needs no
    process group."""
    if world == 1:
        return [local]
    import torch.distributed as dist

    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst)
    return bucket


def concat_scene_results(per_rank):
    """Concatenate per-scene result arrays (batch axis 0) gathered from the ranks, rank order
    = scene order because the shards are contiguous."""
    keys = per_rank[0].keys()
    return {k: np.concatenate([r[k] for r in per_rank], axis=0) for k in keys}


def scatter_box_results(per_rank_rows, per_rank_values, num_boxes):
    """Per-box results (crop outputs) back into the caller's global box order."""
    first = per_rank_values[0]
    out = np.zeros((num_boxes,) + first.shape[1:], dtype=first.dtype)
    for rows, vals in zip(per_rank_rows, per_rank_values):
        out[rows] = vals
    return out
