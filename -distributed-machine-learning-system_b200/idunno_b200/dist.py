"""One process per GPU: shard a query's image range the way the coordinator shards it over VMs,
run the local shard, gather the small per-image results to rank 0.

The data path has no collective (images are independent).  The only exchange is the result
return, the counterpart of the RESULT message at mp4_machinelearning.py:603-613: per image
(index:int64, top1_idx:int32, top1_prob:fp32), gathered with torch.distributed (NCCL over NVLink
on GPUs; gloo in CPU tests).
"""
from __future__ import annotations

import numpy as np

from . import sharding


def shard_for_rank(start: int, end: int, world: int, rank: int, alive=None):
    """This rank's inclusive (s, e) of the query [start, end]; ``alive`` (ordered rank list)
    defaults to all ranks.  Dead ranks get None."""
    alive = list(range(world)) if alive is None else list(alive)
    for (r, s, e) in sharding.split_range(start, end, alive):
        if r == rank:
            return (s, e)
    return None


def gather_results(indices: np.ndarray, idx: np.ndarray, prob: np.ndarray, device=None, dst: int = 0):
    """Gather ragged per-rank results to ``dst``; returns (indices, idx, prob) sorted by image index
    on dst, None elsewhere.  Uses all_gather of sizes + padded all_gather (one small collective)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size()
    rank = dist.get_rank()
    dev = device if device is not None else torch.device("cpu")
    n = torch.tensor([int(indices.shape[0])], dtype=torch.int64, device=dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    buf = torch.zeros(cap, 3, dtype=torch.float64, device=dev)   # exact for int64<2^53, int32, fp32
    if indices.shape[0]:
        buf[: indices.shape[0], 0] = torch.from_numpy(indices.astype(np.float64)).to(dev)
        buf[: indices.shape[0], 1] = torch.from_numpy(idx.astype(np.float64)).to(dev)
        buf[: indices.shape[0], 2] = torch.from_numpy(prob.astype(np.float64)).to(dev)
    outs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf)
    if rank != dst:
        return None
    rows = torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0).cpu().numpy()
    order = np.argsort(rows[:, 0], kind="stable")
    rows = rows[order]
    return rows[:, 0].astype(np.int64), rows[:, 1].astype(np.int32), rows[:, 2].astype(np.float32)


def merge_idempotent(results: dict, indices, idx, prob):
    """Result merge that tolerates a re-run shard (worker-loss path recomputes whole sub-ranges):
    later reports of an image overwrite earlier ones; values are identical by construction."""
    for i, k, p in zip(indices, idx, prob):
        results[int(i)] = (int(k), float(p))
    return results


def merge_chunks(chunks):
    """The same idempotent merge on whole arrays (the box runner's hot path: a per-image Python dict cost
    more than classifying the images): ``chunks`` = [(indices, idx, prob)] in arrival order -> (indices,
    idx, prob) sorted by image index, the LAST report of an image winning."""
    if not chunks:
        return (np.zeros(0, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32))
    ind = np.concatenate([np.asarray(c[0], dtype=np.int64) for c in chunks])
    idx = np.concatenate([np.asarray(c[1], dtype=np.int32) for c in chunks])
    prob = np.concatenate([np.asarray(c[2], dtype=np.float32) for c in chunks])
    order = np.argsort(ind, kind="stable")                 # stable: equal indices stay in arrival order
    ind, idx, prob = ind[order], idx[order], prob[order]
    last = np.ones(len(ind), dtype=bool)
    last[:-1] = ind[1:] != ind[:-1]                        # keep the last of every run of equal indices
    return ind[last], idx[last], prob[last]
