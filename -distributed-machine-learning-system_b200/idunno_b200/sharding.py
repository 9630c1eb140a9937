"""Range sharding of a query over the GPUs of one box, mirroring the coordinator's arithmetic.

One GPU rank plays the part of one worker VM.  Pure host logic (no CUDA); each function cites the
reference site whose behaviour it reproduces.
"""
from __future__ import annotations

RATE_FACTOR = 10          # mp4_machinelearning.py (module constant used at :509-514)
QUERY_BATCH = 400         # ALEXNET_BATCHSIZE = RESNET_BATCHSIZE = 400, mp4_machinelearning.py:45-46


def split_range(start: int, end: int, workers: list) -> list[tuple]:
    """Contiguous inclusive sub-ranges of [start, end] over ``workers`` in the given order:
    ``k_i = round(remaining_items / remaining_workers)`` with Python's banker's rounding, shard
    ``[s, s + k_i - 1]`` (mp4_machinelearning.py:516-536).  A worker may receive an empty range
    (s > e) when there are more workers than items, exactly as the reference's arithmetic does."""
    start, end = int(start), int(end)
    remaining = end - start + 1
    left = len(workers)
    out = []
    s = start
    for wkr in workers:
        k = round(remaining / left)
        out.append((wkr, s, s + k - 1))
        remaining -= k
        left -= 1
        s += k
    return out


def split_queries(start: int, end: int, batch: int = QUERY_BATCH) -> list[tuple[int, int]]:
    """Client-side chopping of ``inference start end model`` into queries:
    ``for i in range(start, end, batch): (i, i + batch - 1)`` (mp4_machinelearning.py:1104-1109).
    Like the reference, the last query may extend past ``end`` and ``end`` itself is a range()
    bound (so ``inference 0 399`` is one query 0..399)."""
    return [(i, i + batch - 1) for i in range(int(start), int(end), int(batch))]


def fair_share(model: str, avg_time: dict, num_workers: int, rate_factor: int = RATE_FACTOR) -> int:
    """Number of workers a query of ``model`` gets when AlexNet and ResNet18 jobs share the box:
    ``round(r / (r + 1) * RATE_FACTOR)`` for AlexNet and the reciprocal form for ResNet18, with
    ``r = avg_time[alexnet] / avg_time[resnet18]``, clamped to [0, num_workers]
    (mp4_machinelearning.py:504-514).  The reference never updates avg_time (it stays 100/100,
    :138,676,807); callers here pass measured per-query times."""
    t_r = avg_time.get("resnet18", 0)
    ratio = (avg_time.get("alexnet", 0) / t_r) if t_r != 0 else 1.0
    n = max(min(round(ratio / (ratio + 1) * rate_factor), num_workers), 0)
    if model != "alexnet":
        rr = 1 / ratio if ratio != 0 else 1.0
        n = max(min(round(rr / (rr + 1) * rate_factor), num_workers), 0)
    return n


def ring_successors(rank: int, world: int):
    """Ranks after ``rank`` in ring order, wrapping (utils.py:30-39 get_replica_neighbors)."""
    for k in range(1, world + 1):
        yield (rank + k) % world


def reassign(shards: list[tuple], dead: int, alive: set, world: int, coordinator: int = 0):
    """Worker-loss path (mp4_machinelearning.py:706-760): every sub-range the dead rank held moves
    WHOLE to the first ring successor that is alive and is not the coordinator; no partial progress
    is kept.  ``shards`` = [(rank, s, e), ...]; returns the new list (same order).  If only the
    coordinator is left it takes the work itself (the reference would drop it; a one-box run must
    still finish)."""
    out = []
    for (r, s, e) in shards:
        if r != dead:
            out.append((r, s, e))
            continue
        target = None
        for cand in ring_successors(dead, world):
            if cand == dead or cand not in alive or cand == coordinator:
                continue
            target = cand
            break
        if target is None:
            if coordinator in alive and coordinator != dead:
                target = coordinator
            else:
                raise RuntimeError("no alive rank left to take over the shard")
        out.append((target, s, e))
    return out
