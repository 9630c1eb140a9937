"""CPU, world_size 2 over gloo: range sharding + result gather + worker-loss reassignment, with
the per-shard compute replaced by a deterministic function of the image index."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")


def _fake_infer(indices):
    idx = (np.asarray(indices, dtype=np.int64) * 7919 % 1000).astype(np.int32)
    prob = (1.0 / (1.0 + np.asarray(indices, dtype=np.float64) % 13)).astype(np.float32)
    return idx, prob


def _worker(rank, world, port, start, end, dead, out_path):
    sys.path.insert(0, PKG)
    import torch.distributed as dist
    from idunno_b200 import dist as idist, sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shards = sharding.split_range(start, end, list(range(world)))
        if dead is not None:     # the dead rank still answers the collective here, with no work
            shards = sharding.reassign(shards, dead, set(range(world)) - {dead}, world)
        mine = [i for (r, s, e) in shards if r == rank for i in range(s, e + 1)]
        indices = np.asarray(mine, dtype=np.int64)
        idx, prob = _fake_infer(indices)
        got = idist.gather_results(indices, idx, prob)
        if rank == 0:
            np.savez(out_path, indices=got[0], idx=got[1], prob=got[2])
        else:
            assert got is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("start,end,dead", [(0, 399, None), (10, 12, None), (0, 99, 1), (5, 5, None)])
def test_shard_gather_world2(tmp_path, start, end, dead):
    out = str(tmp_path / "res.npz")
    port = 29500 + (os.getpid() + start + end) % 2000
    mp.spawn(_worker, args=(2, port, start, end, dead, out), nprocs=2, join=True)
    r = np.load(out)
    want = np.arange(start, end + 1)
    assert np.array_equal(r["indices"], want)
    idx, prob = _fake_infer(want)
    assert np.array_equal(r["idx"], idx) and np.array_equal(r["prob"], prob)


def test_shard_for_rank_and_merge():
    sys.path.insert(0, PKG)
    from idunno_b200 import dist as idist
    assert idist.shard_for_rank(0, 79999, 8, 3) == (30000, 39999)
    assert idist.shard_for_rank(0, 399, 4, 2, alive=[0, 1, 3]) is None
    assert idist.shard_for_rank(0, 399, 4, 3, alive=[0, 1, 3]) == (267, 399)
    res = {}
    idist.merge_idempotent(res, [1, 2], [5, 6], [0.5, 0.25])
    idist.merge_idempotent(res, [2, 3], [6, 7], [0.25, 0.125])     # shard 2.. re-run elsewhere
    assert res == {1: (5, 0.5), 2: (6, 0.25), 3: (7, 0.125)}


def test_merge_chunks_is_the_idempotent_merge():
    from idunno_b200.dist import merge_chunks, merge_idempotent
    rng = np.random.default_rng(3)
    chunks = []
    for s, e in ((10, 19), (0, 9), (10, 19), (25, 29)):          # one shard delivered twice, a gap
        ind = np.arange(s, e + 1)
        chunks.append((ind, rng.integers(0, 1000, len(ind)).astype(np.int32), rng.random(len(ind)).astype(np.float32)))
    want = {}
    for c in chunks:
        merge_idempotent(want, *c)
    ind, idx, prob = merge_chunks(chunks)
    assert list(ind) == sorted(want)
    assert [int(k) for k in idx] == [want[i][0] for i in ind]
    assert np.array_equal(prob, np.asarray([want[i][1] for i in ind], dtype=np.float32))
    assert merge_chunks([])[0].shape == (0,)
