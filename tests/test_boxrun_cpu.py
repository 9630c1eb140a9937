"""CPU: the worker-loss-tolerant range runner (idunno_b200/boxrun.py) with real processes, one of which
dies without a goodbye in the middle of its shard.  The per-shard compute is a deterministic function of
the image index, so the merged result must equal the no-failure result exactly (SURVEY.md 8e "worker
loss", mp4_machinelearning.py:706-760)."""
import multiprocessing as mp
import os
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")
sys.path.insert(0, PKG)


def _fake(indices):
    indices = np.asarray(indices, dtype=np.int64)
    return (indices * 7919 % 1000).astype(np.int32), (1.0 / (1.0 + indices % 13)).astype(np.float32)


def _classify(model, s, e, between=None):
    out_i, out_p = [], []
    for k, b in enumerate(range(int(s), int(e) + 1, 10)):          # "tensor batches" of 10 images
        ind = np.arange(b, min(b + 10, int(e) + 1))
        i, p = _fake(ind)
        out_i.append(i); out_p.append(p)
        time.sleep(0.01)
        if between is not None:
            between(k + 1)
    return np.arange(int(s), int(e) + 1, dtype=np.int64), np.concatenate(out_i), np.concatenate(out_p)


def _worker(port, rank, die_after_batches):
    sys.path.insert(0, PKG)
    from idunno_b200 import boxrun
    cb = boxrun._exit_after(die_after_batches) if die_after_batches else None
    boxrun.worker_loop("127.0.0.1", port, rank, _classify, die_in_task=cb)


def _run(world, queries, die_rank=None, die_after=None, timeout_s=0.6, n_workers=None):
    from idunno_b200 import boxrun
    port = 23000 + (os.getpid() * 7 + int(time.time() * 10)) % 4000
    ctx = mp.get_context("fork")
    runner = boxrun.BoxRunner("127.0.0.1", port, world, _classify, seed=1, loss_timeout_s=timeout_s)
    procs = [ctx.Process(target=_worker, args=(port, r, die_after if r == die_rank else None), daemon=True)
             for r in range(1, world)]
    for p in procs:
        p.start()
    try:
        runner.wait_for_workers(30)
        t0 = time.monotonic()
        done = runner.run_queries(queries, n_workers=n_workers)
        dt = time.monotonic() - t0
    finally:
        runner.shutdown()
        for p in procs:
            p.join(timeout=5)
            if p.is_alive():
                p.kill()
    return runner, done, dt, procs


def _check(done, queries):
    for (model, qnum, s, e) in queries:
        ind, idx, prob = done[(model, qnum)]
        want = np.arange(s, e + 1)
        assert np.array_equal(ind, want)
        wi, wp = _fake(want)
        assert np.array_equal(idx, wi) and np.array_equal(prob, wp)


def test_no_failure_world4():
    q = [("resnet18", 0, 0, 399), ("resnet18", 1, 400, 799), ("resnet18", 2, 800, 1001)]
    runner, done, _, _ = _run(4, q, n_workers=4)
    _check(done, q)
    assert not runner.events
    assert runner.coord.c1()["resnet18"] > 0


def test_worker_dies_mid_shard_world4():
    q = [("resnet18", 0, 0, 799)]
    runner, done, dt, procs = _run(4, q, die_rank=2, die_after=3, n_workers=4)
    _check(done, q)
    lost = [e for e in runner.events if e[0] == "lost"]
    moved = [e for e in runner.events if e[0] == "moved"]
    assert [e[1] for e in lost] == [2]
    # whole sub-range of rank 2 went to its ring successor, rank 3 (not the coordinator)
    assert len(moved) == 1 and moved[0][3] == 3
    assert (moved[0][4], moved[0][5]) in [(0, 199), (200, 399), (400, 599), (600, 799)]   # workers are drawn at random (:517)
    assert procs[1].exitcode == 17
    assert 2 not in runner.coord.alive
    assert dt < 10


def test_last_rank_dies_successor_skips_coordinator():
    q = [("alexnet", 5, 100, 399)]
    runner, done, _, _ = _run(3, q, die_rank=2, die_after=1, n_workers=3)
    _check(done, q)
    moved = [e for e in runner.events if e[0] == "moved"]
    assert len(moved) == 1 and moved[0][3] == 1          # ring successor of 2 is 0 = coordinator: skipped


def test_two_models_concurrently_share_the_ranks():
    q = [("alexnet", i, i * 100, i * 100 + 99) for i in range(4)] + \
        [("resnet18", i, i * 100, i * 100 + 99) for i in range(4)]
    runner, done, _, _ = _run(4, q)
    _check(done, q)
    c2 = runner.coord.c2()
    assert c2["alexnet"] is not None and c2["resnet18"] is not None
