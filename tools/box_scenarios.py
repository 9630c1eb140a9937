"""BASELINE.json configs 3-5 on the GPUs of one box, through idunno_b200.boxrun (one process per GPU,
rank 0 = coordinator + worker).  Prints one JSON object per scenario and writes them to
gpurun_out/box_scenarios.json.

  sharded     ResNet18, one query over N*10000 images, range-split by the reference's rule (config 3)
  concurrent  AlexNet + ResNet18 query streams at the same time, fair-share rank counts from measured
              per-query times, c1/c2 statistics (config 4)
  killsweep   ResNet18 tensor-batch sweep 64..4096, one worker process dies mid-shard, its sub-range is
              re-run by its ring successor; the merged result must equal the no-failure result (config 5)

    python tools/box_scenarios.py --gpus 2 [--scenario all] [--batches 64,256,1024,4096]
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")
sys.path.insert(0, PKG)
os.environ["PYTHONPATH"] = PKG + os.pathsep + os.environ.get("PYTHONPATH", "")
os.environ.setdefault("IDN_SYNTH_BLOCK", "/tmp/idn_synth_block_224.npy")

import numpy as np  # noqa: E402


def start_box(world, max_batch, port, preload, victim=None, die_after_batches=None, loss_timeout_s=2.0):
    from idunno_b200 import boxrun
    ctx = mp.get_context("spawn")
    classify = boxrun.gpu_classifier(0, max_batch)
    for m in preload:
        classify.engine(m)
    runner = boxrun.BoxRunner("127.0.0.1", port, world, classify, seed=0, loss_timeout_s=loss_timeout_s)
    procs = []
    for r in range(1, world):
        p = ctx.Process(target=boxrun.gpu_worker_main,
                        args=("127.0.0.1", port, r, max_batch, die_after_batches if r == victim else None, preload))
        p.start()
        procs.append(p)
    runner.wait_for_workers(300)
    return runner, procs, classify


def stop_box(runner, procs):
    runner.shutdown()
    for p in procs:
        p.join(timeout=20)
        if p.is_alive():
            p.kill()                       # the exact process this script started


def scenario_sharded(world, port, per_rank=10000):
    runner, procs, _ = start_box(world, 400, port, ("resnet18",))
    try:
        total = per_rank * world
        runner.run_queries([("resnet18", "warm", 0, 400 * world - 1)], n_workers=world)
        t0 = time.perf_counter()
        done = runner.run_queries([("resnet18", 0, 0, total - 1)], n_workers=world)
        dt = time.perf_counter() - t0
        ind, idx, prob = done[("resnet18", 0)]
        assert len(ind) == total and np.array_equal(ind, np.arange(total))
        # the set repeats every 400 images: every period must classify identically
        assert np.array_equal(idx.reshape(-1, 400), np.tile(idx[:400], (total // 400, 1)))
        assert np.array_equal(prob.reshape(-1, 400), np.tile(prob[:400], (total // 400, 1)))
        return {"scenario": "sharded", "model": "resnet18", "n_gpus": world, "images": total,
                "tensor_batch": 400, "seconds": round(dt, 4), "images_per_s": round(total / dt, 1),
                "timing": "host wall clock on the coordinator: dispatch + H2D + kernels + result return",
                "periodic_results_identical": True}
    finally:
        stop_box(runner, procs)


def scenario_concurrent(world, port, queries_per_model=40):
    runner, procs, _ = start_box(world, 400, port, ("alexnet", "resnet18"))
    try:
        per_query = 400 * world
        warm = [(m, f"warm{i}", 0, per_query - 1) for m in ("alexnet", "resnet18") for i in range(2)]
        runner.run_queries(warm)
        q = []
        for i in range(queries_per_model):                 # interleaved, as two clients typing at once
            q.append(("alexnet", i, i * per_query, (i + 1) * per_query - 1))
            q.append(("resnet18", i, i * per_query, (i + 1) * per_query - 1))
        shares = []
        t0 = time.perf_counter()
        done = runner.run_queries(q, on_tick=lambda r: shares.append(
            (r.coord.avg_time["alexnet"], r.coord.avg_time["resnet18"])) if len(shares) < 100000 else None)
        dt = time.perf_counter() - t0
        from idunno_b200 import sharding
        avg = dict(runner.coord.avg_time)
        ref = {m: done[(m, 0)] for m in ("alexnet", "resnet18")}
        for (m, i, s, e) in q:                              # every query covers whole periods of the set
            ind, idx, prob = done[(m, i)]
            assert np.array_equal(ind, np.arange(s, e + 1))
            assert np.array_equal(idx, ref[m][1]) and np.array_equal(prob, ref[m][2])
        return {"scenario": "concurrent", "n_gpus": world, "queries_per_model": queries_per_model,
                "images_per_query": per_query, "seconds": round(dt, 4),
                "images_per_s": {"both": round(2 * queries_per_model * per_query / dt, 1)},
                "avg_seconds_per_400": {k: round(v, 6) for k, v in avg.items()},
                "fair_share_ranks": {m: max(1, sharding.fair_share(m, avg, world)) for m in ("alexnet", "resnet18")},
                "c1_images_per_s_in_window": runner.coord.c1(),
                "c2_seconds_per_400[mean,q1,median,q3,std]": runner.coord.c2(),
                "all_queries_identical_to_first": True}
    finally:
        stop_box(runner, procs)


def scenario_kill(world, port, batch, batches_per_rank=8, die_after_batches=3):
    victim = world - 1 if world > 1 else None
    runner, procs, classify = start_box(world, batch, port, ("resnet18",), victim=victim,
                                        die_after_batches=die_after_batches)
    try:
        total = batch * batches_per_rank * world
        _, want_idx, want_prob = classify("resnet18", 0, total - 1)          # no-failure result, one GPU
        t0 = time.perf_counter()
        done = runner.run_queries([("resnet18", batch, 0, total - 1)], n_workers=world)
        dt = time.perf_counter() - t0
        ind, idx, prob = done[("resnet18", batch)]
        same = bool(np.array_equal(ind, np.arange(total)) and np.array_equal(idx, want_idx)
                    and np.array_equal(prob, want_prob))
        lost = [e for e in runner.events if e[0] == "lost"]
        moved = [e for e in runner.events if e[0] == "moved"]
        out = {"scenario": "killsweep", "model": "resnet18", "n_gpus": world, "tensor_batch": batch,
               "images": total, "victim_rank": victim, "victim_exit_code": None,
               "lost": [e[1] for e in lost],
               "moved": [{"to_rank": e[3], "start": e[4], "end": e[5]} for e in moved],
               "seconds_with_loss": round(dt, 4), "loss_timeout_s": runner.loss_timeout_s,
               "identical_to_no_failure_run": same}
        if not same:
            raise AssertionError(f"result after reassignment differs: {out}")
        return out, procs, victim
    finally:
        stop_box(runner, procs)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--scenario", default="all")
    ap.add_argument("--batches", default="64,128,256,512,1024,2048,4096")
    ap.add_argument("--port", type=int, default=24500)
    ap.add_argument("--per-rank", type=int, default=10000,
                    help="images per rank of the sharded scenario (config 3 is 10000; 100000 gives a >= 0.5 s run)")
    a = ap.parse_args()
    results, port = [], a.port
    if a.scenario in ("all", "sharded"):
        results.append(scenario_sharded(a.gpus, port, a.per_rank)); port += 1
        print(json.dumps(results[-1]), flush=True)
    if a.scenario in ("all", "concurrent"):
        results.append(scenario_concurrent(a.gpus, port)); port += 1
        print(json.dumps(results[-1]), flush=True)
    if a.scenario in ("all", "killsweep"):
        for b in [int(x) for x in a.batches.split(",")]:
            out, procs, victim = scenario_kill(a.gpus, port, b); port += 1
            if victim is not None:
                out["victim_exit_code"] = procs[victim - 1].exitcode
            results.append(out)
            print(json.dumps(out), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"box_scenarios_{a.gpus}gpu.json"), "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
