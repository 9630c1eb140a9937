"""GPU: the range runner with real GPU workers.  Two ranks share cuda:0 (the suite runs on one GPU); the
second one dies without a goodbye in the middle of its shard and the coordinator re-runs its sub-range.
Results must be identical to classifying the whole range in one process."""
import multiprocessing as mp
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")
sys.path.insert(0, PKG)

pytestmark = pytest.mark.gpu


def test_synthetic_set_windows_match_explicit_images(built_library):
    from idunno_b200 import Engine, boxrun
    from idunno_b200.synth import make_state_dict, synth_image
    classify = boxrun.gpu_classifier(0, max_batch=48)
    ind, idx, prob = classify("alexnet", 390, 520)            # crosses the 400-image period, ragged tail
    assert np.array_equal(ind, np.arange(390, 521))
    with Engine("alexnet", make_state_dict("alexnet", seed=0), max_batch=48) as eng:
        want_idx, want_prob, _ = eng.infer([synth_image(i % 400, (224, 224)) for i in range(390, 521)])
    assert np.array_equal(idx, want_idx) and np.array_equal(prob, want_prob)


@pytest.mark.parametrize("die_after", [None, 2])
def test_box_of_two_ranks_on_one_gpu(built_library, die_after):
    from idunno_b200 import boxrun
    os.environ["PYTHONPATH"] = PKG + os.pathsep + os.environ.get("PYTHONPATH", "")
    port = 25000 + os.getpid() % 3000 + (7 if die_after else 0)
    batch = 32
    classify = boxrun.gpu_classifier(0, batch)
    total = batch * 12
    _, want_idx, want_prob = classify("resnet18", 100, 100 + total - 1)
    runner = boxrun.BoxRunner("127.0.0.1", port, 2, classify, seed=0, loss_timeout_s=1.0)
    p = mp.get_context("spawn").Process(target=boxrun.gpu_worker_main,
                                        args=("127.0.0.1", port, 1, batch, die_after, ("resnet18",), 0))
    p.start()
    try:
        runner.wait_for_workers(300)
        done = runner.run_queries([("resnet18", 3, 100, 100 + total - 1)], n_workers=2)
    finally:
        runner.shutdown()
        p.join(timeout=30)
        if p.is_alive():
            p.kill()
    ind, idx, prob = done[("resnet18", 3)]
    assert np.array_equal(ind, np.arange(100, 100 + total))
    assert np.array_equal(idx, want_idx) and np.array_equal(prob, want_prob)
    if die_after:
        assert p.exitcode == 17
        assert [e[1] for e in runner.events if e[0] == "lost"] == [1]
        moved = [e for e in runner.events if e[0] == "moved"]
        assert len(moved) == 1 and moved[0][3] == 0       # only the coordinator is left: it takes the shard
    else:
        assert p.exitcode == 0 and not runner.events


def test_staging_pipeline_matches_single_call(built_library, tmp_path, monkeypatch):
    """classify_range (threads -> pinned ring -> submit/collect, several chunks in flight) against one
    blocking call over the same decoded images; mixed sizes, a missing file, a ragged last chunk."""
    from PIL import Image
    from idunno_b200 import Engine, staging
    from idunno_b200.synth import make_state_dict, synth_image
    monkeypatch.chdir(tmp_path)
    os.makedirs("resnet18")
    for i in range(20, 43):
        if i != 31:
            Image.fromarray(synth_image(i, mixed=True)).save(f"resnet18/test_{i}.JPEG", format="PNG")
    with Engine("resnet18", make_state_dict("resnet18", seed=0), max_batch=5) as eng:
        ind, idx, prob = staging.classify_range(eng, "resnet18", "20", 45, threads=4)
        indices, images = staging.load_range("resnet18", 20, 45)
        want_idx, want_prob, _ = eng.infer(images)
        again = staging.classify_range(eng, "resnet18", 20, 45)               # pinned ring is reused
    assert list(ind) == indices == [i for i in range(20, 43) if i != 31]
    assert np.array_equal(idx, want_idx) and np.array_equal(prob, want_prob)
    assert np.array_equal(again[1], idx) and np.array_equal(again[2], prob)


def test_concurrent_jobs_c1_c2_and_measured_fair_share(built_library):
    """SURVEY 8f-3 on the GPU: AlexNet and ResNet18 query streams run concurrently through the box runner
    (two ranks on cuda:0); the coordinator's c1 (images per second in the 30 s window) and c2 (mean /
    quartiles / sigma of seconds per 400-image query) come from MEASURED shard times, the averages the
    reference leaves at the constant 100 (mp4_machinelearning.py:138,676) are updated from them, and the
    worker count of later queries is the fair-share formula (:504-514) evaluated on those averages."""
    from idunno_b200 import boxrun, sharding
    os.environ["PYTHONPATH"] = PKG + os.pathsep + os.environ.get("PYTHONPATH", "")
    port = 25000 + os.getpid() % 3000 + 23
    batch = 40
    classify = boxrun.gpu_classifier(0, batch)
    span = batch * 4
    want = {m: classify(m, 0, span - 1) for m in ("alexnet", "resnet18")}
    runner = boxrun.BoxRunner("127.0.0.1", port, 2, classify, seed=0)
    p = mp.get_context("spawn").Process(target=boxrun.gpu_worker_main,
                                        args=("127.0.0.1", port, 1, batch, None, ("alexnet", "resnet18"), 0))
    p.start()
    try:
        runner.wait_for_workers(300)
        queries = [(m, q, 0, span - 1) for q in range(6) for m in ("alexnet", "resnet18")]
        done = runner.run_queries(queries)                       # fair share decides the worker counts
        coord = runner.coord
        c1, c2, avg = coord.c1(), coord.c2(), dict(coord.avg_time)
        last_counts = {m: len(coord.worker_set[(m, 5)]) for m in ("alexnet", "resnet18")}
    finally:
        runner.shutdown()
        p.join(timeout=30)
        if p.is_alive():
            p.kill()
    assert len(done) == 12
    for (m, q), (ind, idx, prob) in done.items():
        assert np.array_equal(ind, np.arange(span))
        assert np.array_equal(idx, want[m][1]) and np.array_equal(prob, want[m][2])
    for m in ("alexnet", "resnet18"):
        # every image of the six queries finished inside the window: c1 is exactly their count / 30 s
        assert c1[m] == pytest.approx(6 * span / 30.0)
        mean_s, q1, q2, q3, sd = c2[m]
        assert 0 < q1 <= q2 <= q3 and mean_s > 0 and sd >= 0
        assert avg[m] != 100.0 and avg[m] == pytest.approx(mean_s, rel=0.5)   # measured, no longer the constant
        # the last query of each model was split over the fair share computed from the measured averages
        assert 1 <= last_counts[m] <= 2
    print(f"\n[boxrun] c1 = {c1}  c2 mean s/400 = { {m: round(c2[m][0], 4) for m in c2} }  fair share on 2 ranks = "
          f"{ {m: sharding.fair_share(m, avg, 2) for m in avg} }  workers of the last queries = {last_counts}")
    assert p.exitcode == 0
