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
