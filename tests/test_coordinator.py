"""CPU: the in-box coordinator shim (assignment, RESULT wire format, worker loss, statistics)."""
import pytest

from idunno_b200 import sharding
from idunno_b200.coordinator import BUFFER_SIZE, SE, BoxCoordinator


class FakeClock:
    def __init__(self):
        self.t = 1000.0

    def __call__(self):
        return self.t


def test_assign_covers_range_with_fair_share_of_workers():
    clk = FakeClock()
    c = BoxCoordinator(world=8, seed=1, clock=clk)
    entries = c.assign("resnet18", 7, 0, 399)
    assert len(entries) == 5                                    # constant 100/100 times -> 5 workers (:509-514)
    assert sorted((s, e) for _, s, e, *_ in entries) == [(80 * i, 80 * i + 79) for i in range(5)]
    assert all(st == "w" for *_, st, _, _ in [(r, s, e, st, a, b) for (r, s, e, st, a, b) in entries])
    assert not c.query_done("resnet18", 7)
    clk.t += 2.0
    for (r, s, e, *_rest) in entries:
        assert c.mark_finished("resnet18", 7, str(s), str(e))   # the host passes strings
    assert c.query_done("resnet18", 7)
    assert not c.mark_finished("resnet18", 7, 1, 2)
    # honest statistics: 400 images in the 30 s window, 2 s per 80 images -> 10 s per 400
    assert c.c1()["resnet18"] == pytest.approx(400 / 30.0)
    assert c.c2()["resnet18"][0] == pytest.approx(10.0)
    assert c.c2()["alexnet"] is None
    # one model measured only: the ratio is not meaningful yet, both stay at the reference's constant
    assert c.avg_time == {"alexnet": 100.0, "resnet18": 100.0}
    # once both are measured the times drive the fair share (the reference leaves them at 100/100):
    # the slower model gets more ranks
    ea = c.assign("alexnet", 8, 0, 399)
    clk.t += 0.5
    for (r, s, e, *_rest) in ea:
        assert c.mark_finished("alexnet", 8, s, e)
    assert c.avg_time["resnet18"] == pytest.approx(10.0) and c.avg_time["alexnet"] == pytest.approx(2.5)
    assert sharding.fair_share("resnet18", c.avg_time, 8) > sharding.fair_share("alexnet", c.avg_time, 8) >= 1


def test_worker_loss_moves_whole_subranges():
    c = BoxCoordinator(world=4, seed=3, clock=FakeClock())
    c.avg_time = {"alexnet": 1.0, "resnet18": 1.0}
    entries = c.assign("alexnet", 1, 0, 399)
    victim = next(r for (r, *_rest) in entries if r != 0)
    held = [(s, e) for (r, s, e, *_rest) in entries if r == victim]
    moved = c.on_worker_lost(victim)
    assert [(s, e) for (_, _, _, s, e) in moved] == held
    for (_, _, new_rank, s, e) in moved:
        assert new_rank in c.alive and new_rank != victim and new_rank != 0
        assert (new_rank, s, e) in [(r, a, b) for (r, a, b, *_x) in c.worker_set[("alexnet", 1)]]
    assert victim not in c.working_set
    for (r, s, e, *_rest) in list(c.worker_set[("alexnet", 1)]):
        c.mark_finished("alexnet", 1, s, e)
    assert c.query_done("alexnet", 1)


def test_result_messages_fit_the_single_recv():
    res = [(f"test_{i}.JPEG", "great white shark", 0.123456789) for i in range(300, 380)]
    one = BoxCoordinator.result_messages("gpu3", "resnet18", 300, 304, "resnet18", 9, res[:5])
    assert len(one) == 1 and one[0].split(SE)[:7] == ["RESULT", "gpu3", "resnet18", "300", "304", "resnet18", "9"]
    assert one[0].split(SE)[7] == str(res[:5])
    msgs = BoxCoordinator.result_messages("gpu3", "resnet18", 300, 379, "resnet18", 9, res)
    assert len(msgs) > 1 and all(len(m.encode()) <= BUFFER_SIZE for m in msgs)
    covered = []
    for m in msgs:
        f = m.split(SE)
        part = eval(f[7])                                      # the host stores the string; tests may parse it
        assert [int(t[0].split("_")[1].split(".")[0]) for t in part] == list(range(int(f[3]), int(f[4]) + 1))
        covered += part
    assert covered == res
