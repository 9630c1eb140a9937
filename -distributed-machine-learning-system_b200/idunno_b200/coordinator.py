"""In-box coordinator: lets the unchanged host layer drive the GPUs of one box as if they were VMs.

Mirrors, for GPU ranks, the coordinator-side bookkeeping of the reference host
(``mp4_machinelearning.py``) -- pure host logic, no sockets, no CUDA:

* ``assign``           assign_inference_work, :501-539  (fair-share worker count, random choice of
                       workers, contiguous inclusive sub-ranges, 'w' entries with timestamps)
* ``mark_finished``    update_metadata_when_get_inference_result, :623-677  ('w' -> 'f', sliding-window
                       statistics)
* ``on_worker_lost``   transfer_failed_inference_work, :706-760  (whole sub-ranges to the ring successor)
* ``result_messages``  the RESULT wire format of :603-613, split so that each message fits the single
                       ``recv(4096)`` the reference receiver does (:694) -- the reference itself truncates
                       an 80-image shard
* ``c1`` / ``c2``      query rate and processing-time statistics in the shape of the reference's c1/c2
                       commands (:1016-1036, :618-621), computed honestly (the reference prints rescaled
                       numbers)
"""
from __future__ import annotations

import random
import time
from collections import defaultdict
from statistics import mean

import numpy as np

from . import sharding

SE = "<SEPARATOR>"                 # mp4_machinelearning.py:54
BUFFER_SIZE = 4096                 # the receiver's single recv (:694)
SLIDING_WINDOW_SECONDS = 30        # the reference's 10 s x factor 3 window (:652-667)


class BoxCoordinator:
    def __init__(self, world: int, coordinator_rank: int = 0, seed: int | None = None, clock=time.time):
        self.world = world
        self.coordinator_rank = coordinator_rank
        self.alive = set(range(world))
        self.clock = clock
        self.rng = random.Random(seed)
        self.worker_set = defaultdict(list)        # (model, qnum) -> [(rank, s, e, 'w'|'f', t0, t1)]
        self.working_set = defaultdict(list)       # rank -> [(model, qnum, s, e)]
        self.avg_time = {"alexnet": 100.0, "resnet18": 100.0}     # :138 (never updated there)
        self.finished = {"alexnet": [], "resnet18": []}            # (t_finish, images, seconds per 400)

    # ---------------------------------------------------------------- assignment (:501-539)
    def assign(self, model: str, qnum, start: int, end: int, n_workers: int | None = None):
        """``n_workers`` overrides the fair-share count (a box that runs a single job uses every rank;
        the reference always halves the pool because both jobs are assumed active, :504-514)."""
        alive = sorted(self.alive)
        n = sharding.fair_share(model, self.avg_time, len(alive)) if n_workers is None else min(n_workers, len(alive))
        n = max(n, 1) if alive else 0      # with measured times the formula can round to 0: a query still needs a rank
        workers = self.rng.sample(alive, n) if n > 0 else []
        now = self.clock()
        key = (model, qnum)
        self.worker_set[key].clear()
        out = []
        for (rank, s, e) in sharding.split_range(start, end, workers):
            entry = (rank, s, e, "w", now, now)
            self.worker_set[key].append(entry)
            if (model, qnum, s, e) not in self.working_set[rank]:
                self.working_set[rank].append((model, qnum, s, e))
            out.append(entry)
        return out

    # ---------------------------------------------------------------- results (:623-677)
    def mark_finished(self, model: str, qnum, s: int, e: int):
        key = (model, qnum)
        for entry in list(self.worker_set[key]):
            rank, s_i, e_i, state, t0, _ = entry
            if str(s_i) == str(s) and str(e_i) == str(e):
                t1 = self.clock()
                self.worker_set[key].remove(entry)
                self.worker_set[key].append((rank, s_i, e_i, "f", t0, t1))
                if (model, qnum, s_i, e_i) in self.working_set[rank]:
                    self.working_set[rank].remove((model, qnum, s_i, e_i))
                images = int(e_i) - int(s_i) + 1            # the reference counts end-start (off by one, :647)
                if state == "w" and images > 0:
                    self.finished[model].append((t1, images, (t1 - t0) / images * sharding.QUERY_BATCH))
                    self._update_avg(model)
                return True
        return False

    def _window(self, model: str):
        now = self.clock()
        self.finished[model] = [x for x in self.finished[model] if abs(x[0] - now) <= SLIDING_WINDOW_SECONDS]
        return self.finished[model]

    def _update_avg(self, model: str):
        # the update the reference left commented out (:676).  The ratio only means something once BOTH
        # models have been measured; until then both stay at the reference's constant (equal shares).
        wa, wr = self._window("alexnet"), self._window("resnet18")
        if wa and wr:
            self.avg_time["alexnet"] = mean(x[2] for x in wa)
            self.avg_time["resnet18"] = mean(x[2] for x in wr)

    def query_done(self, model: str, qnum) -> bool:
        entries = self.worker_set[(model, qnum)]
        return bool(entries) and all(x[3] == "f" for x in entries)

    # ---------------------------------------------------------------- worker loss (:706-760)
    def on_worker_lost(self, rank: int):
        """Every sub-range the dead rank still works on moves whole to the ring successor; returns the
        new assignments [(model, qnum, new_rank, s, e)] for the caller to dispatch."""
        self.alive.discard(rank)
        moved = []
        for (model, qnum, s, e) in list(self.working_set.get(rank, [])):
            key = (model, qnum)
            self.worker_set[key] = [x for x in self.worker_set[key]
                                    if not (x[0] == rank and str(x[1]) == str(s) and str(x[2]) == str(e))]
            new_rank = sharding.reassign([(rank, s, e)], rank, self.alive, self.world, self.coordinator_rank)[0][0]
            now = self.clock()
            self.worker_set[key].append((new_rank, s, e, "w", now, now))
            self.working_set[new_rank].append((model, qnum, s, e))
            moved.append((model, qnum, new_rank, s, e))
        self.working_set.pop(rank, None)
        return moved

    # ---------------------------------------------------------------- wire format (:603-613, :694)
    @staticmethod
    def result_messages(sender, sdfs_file, start, end, model, qnum, inference_result, limit: int = BUFFER_SIZE):
        """RESULT messages in the reference's format.  The result list of a long shard is cut into
        consecutive pieces so that every encoded message is <= ``limit`` bytes; each piece carries the
        index range it covers (parsed from the image names), so the receiver's bookkeeping still works
        piecewise.  A shard that fits yields exactly the reference's single message."""
        def encode(s, e, part):
            return f"RESULT{SE}{sender}{SE}{sdfs_file}{SE}{s}{SE}{e}{SE}{model}{SE}{qnum}{SE}{part}"
        whole = encode(start, end, inference_result)
        if len(whole.encode()) <= limit or len(inference_result) <= 1:
            return [whole]
        out, part = [], []
        def idx(t):
            return int(t[0].split("_")[1].split(".")[0])
        for item in inference_result:
            trial = part + [item]
            if part and len(encode(idx(part[0]), idx(item), trial).encode()) > limit:
                out.append(encode(idx(part[0]), idx(part[-1]), part))
                part = [item]
            else:
                part = trial
        if part:
            out.append(encode(idx(part[0]), idx(part[-1]), part))
        return out

    # ---------------------------------------------------------------- statistics (:1016-1036, :618-621)
    def c1(self):
        """Images finished in the window per second, per model (the reference divides a 30 s count by
        model-specific constants and then prints rescaled numbers)."""
        out = {}
        for model in ("alexnet", "resnet18"):
            w = self._window(model)
            out[model] = sum(x[1] for x in w) / float(SLIDING_WINDOW_SECONDS)
        return out

    def c2(self):
        """[mean, Q1, median, Q3, std] of seconds per 400-image query in the window, per model."""
        out = {}
        for model in ("alexnet", "resnet18"):
            v = [x[2] for x in self._window(model)]
            out[model] = ([mean(v), float(np.percentile(v, 25)), float(np.percentile(v, 50)),
                           float(np.percentile(v, 75)), float(np.std(v))] if v else None)
        return out
