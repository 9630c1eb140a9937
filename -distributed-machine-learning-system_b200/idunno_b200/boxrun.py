"""Range runner for the GPUs of one box that survives the loss of a worker process.

The reference detects a dead VM by missed pings (2 s timeout, mp4_machinelearning.py:832-884) and then
re-sends every sub-range that VM was working on to its ring successor
(transfer_failed_inference_work, :706-760); results arrive as RESULT messages over fresh TCP
connections (:603-613, :679-704), so a dead peer never blocks the others.  An NCCL communicator does
not have that property (a collective with a dead rank hangs), so inside a box the *tiny* result path
goes through a key-value rendezvous (``torch.distributed.TCPStore``: plain TCP, no collectives, a
vanished client is simply a closed socket) while the GPUs never talk to each other at all:

    task/<rank>/<seq>            coordinator -> worker   (model, qnum, start, end) or STOP
    res/<model>/<qnum>/<s>/<e>   worker -> coordinator   int64 indices | int32 top-1 | fp32 prob
    hb/<rank>                    worker heartbeat counter, bumped every ``HEARTBEAT_S``

The coordinator is rank 0 and, like the reference's (:762-807), classifies a shard of its own.  All
bookkeeping (who works on what, 'w'/'f' states, statistics, the ring-successor rule) is
``coordinator.BoxCoordinator``; this file only moves bytes and watches heartbeats.  Results merge
idempotently: a shard that is delivered twice (the "dead" worker was merely slow) is harmless.

``classify(model, start, end) -> (indices, top1_idx, top1_prob)`` is supplied by the caller: the
product passes a function that drives ``Engine`` on this rank's GPU (see ``gpu_classifier``); the CPU
tests pass a deterministic function of the image index, so the host logic is testable without a GPU.
"""
from __future__ import annotations

import json
import threading
import time
from datetime import timedelta

import numpy as np

from .coordinator import BoxCoordinator
from .dist import merge_chunks

HEARTBEAT_S = 0.25
LOSS_TIMEOUT_S = 2.0          # the reference's ping timeout (mp4_machinelearning.py:847)
POLL_S = 0.001
STOP = "STOP"


def open_store(host: str, port: int, is_master: bool, timeout_s: float = 60.0):
    from torch.distributed import TCPStore
    return TCPStore(host, port, None, is_master, timedelta(seconds=timeout_s), wait_for_workers=False)


def _pack(indices, idx, prob) -> bytes:
    indices = np.ascontiguousarray(indices, dtype=np.int64)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    prob = np.ascontiguousarray(prob, dtype=np.float32)
    assert len(indices) == len(idx) == len(prob)
    return np.int64(len(indices)).tobytes() + indices.tobytes() + idx.tobytes() + prob.tobytes()


def _unpack(b: bytes):
    n = int(np.frombuffer(b[:8], dtype=np.int64)[0])
    o = 8
    indices = np.frombuffer(b[o:o + 8 * n], dtype=np.int64); o += 8 * n
    idx = np.frombuffer(b[o:o + 4 * n], dtype=np.int32); o += 4 * n
    prob = np.frombuffer(b[o:o + 4 * n], dtype=np.float32)
    return indices, idx, prob


def _res_key(model, qnum, s, e) -> str:
    return f"res/{model}/{qnum}/{s}/{e}"


class _Heartbeat(threading.Thread):
    """Bumps hb/<rank>.  GPU work happens inside ctypes calls, which release the GIL, so the beat
    keeps its pace while a shard is being classified -- the same reason the reference's ping
    thread keeps running during deeplearning()."""

    def __init__(self, store, rank: int):
        super().__init__(daemon=True)
        self.store, self.rank, self.stop_flag = store, rank, threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                self.store.add(f"hb/{self.rank}", 1)
            except Exception:
                return
            self.stop_flag.wait(HEARTBEAT_S)


def worker_loop(host: str, port: int, rank: int, classify, die_after_tasks: int | None = None,
                die_in_task=None):
    """A GPU worker: takes (model, qnum, start, end) tasks in order, classifies, posts the result.
    ``die_after_tasks`` / ``die_in_task`` exist for the loss tests and the kill scenario of BASELINE
    config 5: the process leaves with os._exit -- no goodbye, exactly what a crashed worker looks like
    -- either before starting task number ``die_after_tasks`` or from inside classify (``die_in_task``
    is passed through to it as a callback the classifier calls between tensor batches)."""
    import os
    store = open_store(host, port, False)
    beat = _Heartbeat(open_store(host, port, False), rank)     # own connection: no shared socket
    beat.start()
    seq = 0
    try:
        while True:
            key = f"task/{rank}/{seq}"
            while not store.check([key]):
                time.sleep(POLL_S)
            task = json.loads(store.get(key).decode())
            if task == STOP:
                return
            if die_after_tasks is not None and seq >= die_after_tasks:
                os._exit(17)
            model, qnum, s, e = task
            if die_in_task is not None:
                out = classify(model, s, e, die_in_task)
            else:
                out = classify(model, s, e)
            store.set(_res_key(model, qnum, s, e), _pack(*out))
            seq += 1
    finally:
        beat.stop_flag.set()


class BoxRunner:
    """Coordinator side.  One instance per box, on rank 0."""

    def __init__(self, host: str, port: int, world: int, classify, seed: int | None = 0,
                 loss_timeout_s: float = LOSS_TIMEOUT_S):
        self.store = open_store(host, port, True)
        self.world = world
        self.classify = classify
        self.coord = BoxCoordinator(world, coordinator_rank=0, seed=seed)
        self.loss_timeout_s = loss_timeout_s
        self.seq = {r: 0 for r in range(world)}
        self.beat_seen = {}                    # rank -> (counter, local time it last changed)
        self.events = []                       # ("lost", rank, t) / ("moved", model, qnum, new_rank, s, e, t)
        self._local = []                       # shards the coordinator classifies itself
        self._lock = threading.Lock()

    # ---------------------------------------------------------------- liveness
    def wait_for_workers(self, timeout_s: float = 120.0):
        """Blocks until every worker's heartbeat has been seen once (process start + CUDA context +
        weight upload can take a while; the loss timeout only applies from then on)."""
        t_end = time.monotonic() + timeout_s
        for r in range(1, self.world):
            while self.store.add(f"hb/{r}", 0) == 0:
                if time.monotonic() > t_end:
                    raise TimeoutError(f"worker {r} never started")
                time.sleep(POLL_S)
            self.beat_seen[r] = (self.store.add(f"hb/{r}", 0), time.monotonic())

    def _dead_ranks(self):
        now = time.monotonic()
        dead = []
        for r in sorted(self.coord.alive):
            if r == 0:
                continue
            c = self.store.add(f"hb/{r}", 0)
            last_c, last_t = self.beat_seen.get(r, (0, now))
            if c != last_c:
                self.beat_seen[r] = (c, now)
            elif now - last_t > self.loss_timeout_s:
                dead.append(r)
        return dead

    # ---------------------------------------------------------------- dispatch
    def _dispatch(self, rank: int, model: str, qnum, s: int, e: int):
        if rank == 0:
            with self._lock:
                self._local.append((model, qnum, s, e))
        else:
            self.store.set(f"task/{rank}/{self.seq[rank]}", json.dumps([model, qnum, s, e]))
            self.seq[rank] += 1

    def _local_worker(self, stop: threading.Event, results: dict, errors: list):
        while not stop.is_set():
            with self._lock:
                item = self._local.pop(0) if self._local else None
            if item is None:
                time.sleep(POLL_S)
                continue
            model, qnum, s, e = item
            try:
                results[(model, qnum, s, e)] = self.classify(model, s, e)
            except Exception as ex:              # surfaces in run_queries, not in a dead thread
                errors.append(ex)
                return

    # ---------------------------------------------------------------- the query loop
    def run_queries(self, queries, max_in_flight: int = 2, n_workers=None, on_tick=None):
        """``queries`` = [(model, qnum, start, end)] (inclusive ranges, the reference's 400-image
        queries or anything else).  At most ``max_in_flight`` queries per model are outstanding;
        each is split over the alive ranks by ``BoxCoordinator.assign`` (fair share between the two
        models, the reference's rounding rule).  Returns {(model, qnum): (indices, top1, prob)} with
        indices ascending.  ``on_tick(runner)`` is called once per monitor iteration (the kill
        scenario uses it to shoot a worker at a chosen moment)."""
        todo = list(queries)
        open_q = {}                                   # (model, qnum) -> [(indices, top1, prob)] in arrival order
        done = {}
        local_results, errors = {}, []
        stop = threading.Event()
        th = threading.Thread(target=self._local_worker, args=(stop, local_results, errors), daemon=True)
        th.start()
        try:
            while todo or open_q:
                if errors:
                    raise errors[0]
                # admit
                for q in list(todo):
                    model, qnum, s, e = q
                    if sum(1 for k in open_q if k[0] == model) >= max_in_flight:
                        continue
                    todo.remove(q)
                    open_q[(model, qnum)] = []
                    kw = {} if n_workers is None else {"n_workers": n_workers}
                    for (rank, ss, ee, _, _, _) in self.coord.assign(model, qnum, s, e, **kw):
                        self._dispatch(rank, model, qnum, ss, ee)
                # harvest
                for key in list(open_q):
                    model, qnum = key
                    for (rank, ss, ee, state, _, _) in list(self.coord.worker_set[key]):
                        if state != "w":
                            continue
                        got = None
                        if (model, qnum, ss, ee) in local_results:
                            got = local_results.pop((model, qnum, ss, ee))
                        elif self.store.check([_res_key(model, qnum, ss, ee)]):
                            got = _unpack(self.store.get(_res_key(model, qnum, ss, ee)))
                        if got is not None:
                            open_q[key].append(got)     # idempotent: a shard delivered twice merges to one
                            self.coord.mark_finished(model, qnum, ss, ee)
                    if self.coord.query_done(model, qnum):
                        done[key] = merge_chunks(open_q.pop(key))
                # liveness
                for r in self._dead_ranks():
                    t = time.monotonic()
                    self.events.append(("lost", r, t))
                    for (model, qnum, new_rank, ss, ee) in self.coord.on_worker_lost(r):
                        self.events.append(("moved", model, qnum, new_rank, ss, ee, t))
                        self._dispatch(new_rank, model, qnum, ss, ee)
                if on_tick is not None:
                    on_tick(self)
                time.sleep(POLL_S)
        finally:
            stop.set()
            th.join(timeout=5)
        return done

    def shutdown(self):
        for r in range(1, self.world):
            try:
                self.store.set(f"task/{r}/{self.seq[r]}", json.dumps(STOP))
            except Exception:
                pass


# ---------------------------------------------------------------------------------------------
# The product classifier: Engine on this rank's GPU, images of the synthetic benchmark set
# ---------------------------------------------------------------------------------------------
SET_PERIOD = 400        # the benchmark set repeats with this period: image i == synth_image(i % 400)


class SyntheticImageSet:
    """Stands in for the image folder when there is no dataset (BASELINE.json: synthetic
    ImageNet-shaped uint8 images).  Image ``i`` is ``synth.synth_image(i % SET_PERIOD)``; the set is
    laid out period after period in ONE pinned buffer long enough that any ``n <= max_batch``
    consecutive indices are a contiguous slice of it -- a tensor batch is staged without a copy."""

    def __init__(self, max_batch: int, device: int, hw=(224, 224)):
        from .engine import PinnedBuffer
        from .synth import synth_image
        self.h, self.w = hw
        self.img_bytes = self.h * self.w * 3
        periods = (max_batch + SET_PERIOD - 1) // SET_PERIOD + 1
        self.buf = PinnedBuffer(periods * SET_PERIOD * self.img_bytes, device)
        import os
        cache = os.environ.get("IDN_SYNTH_BLOCK")          # generating 400 images takes seconds: share it
        if cache and os.path.exists(cache):
            block = np.load(cache)
        else:
            block = np.concatenate([synth_image(i, (self.h, self.w)).reshape(-1) for i in range(SET_PERIOD)])
            if cache:
                np.save(cache + f".{os.getpid()}.npy", block)
                os.replace(cache + f".{os.getpid()}.npy", cache)
        assert block.size == SET_PERIOD * self.img_bytes
        view = self.buf.array.reshape(periods, SET_PERIOD * self.img_bytes)
        view[:] = block[None, :]
        self.max_batch = max_batch

    def window(self, first_index: int, n: int):
        """-> (host pointer, offsets[n], hw[n,2]) for images first_index .. first_index+n-1."""
        assert 0 < n <= self.max_batch
        base = (first_index % SET_PERIOD) * self.img_bytes
        offsets = np.arange(n, dtype=np.int64) * self.img_bytes
        hw = np.tile(np.asarray([[self.h, self.w]], dtype=np.int32), (n, 1))
        return self.buf.ptr + base, offsets, hw


def gpu_classifier(device: int, max_batch: int = 400, seed: int = 0, variant: str = "default"):
    """classify(model, start, end[, between_batches]) over the synthetic set, tensor batches of
    ``max_batch`` images, up to three batches in flight (copies of k+1, k+2 under the kernels of k).  Engines are
    built on first use per model and stay resident, as in the drop-in ``deeplearning``."""
    from .engine import Engine, MAX_IN_FLIGHT
    from .synth import make_state_dict
    images = SyntheticImageSet(max_batch, device)
    engines = {}

    def engine(model):
        if model not in engines:
            engines[model] = Engine(model, make_state_dict(model, seed=seed, variant=variant), device=device,
                                    max_batch=max_batch)
        return engines[model]

    def classify(model, start, end, between_batches=None):
        eng = engine(model)
        start, end = int(start), int(end)
        total = end - start + 1
        idx = np.zeros(total, dtype=np.int32)
        prob = np.zeros(total, dtype=np.float32)
        pending = []                                        # (ticket, position)
        pos, batches = 0, 0
        while pos < total or pending:
            while pos < total and len(pending) < MAX_IN_FLIGHT:
                n = min(max_batch, total - pos)
                ptr, offsets, hw = images.window(start + pos, n)
                pending.append((eng.submit_packed(offsets, hw, ptr), pos, n))
                pos += n
            ticket, at, n = pending.pop(0)
            i, p, _ = eng.collect(ticket)
            idx[at:at + n], prob[at:at + n] = i, p
            batches += 1
            if between_batches is not None:
                between_batches(batches)
        return np.arange(start, end + 1, dtype=np.int64), idx, prob

    classify.engine = engine
    return classify


def _exit_after(n_batches: int):
    def cb(done: int):
        if done >= n_batches:
            import os
            os._exit(17)
    return cb


def gpu_worker_main(host: str, port: int, rank: int, max_batch: int, die_after_batches: int | None = None,
                    preload=("resnet18",), device: int | None = None):
    """Entry point of one GPU worker process (CUDA device index = rank unless ``device`` says otherwise:
    the single-GPU test runs two ranks on one device)."""
    from .engine import bind_host_thread_to_gpu
    bind_host_thread_to_gpu(rank if device is None else device)     # staging memory on the GPU's NUMA node
    classify = gpu_classifier(rank if device is None else device, max_batch)
    for m in preload:                      # weights resident before the first heartbeat is sent
        classify.engine(m)
    worker_loop(host, port, rank, classify,
                die_in_task=_exit_after(die_after_batches) if die_after_batches else None)
