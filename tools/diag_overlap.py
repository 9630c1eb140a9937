"""NOTE: superseded by tools/diag_overlap_steps.py and tools/diag_e2e_timing.py.  The "overlapped" copy
time printed here includes the time the host needs to enqueue the kernel passes (the engine lets the host
run at most 4 passes ahead), so it overstates the slowdown; kept because its FIRST result -- kernels twice
as slow under a copy -- is what exposed the small per-batch cudaMemcpyAsync queued behind the image copy.

Does a pinned host->device copy slow down while the inference kernels run (and the other way round)?
Times, with CUDA events: (1) a 60 MB H2D copy alone, (2) a 400-image device-resident pass alone,
(3) both at once on different streams.  Prints one JSON object."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))
import torch  # noqa: E402

from idunno_b200 import Engine  # noqa: E402
from idunno_b200.synth import make_state_dict  # noqa: E402

dev = torch.device("cuda:0")
N = 400
IMG = 224 * 224 * 3
host = torch.randint(0, 255, (N * IMG,), dtype=torch.uint8).pin_memory()
d_in = torch.empty(N * IMG, dtype=torch.uint8, device=dev)
d_in.copy_(host)
d_copy = torch.empty_like(d_in)
offsets = np.arange(N, dtype=np.int64) * IMG
hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (N, 1))
d_idx = torch.zeros(N, dtype=torch.int32, device=dev)
d_prob = torch.zeros(N, dtype=torch.float32, device=dev)
out = {}
for model in ("alexnet", "resnet18"):
    eng = Engine(model, make_state_dict(model, seed=0), max_batch=N)
    s_k, s_c = torch.cuda.Stream(), torch.cuda.Stream()
    eng.set_stream(s_k.cuda_stream)

    def kernels(reps):
        for _ in range(reps):
            eng.infer_device(d_in.data_ptr(), offsets, hw, d_idx.data_ptr(), d_prob.data_ptr())

    def copies(reps):
        with torch.cuda.stream(s_c):
            for _ in range(reps):
                d_copy.copy_(host, non_blocking=True)

    def timed(fn_list, reps):
        torch.cuda.synchronize()
        ev = []
        for (stream, fn) in fn_list:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            ev.append((a, b, stream, fn))
        for (a, b, stream, fn) in ev:
            fn(reps)
        for (a, b, stream, fn) in ev:
            b.record(stream)
        torch.cuda.synchronize()
        return [a.elapsed_time(b) / reps for (a, b, _, _) in ev]

    kernels(3); copies(2)
    k_alone = timed([(s_k, kernels)], 20)[0]
    c_alone = timed([(s_c, copies)], 20)[0]
    # balance repetitions so that both streams stay busy for about the same time
    rk = max(1, int(round(20 * c_alone / k_alone)))
    both = timed([(s_k, lambda r: kernels(rk)), (s_c, lambda r: copies(20))], 1)
    out[model] = {"kernels_alone_ms": round(k_alone, 4), "copy_alone_ms": round(c_alone, 4),
                  "copy_alone_GBps": round(N * IMG / c_alone / 1e6, 1),
                  "kernels_overlapped_ms": round(both[0] / rk, 4), "copy_overlapped_ms": round(both[1] / 20, 4),
                  "copy_overlapped_GBps": round(N * IMG / (both[1] / 20) / 1e6, 1)}
    # the same copy split over several streams (several copy engines may then work on it)
    for parts in (2, 4, 8):
        streams = [torch.cuda.Stream() for _ in range(parts)]
        chunk = (N * IMG + parts - 1) // parts

        def split_copies(reps, streams=streams, chunk=chunk):
            for _ in range(reps):
                for j, st in enumerate(streams):
                    with torch.cuda.stream(st):
                        d_copy[j * chunk:(j + 1) * chunk].copy_(host[j * chunk:(j + 1) * chunk], non_blocking=True)

        def split_timed(with_kernels):
            torch.cuda.synchronize()
            a = torch.cuda.Event(enable_timing=True)
            ends = [torch.cuda.Event(enable_timing=True) for _ in streams]
            a.record(torch.cuda.current_stream())
            for st in streams:
                st.wait_event(a)
            if with_kernels:
                kernels(rk)
            split_copies(20)
            for e, st in zip(ends, streams):
                e.record(st)
            torch.cuda.synchronize()
            return max(a.elapsed_time(e) for e in ends) / 20
        split_copies(2)
        out[model][f"copy_{parts}streams_alone_ms"] = round(split_timed(False), 4)
        out[model][f"copy_{parts}streams_overlapped_ms"] = round(split_timed(True), 4)
    eng.close()
print(json.dumps(out))
