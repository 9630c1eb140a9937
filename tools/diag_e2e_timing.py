"""Per-step kernel durations while the host-buffer pipeline (submit/collect, 3 in flight) is running,
next to the same durations with device-resident inputs: shows whether the H2D copies slow the kernels
or whether time is lost between passes."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "-distributed-machine-learning-system_b200"))
import torch  # noqa: E402

from idunno_b200 import Engine  # noqa: E402
from idunno_b200.synth import make_state_dict  # noqa: E402

dev = torch.device("cuda:0")
N, IMG = 400, 224 * 224 * 3
hosts = [torch.randint(0, 255, (N * IMG,), dtype=torch.uint8).pin_memory() for _ in range(4)]
d_in = torch.empty(N * IMG, dtype=torch.uint8, device=dev); d_in.copy_(hosts[0])
offsets = np.arange(N, dtype=np.int64) * IMG
hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (N, 1))
d_idx = torch.zeros(N, dtype=torch.int32, device=dev); d_prob = torch.zeros(N, dtype=torch.float32, device=dev)
out = {}
for model in ("alexnet", "resnet18"):
    eng = Engine(model, make_state_dict(model, seed=0), max_batch=N)
    s_k = torch.cuda.Stream(); eng.set_stream(s_k.cuda_stream)
    eng.set_timing(1)
    for _ in range(5):
        eng.infer_device(d_in.data_ptr(), offsets, hw, d_idx.data_ptr(), d_prob.data_ptr())
    eng.sync()
    alone = dict(eng.step_times_ms())

    def pipeline(k, depth):
        pend = []
        for i in range(k):
            if len(pend) == depth:
                eng.collect(pend.pop(0))
            pend.append(eng.submit_packed(offsets, hw, hosts[i % 4].data_ptr()))
        while pend:
            eng.collect(pend.pop(0))
    res = {"alone_sum_ms": round(sum(alone.values()), 4)}
    for depth in (1, 2, 3):
        pipeline(6, depth)
        t0 = time.perf_counter(); pipeline(40, depth); dt = (time.perf_counter() - t0) / 40 * 1e3
        piped = dict(eng.step_times_ms())           # the last pass of the pipeline (copies were still running for depth>1 until just before)
        res[f"depth{depth}"] = {"ms_per_step_wall": round(dt, 4), "last_pass_kernel_sum_ms": round(sum(piped.values()), 4)}
    # kernels timed in the middle of the pipeline: stop submitting after 20 and read the pass that overlapped a copy
    pend = [eng.submit_packed(offsets, hw, hosts[i % 4].data_ptr()) for i in range(3)]
    eng.collect(pend.pop(0))                        # pass 0 done; passes 1, 2 queued; copy 2 may still run during pass 1
    eng.collect(pend.pop(0))
    mid = dict(eng.step_times_ms())                 # = pass 2's events? (events are per step, overwritten by later passes)
    eng.collect(pend.pop(0))
    worst = {k: round(v / alone[k], 2) for k, v in piped.items() if alone[k] > 0 and v / alone[k] > 1.15}
    res["steps_slower_than_1.15x_in_last_pipelined_pass"] = worst
    eng.set_timing(0)
    for depth in (2, 3):
        pipeline(6, depth)
        t0 = time.perf_counter(); pipeline(60, depth); dt = (time.perf_counter() - t0) / 60 * 1e3
        res[f"depth{depth}_timing_off_ms_per_step"] = round(dt, 4)
    out[model] = res
    eng.close()
print(json.dumps(out, indent=1))
