"""Which plan steps slow a concurrent pinned H2D copy?  Runs the plan prefix 0..k in a loop (stop_after)
while 20 copies of 60 MB are timed on another stream."""
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
N, IMG = 400, 224 * 224 * 3
host = torch.randint(0, 255, (N * IMG,), dtype=torch.uint8).pin_memory()
d_in = torch.empty(N * IMG, dtype=torch.uint8, device=dev); d_in.copy_(host)
d_copy = torch.empty_like(d_in)
offsets = np.arange(N, dtype=np.int64) * IMG
hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (N, 1))
d_idx = torch.zeros(N, dtype=torch.int32, device=dev); d_prob = torch.zeros(N, dtype=torch.float32, device=dev)
out = {}
for model in sys.argv[1:] or ["alexnet", "resnet18"]:
    eng = Engine(model, make_state_dict(model, seed=0), max_batch=N)
    s_k, s_c = torch.cuda.Stream(), torch.cuda.Stream()
    eng.set_stream(s_k.cuda_stream)
    names = [n for n, _ in eng.steps()]
    res = []
    for k in range(len(names)):
        eng.stop_after(k)
        def kernels(reps):
            for _ in range(reps):
                eng.infer_device(d_in.data_ptr(), offsets, hw, d_idx.data_ptr(), d_prob.data_ptr())
        kernels(2); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s_k); kernels(10); b.record(s_k); torch.cuda.synchronize()
        t_pass = a.elapsed_time(b) / 10
        reps = max(2, int(30.0 / t_pass))
        ca, cb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s_k); kernels(reps); b.record(s_k)
        with torch.cuda.stream(s_c):
            ca.record()
            for _ in range(20):
                d_copy.copy_(host, non_blocking=True)
            cb.record()
        torch.cuda.synchronize()
        res.append({"upto": names[k], "pass_ms": round(t_pass, 4), "pass_ms_overlapped": round(a.elapsed_time(b) / reps, 4),
                    "copy_GBps": round(N * IMG / (ca.elapsed_time(cb) / 20) / 1e6, 1)})
    out[model] = res
    eng.close()
print(json.dumps(out, indent=0))
