"""Does running a 400-image batch as several smaller passes pay (activations of the early layers then fit
the 126 MB L2)?  Device-resident inputs, CUDA-event timing, same engine (max_batch 400)."""
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
bufs = [torch.randint(0, 255, (N * IMG,), dtype=torch.uint8, device=dev) for _ in range(3)]
hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (N, 1))
d_idx = torch.zeros(N, dtype=torch.int32, device=dev); d_prob = torch.zeros(N, dtype=torch.float32, device=dev)
out = {}
for model in ("alexnet", "resnet18"):
    eng = Engine(model, make_state_dict(model, seed=0), max_batch=N)
    st = torch.cuda.Stream(); eng.set_stream(st.cuda_stream)
    res = {}
    for sub in (400, 200, 100, 50):
        offs = np.arange(sub, dtype=np.int64) * IMG
        def step(i):
            b = bufs[i % 3]
            for k in range(0, N, sub):
                eng.infer_device(b.data_ptr() + k * IMG, offs, hw[:sub], d_idx.data_ptr() + 4 * k, d_prob.data_ptr() + 4 * k)
        for i in range(3):
            step(i)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for i in range(20):
            step(i)
        b.record(st); torch.cuda.synchronize()
        res[sub] = round(a.elapsed_time(b) / 20, 4)
    out[model] = res
    eng.close()
print(json.dumps(out))
