"""Prints per-step kernel times (and pipeline counters) of one model from bench.py's detail file:
    python tools/kms.py gpurun_out/bench_detail.json resnet18 [step names ...]"""
import json
import sys

d = json.load(open(sys.argv[1]))
m = sys.argv[2]
r = d["models"][m]
keys = sys.argv[3:]
print(m, "ms/step", round(r["ms_per_step"], 4), {k: v for k, v in r["kernel_ms"].items() if not keys or k in keys},
      {k: v for k, v in r.get("wait_cycles", {}).items() if k in keys})
