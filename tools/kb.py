"""Experiment helper: per-step kernel times of a bench_detail.json, scaled to a 400-image step."""
import json
import sys

d = json.load(open(sys.argv[1]))
b = int(sys.argv[2])
for m, r in d["models"].items():
    k = r["kernel_ms"]
    print(m, "batch", b, "ms/400 img", round(r["ms_per_step"] * 400 / b, 4),
          {n: round(v * 400 / b, 4) for n, v in k.items() if v * 400 / b > 0.004})
