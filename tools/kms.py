"""Prints per-step kernel times (and wait-cycle counters) of one model from a bench.py line."""
import json, sys
d = json.load(open(sys.argv[1]))
m = sys.argv[2] if len(sys.argv) > 2 else None
r = d.get(m, d) if m and m in d else d
keys = sys.argv[3:]
print(r.get("value"), r.get("ms_per_step"), {k: v for k, v in r["kernel_ms"].items() if not keys or k in keys},
      {k: v for k, v in r.get("wait_cycles", {}).items() if k in keys})
