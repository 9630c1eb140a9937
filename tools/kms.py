"""Prints the per-step kernel times of one model from a bench.py line (JSON on stdin or a file)."""
import json, sys
d = json.load(open(sys.argv[1]))
m = sys.argv[2] if len(sys.argv) > 2 else None
r = d.get(m, d) if m and m in d else d
print(r.get("value"), r.get("ms_per_step"), {k: v for k, v in r["kernel_ms"].items() if not sys.argv[3:] or k in sys.argv[3:]})
