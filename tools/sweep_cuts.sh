#!/bin/bash
# Experiment: blocking-call throughput (single_call_value) for several sub-batch cuts of one 400-image call
# (IDN_SINGLE_CALL_CUTS); run under gpurun from the repo root.
B="--steps 5 --warmup 3 --sustain 0 --range-images 0 --no-cpu"
for c in "34,33,33" "25,25,25,25" "40,30,30" "28,36,36" "20,40,40" "50,50" "20,20,20,20,20" "10,30,60" "8,23,69"; do
  IDN_SINGLE_CALL_CUTS=$c python bench.py $B > gpurun_out/b1.json 2>gpurun_out/b1.err
  python - "$c" <<'PY'
import json,sys
d=json.load(open('gpurun_out/b1.json'))
print(sys.argv[1], 'alexnet single', d['e2e']['single_call_value'], 'resnet18 single', d['resnet18']['single_call'])
PY
done
