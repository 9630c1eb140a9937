#!/bin/bash
# Runs the ResNet18 / AlexNet device-resident bench under a few environment settings (kernel-selection
# experiments) and prints images/s + ms per step for each.  Usage: bash tools/sweep_env.sh "A=1 B=2" "C=3" ...
mkdir -p gpurun_out
for cfg in "$@"; do
  for m in ${SWEEP_MODELS:-resnet18}; do
    out=$(env $cfg python bench.py --models $m --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'])")
    echo "[$cfg] $m: $out"
  done
done
