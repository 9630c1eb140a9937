#!/bin/bash
# Per-step kernel times of one model under environment settings: bash tools/kernel_ms_env.sh alexnet "A=1" "B=2"
m=$1; shift
for cfg in "$@"; do
  env $cfg python bench.py --models $m --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('[$cfg]', d['ms_per_step'], {k: v for k, v in d['kernel_ms'].items()})
        w=d.get('wait_cycles',{})
        for k,v in w.items():
            if 'classifier' in k: print('   wait', k, [round(x/148) for x in v[:4]])"
done
