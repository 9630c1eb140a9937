#!/bin/bash
# One GPU-box session: parity tests (one pytest process per file), diagnostics, bench, profiles.
# Usage (from the repo root, under gpurun):  bash tools/gpu_round.sh [tests|bench|ncu ...]
set -u
mkdir -p gpurun_out
WHAT="${*:-tests bench}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.csv 2>&1
for w in $WHAT; do
case $w in
diag)
  timeout 600 python tools/gpu_diag.py > gpurun_out/diag.log 2>&1; echo "diag rc=$?" ;;
tests)
  for f in tests/test_gpu_preprocess.py tests/test_gpu_conv.py tests/test_gpu_model.py tests/test_gpu_boxrun.py tests/test_zz_no_alloc.py; do
    b=$(basename $f .py)
    timeout 1500 python -m pytest $f -m gpu -q -rA -s -p no:cacheprovider > gpurun_out/$b.log 2>&1
    echo "$b rc=$? $(tail -1 gpurun_out/$b.log)"
  done ;;
suite)
  timeout 1500 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/suite.log 2>&1
  echo "suite rc=$? $(tail -1 gpurun_out/suite.log)" ;;
bench)
  timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
  echo "bench rc=$?"; cat gpurun_out/bench.json | cut -c1-1500
  for m in alexnet resnet18; do python tools/kms.py gpurun_out/bench_detail.json $m; done ;;
benchref)
  timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
  echo "benchref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json ;;
ncu)
  for m in alexnet resnet18; do
    # one pass = 12 launches (alexnet) / 20 (resnet18) incl. the record copy; skip the 3 warm-up passes, list 2 passes
    if [ $m = alexnet ]; then SK=36; CN=24; else SK=60; CN=40; fi
    B="--steps 2 --warmup 3 --sustain 0 --range-images 0 --no-cpu --models $m"
    timeout 600 python bench.py $B > gpurun_out/ncu_plain_$m.log 2>&1 &&
    timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $SK -c $CN --csv \
        --log-file gpurun_out/launches_$m.csv python bench.py $B \
        > gpurun_out/ncu_launch_$m.log 2>&1
    echo "ncu launches $m rc=$?"
  done ;;
ncufull)
  m=${NCU_MODEL:-resnet18}
  B="--steps 2 --warmup 3 --sustain 0 --range-images 0 --no-cpu --models $m"
  timeout 600 python bench.py $B > gpurun_out/ncu_plain_full.log 2>&1 &&
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:${NCU_REGEX:-kernel} \
      -s ${NCU_SKIP:-60} -c ${NCU_COUNT:-20} -o gpurun_out/prof_${NCU_TAG:-$m} -f python bench.py $B \
      > gpurun_out/ncu_full_$m.log 2>&1
  echo "ncu full $m rc=$?" ;;
smoke)
  timeout 900 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; grep "^smoke" gpurun_out/smoke.log ;;
scen)
  timeout 1500 python tools/box_scenarios.py --gpus ${SCEN_GPUS:-2} --scenario ${SCEN_WHAT:-all} \
      --batches ${SCEN_BATCHES:-64,128,256,512,1024,2048,4096} --per-rank ${SCEN_PER_RANK:-10000} \
      > gpurun_out/scen_${SCEN_GPUS:-2}gpu.log 2>&1
  echo "scen rc=$?"; grep '^{' gpurun_out/scen_${SCEN_GPUS:-2}gpu.log | cut -c1-700 ;;
benchN)
  N=${BENCH_GPUS:-2}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
  echo "bench$N rc=$?"; cut -c1-900 gpurun_out/bench_${N}gpu.json ;;
esac
done
