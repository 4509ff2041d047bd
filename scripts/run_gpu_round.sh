#!/usr/bin/env bash
# Standard single-box check used during development (run on a B200 box from the repo root):
#   GPU test suite, one bench line per workload, per-kernel step profile of the default workload.
# Results land in gpurun_out/ (scratch); copy what should be kept into profiles/.
#   usage: bash scripts/run_gpu_round.sh [tag]
tag=${1:-check}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${tag}_tests.log
tail -3 gpurun_out/${tag}_tests.log | cut -c1-250
grep -n "^E " gpurun_out/${tag}_tests.log | head -8
for wl in resnet50 bert-base bert-large; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 > gpurun_out/${tag}_bench_${wl}.json 2> gpurun_out/${tag}_bench_${wl}.err
  tail -1 gpurun_out/${tag}_bench_${wl}.json | cut -c1-220
done
timeout 600 python benchmarks/profile_step.py --workload resnet50 --top 30 > gpurun_out/${tag}_profile_resnet50.log 2>&1
grep -A12 "^# resnet" gpurun_out/${tag}_profile_resnet50.log | cut -c1-180
