mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t52_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t52_tests.log
tail -4 gpurun_out/t52_tests.log | cut -c1-250; grep -n "^E " gpurun_out/t52_tests.log | head -5
timeout 600 python bench.py --workload bert-base --steps 20 --warmup 5 > gpurun_out/t52_bench_bertbase.json 2>/dev/null; tail -1 gpurun_out/t52_bench_bertbase.json | cut -c80-200
timeout 600 python benchmarks/profile_step.py --workload bert-base --top 30 2>&1 | grep -A30 "^# bert" | cut -c1-150 | grep -i "embed\|index\|Fill\|^#\|copy\|multi_tensor"
