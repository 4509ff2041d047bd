set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t36_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t36_tests.log
tail -4 gpurun_out/t36_tests.log | cut -c1-250
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/t36_bench_resnet.json 2> gpurun_out/t36_bench_resnet.err; tail -1 gpurun_out/t36_bench_resnet.json
DDL_STEM=cudnn timeout 600 python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 | cut -c1-200
timeout 600 python benchmarks/profile_step.py --workload resnet50 --top 60 > gpurun_out/prof36_resnet.log 2>&1; grep "stem_\|^# resnet" gpurun_out/prof36_resnet.log | cut -c1-200
