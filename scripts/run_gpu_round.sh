mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t55_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t55_tests.log
tail -3 gpurun_out/t55_tests.log | cut -c1-250; grep -n "^E " gpurun_out/t55_tests.log | head -8
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/t55_bench_resnet.json 2>/dev/null; tail -1 gpurun_out/t55_bench_resnet.json | cut -c80-200
timeout 600 python bench.py --workload bert-base --steps 30 --warmup 5 > gpurun_out/t55_bench_bertbase.json 2>/dev/null; tail -1 gpurun_out/t55_bench_bertbase.json | cut -c80-200
timeout 600 python bench.py --workload bert-large --steps 20 --warmup 5 > gpurun_out/t55_bench_bertlarge.json 2>/dev/null; tail -1 gpurun_out/t55_bench_bertlarge.json | cut -c80-200
