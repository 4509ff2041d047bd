set -x
mkdir -p gpurun_out
timeout 300 python scripts/debug_stem.py > gpurun_out/debug_stem.log 2>&1; cat gpurun_out/debug_stem.log | tail -40
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "pool" > gpurun_out/t31_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t31_tests.log
tail -5 gpurun_out/t31_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/t31_bench_resnet.json 2> gpurun_out/t31_bench_resnet.err; tail -1 gpurun_out/t31_bench_resnet.json
