mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "conv1x1_weight or resnet" > gpurun_out/t44_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t44_tests.log
tail -5 gpurun_out/t44_tests.log | cut -c1-250
timeout 900 python benchmarks/conv_bench.py > gpurun_out/conv_bench.log 2>&1; grep "ws.fwd\|k1/s1" gpurun_out/conv_bench.log | head -8
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/t44_bench_resnet.json 2> gpurun_out/t44_bench_resnet.err; tail -1 gpurun_out/t44_bench_resnet.json | cut -c80-200
