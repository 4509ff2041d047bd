mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t46_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t46_tests.log
tail -4 gpurun_out/t46_tests.log | cut -c1-250
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/t46_bench_resnet50_2gpu.json 2> gpurun_out/t46_bench_resnet50_2gpu.err; tail -1 gpurun_out/t46_bench_resnet50_2gpu.json | cut -c1-330
