mkdir -p gpurun_out
N=${NGPU:-8}
run() { # impl workload
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --impl $1 --workload $2 --steps 20 --warmup 5 > gpurun_out/t60_bench_$2_$1_${N}gpu.json 2> gpurun_out/t60_bench_$2_$1_${N}gpu.err
  echo "rc=$? $1 $2"; tail -1 gpurun_out/t60_bench_$2_$1_${N}gpu.json | cut -c1-300
}
run ours resnet50
run ours bert-base
run ours bert-large
timeout 900 python -m pytest tests/test_gpu_collectives.py -m gpu -x -q 2>&1 | tail -2
