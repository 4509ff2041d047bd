mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "attention or attn or bert" > gpurun_out/t48_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t48_tests.log
tail -6 gpurun_out/t48_tests.log | cut -c1-250
timeout 600 python bench.py --workload bert-large --steps 20 --warmup 5 2>/dev/null | tail -1 | cut -c80-200
DDL_ATTN_FWD=resident timeout 600 python bench.py --workload bert-large --steps 20 --warmup 5 2>/dev/null | tail -1 | cut -c80-200
timeout 600 python bench.py --workload bert-base --steps 20 --warmup 5 2>/dev/null | tail -1 | cut -c80-200
timeout 600 python benchmarks/profile_step.py --workload bert-large --top 8 2>&1 | grep "attn" | cut -c1-150
timeout 600 python benchmarks/profile_step.py --workload bert-base --top 12 2>&1 | grep "attn" | cut -c1-150
