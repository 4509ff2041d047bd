mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t57_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t57_tests.log
tail -3 gpurun_out/t57_tests.log | cut -c1-250; grep -n "^E " gpurun_out/t57_tests.log | head -8
timeout 600 python benchmarks/profile_step.py --workload bert-base --top 8 2>&1 | grep "^#\|ln_" | cut -c1-150
timeout 600 python benchmarks/profile_step.py --workload bert-large --top 12 2>&1 | grep "^#\|ln_" | cut -c1-150
