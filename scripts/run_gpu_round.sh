mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t54_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t54_tests.log
tail -4 gpurun_out/t54_tests.log | cut -c1-250; grep -n "^E " gpurun_out/t54_tests.log | head -8
