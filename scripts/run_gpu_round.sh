set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t38_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/t38_tests.log
tail -4 gpurun_out/t38_tests.log | cut -c1-250
timeout 600 python bench.py --workload bert-large --steps 20 --warmup 5 > gpurun_out/t38_bench_bertlarge.json 2> gpurun_out/t38_bench_bertlarge.err; tail -1 gpurun_out/t38_bench_bertlarge.json | cut -c1-300
timeout 600 python bench.py --workload bert-base --steps 20 --warmup 5 > gpurun_out/t38_bench_bertbase.json 2> gpurun_out/t38_bench_bertbase.err; tail -1 gpurun_out/t38_bench_bertbase.json | cut -c1-300
timeout 300 python benchmarks/ncu_target.py stem && timeout 900 ncu --set full --clock-control none --import-source on -k regex:stem_ -c 4 -o gpurun_out/prof_stem -f python benchmarks/ncu_target.py stem > gpurun_out/ncu_stem.log 2>&1; tail -3 gpurun_out/ncu_stem.log
