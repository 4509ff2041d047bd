mkdir -p gpurun_out
timeout 300 python benchmarks/ncu_target.py conv1x1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv1x1_fwd -c 3 -s 3 -o gpurun_out/prof_conv1x1 -f python benchmarks/ncu_target.py conv1x1 > gpurun_out/ncu_conv1x1.log 2>&1; tail -2 gpurun_out/ncu_conv1x1.log
timeout 300 python benchmarks/ncu_target.py attention && timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_ -c 6 -s 6 -o gpurun_out/prof_attn2 -f python benchmarks/ncu_target.py attention > gpurun_out/ncu_attn2.log 2>&1; tail -2 gpurun_out/ncu_attn2.log
