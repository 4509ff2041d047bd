mkdir -p gpurun_out
for mb in 0 16 32 64 128; do
echo "max_mb=$mb"; DDL_CONV_BWD_OVERLAP_MAX_MB=$mb timeout 600 python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 | cut -c80-190
done
