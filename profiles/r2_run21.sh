mkdir -p gpurun_out
CMD5="python bench.py --workload c5 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-secondary"
$CMD5 > gpurun_out/r2_plain_c5.log 2>&1 && ncu --set full --clock-control none -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2d_c5 $CMD5 > gpurun_out/r2d_ncu_c5.log 2>&1
