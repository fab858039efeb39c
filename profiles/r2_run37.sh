mkdir -p gpurun_out
W=$1
CMD="python bench.py --workload $W --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r2_plain_$W.log 2>&1 && ncu --set full --clock-control none -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2d_$W $CMD > gpurun_out/r2d_ncu_$W.log 2>&1
tail -2 gpurun_out/r2d_ncu_$W.log
