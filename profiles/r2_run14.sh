mkdir -p gpurun_out
for w in c5 c4; do
CMD="python bench.py --workload $w --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r2_plain_$w.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2b_$w $CMD > gpurun_out/r2b_ncu_$w.log 2>&1
tail -2 gpurun_out/r2b_ncu_$w.log
done
