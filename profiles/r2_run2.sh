# round 2, second GPU session: parity suite; ncu launch list + full capture of the three c4 stage kernels; launch-config sweeps
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2_pytest2.log
tail -4 gpurun_out/r2_pytest2.log
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/r2_plain_c4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_c4.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2_c4 $CMD > gpurun_out/r2_ncu_full.log 2>&1
tail -3 gpurun_out/r2_ncu_full.log
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s tune %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
export DG1D_LIB=$PWD/fem_b200/libdg1d_tune.so
for id in 0 11 12 13 14 15 16 17 18 19 1 4 7 0; do
  DG1D_TUNE=$id python bench.py --workload c5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" c5 $id
done 2>&1 | tee gpurun_out/r2_sweep_c5.txt
for id in 0 1 2 3 4 5 7 9 10 11 13 14 15 16 17 18 0; do
  DG1D_TUNE=$id python bench.py --workload c4 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" c4 $id
done 2>&1 | tee gpurun_out/r2_sweep_c4.txt
for id in 0 1 2 3 14 15 0; do
  DG1D_TUNE=$id python bench.py --workload c4roe --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" c4roe $id
done 2>&1 | tee gpurun_out/r2_sweep_c4roe.txt
