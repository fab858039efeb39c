mkdir -p gpurun_out
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s tune %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
export DG1D_LIB=$PWD/fem_b200/libdg1d_tune.so
for id in 0 1 2 4 5 6 7 8 9 10 11 14 15 0; do
  DG1D_TUNE=$id python bench.py --workload c4 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" c4 $id
done 2>&1 | tee gpurun_out/r2_sweep2_c4.txt
for id in 0 1 2 4 5 6 7 8 10 11 13 17 0; do
  DG1D_TUNE=$id python bench.py --workload c5 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" c5 $id
done 2>&1 | tee gpurun_out/r2_sweep2_c5.txt
for id in 0 1 2 14 4 5 0; do
  DG1D_TUNE=$id python bench.py --workload c4roe --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" c4roe $id
done 2>&1 | tee gpurun_out/r2_sweep2_c4roe.txt
