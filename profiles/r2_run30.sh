mkdir -p gpurun_out
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s tune0 %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
export DG1D_LIB=$PWD/fem_b200/libdg1d_tune.so
for id in 0 15 1 8 9 4 5 10 0; do
  DG1D_TUNE=0 DG1D_TUNE0=$id python bench.py --workload c4 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" c4 $id
done 2>&1 | tee gpurun_out/r2_sweep3_c4_stage0.txt
for id in 0 15 1 8 4; do
  DG1D_TUNE=0 DG1D_TUNE0=$id python bench.py --workload c4roe --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" c4roe $id
done 2>&1 | tee -a gpurun_out/r2_sweep3_c4_stage0.txt
