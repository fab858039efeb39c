# launch-configuration sweep of the headline Euler kernel after the instruction removals (tuning build: DG1D_TUNE_VARIANTS)
P='import json,sys
d=json.loads(sys.stdin.read()); print("c4 tune %s: %.3e DOF/s  %.4f ms/step  frac %.3f"%(sys.argv[1],d["value"],d["ms_per_step"],d["roofline"]["frac"]))'
for id in 0 1 2 3 4 5 6 7 8 9 10 11 12 1 0; do
  DG1D_LIB=$PWD/fem_b200/libdg1d_tune.so DG1D_TUNE=$id python bench.py --workload c4 --no-e2e --no-cpu-baseline --steps 30 2>/dev/null | python -c "$P" $id
done
