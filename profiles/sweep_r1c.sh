# launch-configuration sweep of the ensemble kernel (Euler k=1, conserved-variable limiter; tuning build)
P='import json,sys
d=json.loads(sys.stdin.read()); print("c5 tune %s: %.3e DOF/s  %.4f ms/step  frac %.3f"%(sys.argv[1],d["value"],d["ms_per_step"],d["roofline"]["frac"]))'
for id in 0 1 2 3 4 5 6 7 8 9 10 0; do
  DG1D_LIB=$PWD/fem_b200/libdg1d_tune.so DG1D_TUNE=$id python bench.py --workload c5 --no-e2e --no-cpu-baseline --steps 50 2>/dev/null | python -c "$P" $id
done
