# same-box A/B of two builds on every headline workload: DG1D_LIB selects the library; alternate to average out drift
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"]))'
for rep in 1 2; do for w in c4 c5 c3 c2; do
  DG1D_LIB=$PWD/fem_b200/libdg1d_old.so python bench.py --workload $w --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" old "$w"
  python bench.py --workload $w --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" new "$w"
done; done
