# same-box A/B: bash profiles/r2_ab.sh "<libs>" "<workloads>" [reps]; lib "new" = fem_b200/libdg1d.so, else fem_b200/libdg1d_<lib>.so
libs=${1:-"prev new"}; wls=${2:-"c4 c4roe c5 c3"}; reps=${3:-2}
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
for rep in $(seq $reps); do for w in $wls; do for lib in $libs; do
  if [ $lib = new ]; then unset DG1D_LIB; else export DG1D_LIB=$PWD/fem_b200/libdg1d_$lib.so; fi
  python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" $lib $w
done; done; done
unset DG1D_LIB
