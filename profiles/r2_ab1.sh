# round 2, first GPU session: parity suite, then same-box A/B of round-1 build / no-shortcut build / current build
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2_pytest1.log
tail -5 gpurun_out/r2_pytest1.log
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
for rep in 1 2; do for w in c4 c4roe c5 c3; do
  for lib in old noskip new; do
    if [ $lib = new ]; then unset DG1D_LIB; else export DG1D_LIB=$PWD/fem_b200/libdg1d_$lib.so; fi
    python bench.py --workload $w --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" $lib $w
  done
done; done 2>&1 | tee gpurun_out/r2_ab1.txt
unset DG1D_LIB
python bench.py > gpurun_out/r2_bench_c4_first.json 2> gpurun_out/r2_bench_c4_first.err; tail -c 1500 gpurun_out/r2_bench_c4_first.json
