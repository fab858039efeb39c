mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest36.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest36.log
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
for rep in 1 2; do for w in c4 c4roe c5 c3; do
DG1D_PIPE=0 python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" nopipe $w
python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" pipe $w
done; done 2>&1 | tee gpurun_out/r2_ab36.txt
