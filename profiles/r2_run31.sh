mkdir -p gpurun_out
P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f sm %s %s"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"],d["clocks"]["sm_mhz"],d["clocks"]["reasons"]))'
python bench.py --workload c3 --post-shock --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | tee gpurun_out/r2_bench_c3_postshock.json | python -c "$P" new c3-post-shock
bash profiles/r2_ab.sh "noskip new" "c4 c4roe c5 c3" 1
