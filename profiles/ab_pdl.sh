P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.3e DOF/s  %.4f ms/step  frac %.3f"%(sys.argv[1],sys.argv[2],d["value"],d["ms_per_step"],d["roofline"]["frac"]))'
for rep in 1 2; do for w in c2 c5 c3 c4; do
  DG1D_PDL=0 python bench.py --workload $w --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" pdl=0 $w
  DG1D_PDL=1 python bench.py --workload $w --no-e2e --no-cpu-baseline 2>/dev/null | python -c "$P" pdl=1 $w
done; done
DG1D_PDL=0 python profiles/time_c1.py 2>/dev/null | tail -2
DG1D_PDL=1 python profiles/time_c1.py 2>/dev/null | tail -2
