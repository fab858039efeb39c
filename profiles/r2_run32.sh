P='import json,sys
d=json.loads(sys.stdin.read()); print("%s %s: %.4f ms/step"%(sys.argv[1],sys.argv[2],d["ms_per_step"]))'
for w in c3 c4 c2; do
python bench.py --workload $w --steps 100 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" plain $w
DG1D_DBG_RECORD=1 python bench.py --workload $w --steps 100 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" record_between $w
DG1D_PDL=0 python bench.py --workload $w --steps 100 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "$P" no_pdl $w
done
