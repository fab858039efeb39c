for w in c5; do for id in 0 1 2 3 4 5 6 7; do
  DG1D_TUNE=$id python bench.py --workload $w --no-e2e --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w tune $id: %.3e DOF/s  %.3f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"
done; done
for k in 1 2 3 4; do python bench.py --workload c2 --degree $k --cells 20000000 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('linear k=$k 2e7 cells: %.3e DOF/s  %.3f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"; done
for w in c4 c3; do python bench.py --workload $w --no-e2e --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w default: %.3e DOF/s  %.3f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"; done
