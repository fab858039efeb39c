for id in 0 1 2 3 4 5 6 7; do
  DG1D_TUNE=$id python bench.py --workload c2 --cells 20000000 --no-e2e --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('linear k=1 2e7 cells tune $id: %.3e DOF/s  %.3f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"
  DG1D_TUNE=$id python bench.py --workload c2 --no-e2e --no-cpu-baseline --steps 50 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('linear k=1 1e6 cells tune $id: %.3e DOF/s  %.4f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"
done
