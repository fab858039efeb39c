for w in c4 c3; do for id in 0 1 2 3 4 5 6 7 8 9 10 11; do
  DG1D_TUNE=$id python bench.py --workload $w --no-e2e --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w tune $id: %.3e DOF/s  %.3f ms/step  frac %.3f'%(d['value'],d['ms_per_step'],d['roofline']['frac']))"
done; done
