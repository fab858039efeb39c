for mb in 32 128 512 4096; do DG1D_STAGE_MB=$mb python bench.py --steps 20 --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('stage MB $mb', 'steps', d['steps'], 'ms/step', d['ms_per_step'], 'e2e total ms', d['e2e']['ms_total'], 'copies ms', d['e2e']['ms_total']-20*d['ms_per_step'])"; done
