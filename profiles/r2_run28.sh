mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest28.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2_pytest28.log
for k in 20 50; do python bench.py --steps $k --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('steps', d['steps'], 'ms/step', d['ms_per_step'], 'e2e total ms', d['e2e']['ms_total'], 'e2e/value', d['e2e']['value']/d['value'])"; done
