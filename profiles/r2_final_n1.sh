mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_final.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest_final.log
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_default.json').readline()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_total'], {k:(round(v['roofline_frac'],3), round(v['ms_per_step'],4)) for k,v in d['workloads'].items()}, d['clocks'], d['gpu_launches'])"
python bench.py --impl reference > gpurun_out/r2_bench_refarm.json 2>/dev/null; echo "ref rc $?"
