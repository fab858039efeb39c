mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest4.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2_pytest4.log
tail -6 gpurun_out/r2_pytest4.log
( time python bench.py ) > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc $?"; tail -c 400 gpurun_out/r2_bench_default.err
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_default.json').readline()); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_total'], {k:(round(v['roofline_frac'],3), v['ms_per_step']) for k,v in d['workloads'].items()}, d['cpu_baseline'])"
( time python bench.py --impl reference ) > gpurun_out/r2_bench_refarm.json 2> gpurun_out/r2_bench_refarm.err; tail -c 300 gpurun_out/r2_bench_refarm.err; head -c 600 gpurun_out/r2_bench_refarm.json
