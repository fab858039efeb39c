mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest3.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2_pytest3.log
tail -4 gpurun_out/r2_pytest3.log
bash profiles/r2_ab.sh "prev new" "c4 c4roe c5 c3" 2 2>&1 | tee gpurun_out/r2_ab3.txt
python bench.py --steps 20 > gpurun_out/r2_bench_full.json 2> gpurun_out/r2_bench_full.err; echo "bench rc $?"; tail -c 600 gpurun_out/r2_bench_full.err; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_full.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], {k:(v['roofline_frac'], v['ms_per_step']) for k,v in d['workloads'].items()}, d['cpu_baseline'])"
