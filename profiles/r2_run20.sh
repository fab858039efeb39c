mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc $?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_default.json').readline()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_total'], {k:(round(v['roofline_frac'],3), round(v['ms_per_step'],4)) for k,v in d['workloads'].items()}, d['clocks'])"
python bench.py --impl reference > gpurun_out/r2_bench_refarm.json 2>/dev/null
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r2_plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_c4.csv $CMD > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2d_c4 $CMD > gpurun_out/r2d_ncu_c4.log 2>&1
