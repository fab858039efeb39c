mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
( time $TR bench.py --gpus $N --steps 20 --warmup 5 ) > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench rc $?"
grep "bit-identical" gpurun_out/r2_bench_n$N.err > gpurun_out/r2_dist_check_n$N.log; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n$N.log
python -c "
import json
for l in open('gpurun_out/r2_bench_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d.get('ms_per_step_nccl'), d['roofline']['frac'], d.get('dist_parity'), {k:(round(v['roofline_frac'],3), round(v['ms_per_step'],4)) for k,v in d['workloads'].items()}, d['e2e']['value'], d['clocks'])"
tail -5 gpurun_out/r2_bench_n$N.err | cut -c1-300
$TR bench.py --gpus $N --steps 20 --warmup 5 --impl reference > gpurun_out/r2_bench_n${N}_ref.json 2>/dev/null; echo "ref rc $?"; head -c 200 gpurun_out/r2_bench_n${N}_ref.json
