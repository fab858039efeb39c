mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
$TR tests/dist_check.py > gpurun_out/r2_dist_check_n2.log 2>&1; echo "dist_check rc $?"; grep -v "^W\|^\[" gpurun_out/r2_dist_check_n2.log | tail -14
( time $TR bench.py --gpus 2 --steps 20 --warmup 5 ) > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench rc $?"; tail -c 1200 gpurun_out/r2_bench_n2.err
python -c "
import json
for l in open('gpurun_out/r2_bench_n2.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d.get('ms_per_step_nccl'), d['roofline']['frac'], d.get('dist_parity'), {k:(round(v['roofline_frac'],3), v['ms_per_step']) for k,v in d['workloads'].items()}, d['e2e']['value'])"
$TR bench.py --gpus 2 --steps 20 --warmup 5 --impl reference > gpurun_out/r2_bench_n2_ref.json 2> gpurun_out/r2_bench_n2_ref.err; echo "ref rc $?"; head -c 300 gpurun_out/r2_bench_n2_ref.json
