mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
( time $TR bench.py --gpus 2 --steps 50 --warmup 5 ) > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench rc $?"
grep "bit-identical" gpurun_out/r2_bench_n2.err > gpurun_out/r2_dist_check_n2.log; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n2.log
python -c "
import json
for l in open('gpurun_out/r2_bench_n2.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d.get('ms_per_step_nccl'), d['roofline']['frac'], {k:(round(v['roofline_frac'],3), round(v['ms_per_step'],4)) for k,v in d['workloads'].items()}, d['e2e']['value'])"
DG1D_PEER_FUSED=1 timeout 600 $TR tests/dist_check.py --peer > gpurun_out/r2_dist_check_n2_fused.log 2>&1; echo "fused rc $?"; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n2_fused.log
python bench.py --steps 50 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('n1', d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['workloads'].items()})"
