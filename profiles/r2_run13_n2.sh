mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dist_check.py > gpurun_out/r2_dist_check_n2.log 2>&1; echo "dist_check rc $?"; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n2.log
DG1D_PEER_FUSED=1 timeout 600 $TR tests/dist_check.py --peer > gpurun_out/r2_dist_check_n2_fused.log 2>&1; echo "dist_check fused rc $?"; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n2_fused.log
S='import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(sys.argv[1], d["config"]["name"], round(d["ms_per_step"],4), d["state_finite"])'
for w in c3 c4 c2; do
python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary --steps 100 2>/dev/null | python -c "$S" n1
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" default
DG1D_PEER_FUSED=1 $TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" fused
done
( time $TR bench.py --gpus 2 --steps 20 --warmup 5 ) > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench rc $?"
python -c "
import json
for l in open('gpurun_out/r2_bench_n2.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['value'], d['ms_per_step'], d.get('ms_per_step_nccl'), d['roofline']['frac'], {k:(round(v['roofline_frac'],3), v['ms_per_step']) for k,v in d['workloads'].items()}, d['e2e']['value'])"
tail -4 gpurun_out/r2_bench_n2.err
