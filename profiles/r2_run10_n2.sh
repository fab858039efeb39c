mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dist_check.py --peer > gpurun_out/r2_dist_check_n2_fused.log 2>&1; echo "dist_check fused rc $?"; grep "bit-identical\|Error\|error" gpurun_out/r2_dist_check_n2_fused.log | tail -6
S='import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(sys.argv[1], d["config"]["name"], round(d["ms_per_step"],4), d["state_finite"])'
for w in c3 c2 c4; do
python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary --steps 100 2>/dev/null | python -c "$S" n1
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" fused
DG1D_DBG_NO_REDUCE=1 $TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" fused_noreduce
DG1D_PEER_FUSED=0 $TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" unfused
done
