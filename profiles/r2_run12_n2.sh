mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest12.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest12.log
timeout 600 $TR tests/dist_check.py > gpurun_out/r2_dist_check_n2_fused.log 2>&1; echo "dist_check rc $?"; grep "bit-identical\|Error\|error" gpurun_out/r2_dist_check_n2_fused.log | grep -c True
S='import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(sys.argv[1], d["config"]["name"], round(d["ms_per_step"],4), d["state_finite"])'
for w in c3 c2 c4 c2big; do
python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary --steps 100 2>/dev/null | python -c "$S" n1
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" fused
DG1D_PEER_FUSED=0 $TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" unfused
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary --transport nccl 2>/dev/null | python -c "$S" nccl
done
