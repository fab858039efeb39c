mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
python -m pytest tests -m gpu -x -q -k "two_parts or halo" > gpurun_out/r2_pytest25.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest25.log
timeout 600 $TR tests/dist_check.py > gpurun_out/r2_dist_check_n2.log 2>&1; echo "dist_check rc $?"; grep -c "bit-identical=True" gpurun_out/r2_dist_check_n2.log
S='import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(sys.argv[1], d["config"]["name"], round(d["ms_per_step"],4), d["state_finite"])'
for w in c3 c4 c2big c2; do
python bench.py --workload $w --no-e2e --no-cpu-baseline --no-secondary --steps 100 2>/dev/null | python -c "$S" n1
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" deep
DG1D_DIST_DEEP_EDGES=0 $TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary 2>/dev/null | python -c "$S" flat
$TR bench.py --gpus 2 --workload $w --steps 100 --no-e2e --no-secondary --transport nccl 2>/dev/null | python -c "$S" deep_nccl
done
