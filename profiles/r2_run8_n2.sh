mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/dist_check.py > gpurun_out/r2_dist_check_n2_fused.log 2>&1; echo "dist_check fused rc $?"; grep "bit-identical\|Error\|error" gpurun_out/r2_dist_check_n2_fused.log | tail -14
S='import json,sys
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(sys.argv[1], d["value"], d["ms_per_step"], d.get("ms_per_step_nccl"), round(d["roofline"]["frac"],3), {k:(round(v["roofline_frac"],3), round(v["ms_per_step"],4)) for k,v in d.get("workloads",{}).items()})'
for rep in 1 2; do
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e 2> gpurun_out/r2_bench_n2_fused.err | tee gpurun_out/r2_bench_n2_fused.json | python -c "$S" fused
DG1D_PEER_FUSED=0 timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e 2> gpurun_out/r2_bench_n2_unfused.err | python -c "$S" unfused
done
python bench.py --no-e2e --no-cpu-baseline --steps 20 2>/dev/null | python -c "$S" n1
tail -c 300 gpurun_out/r2_bench_n2_fused.err
