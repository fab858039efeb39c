mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest18.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest18.log
bash profiles/r2_ab.sh "new" "c4 c4roe c5 c3" 1 2>&1 | tee gpurun_out/r2_ab18.txt
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/r2_plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stage_kernel -s 9 -c 3 -f -o gpurun_out/r2c_c4 $CMD > gpurun_out/r2c_ncu_c4.log 2>&1
tail -2 gpurun_out/r2c_ncu_c4.log
