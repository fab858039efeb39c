mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest19.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest19.log
bash profiles/r2_ab.sh "prev new" "c4 c4roe c5 c3" 2 2>&1 | tee gpurun_out/r2_ab19.txt
