mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest26.log 2>&1; echo "pytest rc $?"; tail -2 gpurun_out/r2_pytest26.log
bash profiles/r2_run15_n8.sh 8
