cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/r2_gpu_suite.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2_gpu_suite.log
grep -E "^(FAILED|ERROR)|Error|assert" gpurun_out/r2_gpu_suite.log | head -20
