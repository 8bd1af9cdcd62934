cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_suite.log 2>&1 ) 2>&1 | tail -3; echo "suite rc=$?"
tail -15 gpurun_out/r2_gpu_suite.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
