set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_x3_gpu.py -x -q -s > gpurun_out/r2_x3_layers.log 2>&1; echo "x3 layers rc=$?"
tail -5 gpurun_out/r2_x3_layers.log
timeout 1200 python -m pytest tests/test_neck_gpu.py -q -s -k "full_size" > gpurun_out/r2_neck_full.log 2>&1; echo "neck full rc=$?"
grep -E "max err|passed|failed|Error" gpurun_out/r2_neck_full.log | tail -60
for n in KittiImVoxelNeck:8 FastIndoorImVoxelNeck:8 NuScenesImVoxelNeck:4 ImVoxelNeck:32; do
  timeout 300 python tools/neck_bench.py --neck ${n%%:*} --batch ${n##*:} --precision fp32 --layers > gpurun_out/r2_neckbench_fp32_${n%%:*}.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/r2_neckbench_fp32_${n%%:*}.log
done
