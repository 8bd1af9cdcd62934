cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_x3_gpu.py -q -s > gpurun_out/r2_x3_layers.log 2>&1; echo "x3 layers rc=$?"
grep -E "3xTF32|one-signed|passed|failed" gpurun_out/r2_x3_layers.log | tail -40
timeout 1200 python -m pytest tests/test_neck_gpu.py -q -s -k "full_size and fp32" > gpurun_out/r2_neck_full.log 2>&1; echo "neck full rc=$?"
grep -E "max err /|passed|failed" gpurun_out/r2_neck_full.log | tail -60
for g in 1 2 4 100000; do
  echo "== drain_groups $g"
  IVX_X3_DRAIN_GROUPS=$g timeout 300 python tools/neck_bench.py --neck KittiImVoxelNeck --batch 8 --precision fp32 --layers 2>&1 | tail -12
done
for g in 4 16 100000; do
  echo "== drain_steps $g"
  IVX_X3_DRAIN_STEPS=$g timeout 300 python tools/neck_bench.py --neck FastIndoorImVoxelNeck --batch 8 --precision fp32 --layers 2>&1 | tail -18
done
