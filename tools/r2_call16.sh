cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out /tmp/ncu
B="python bench.py --steps 2 --warmup 3 --no-extras"
$B > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_kitti_r2.csv $B > gpurun_out/ncu_ll.log 2>&1
echo "ll rc=$?"
N="python tools/neck_bench.py --neck KittiImVoxelNeck --batch 8 --precision bf16"
$N > gpurun_out/plain_n1.log 2>&1 && ncu --set full --clock-control none -k regex:conv3d_tc_halo -s 18 -c 9 -o /tmp/ncu/bf16 -f $N > gpurun_out/ncu_n1.log 2>&1
echo "bf16 rc=$?"
python tools/ncu_summary.py /tmp/ncu/bf16.ncu-rep --json gpurun_out/traffic_neck_imvoxelnet_kitti.json > gpurun_out/ncu_kitti_neck_r2.txt 2>&1
N="python tools/neck_bench.py --neck KittiImVoxelNeck --batch 8 --precision fp32"
$N > gpurun_out/plain_n2.log 2>&1 && ncu --set full --clock-control none -k regex:conv3d_tc_halo -s 18 -c 9 -o /tmp/ncu/x3 -f $N > gpurun_out/ncu_n2.log 2>&1
echo "x3 rc=$?"
python tools/ncu_summary.py /tmp/ncu/x3.ncu-rep > gpurun_out/ncu_kitti_neck_x3_r2.txt 2>&1
ls -la gpurun_out/ | tail -12
