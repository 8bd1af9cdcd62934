cd $GRAFT_REPO_ROOT
run() { # neck batch plan skip count
  CMD="python tools/neck_bench.py --neck $1 --batch $2 --precision bf16"
  IVX_HALO_PLAN=$3 $CMD > gpurun_out/plain.log 2>&1 && IVX_HALO_PLAN=$3 ncu --metrics gpu__time_duration.sum --clock-control none -s $4 -c $5 --csv --log-file gpurun_out/r2_ll_$1_$2_$3.csv $CMD > /dev/null 2>&1
  echo "$1 $2 $3 rc=$?"
}
for plan in score cost; do
  run FastIndoorImVoxelNeck 1 $plan 200 15
  run FastIndoorImVoxelNeck 2 $plan 200 15
  run FastIndoorImVoxelNeck 8 $plan 200 15
  run ImVoxelNeck 32 $plan 560 47
done
