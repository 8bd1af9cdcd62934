cd $GRAFT_REPO_ROOT
run() {
  CMD="python tools/neck_bench.py --neck $1 --batch $2 --precision bf16"
  $CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s $3 -c $4 --csv --log-file gpurun_out/r2_ll2_$1_$2.csv $CMD > /dev/null 2>&1
  echo "$1 $2 rc=$?"
}
run FastIndoorImVoxelNeck 1 200 15
run ImVoxelNeck 1 560 47
