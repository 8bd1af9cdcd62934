cd $GRAFT_REPO_ROOT
CMD="python tools/neck_bench.py --neck FastIndoorImVoxelNeck --batch 1 --precision bf16"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 30 --csv --log-file gpurun_out/r2_launches_fast_b1.csv $CMD > gpurun_out/ncu_fast_b1.log 2>&1
echo rc=$?
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r2_launches_fast_b1.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); gi=hdr.index('Grid Size')
tot=0
for r in rows[1:]:
    print(r[ki][:60], r[gi], r[vi]); 
PY
