cd $GRAFT_REPO_ROOT
for b in 1 2; do
 timeout 200 python tools/neck_bench.py --neck FastIndoorImVoxelNeck --batch $b --precision bf16 --layers --graph 2>&1 | tail -19
done
