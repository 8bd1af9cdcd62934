cd $GRAFT_REPO_ROOT
for n in FastIndoorImVoxelNeck:1 FastIndoorImVoxelNeck:8 KittiImVoxelNeck:8 NuScenesImVoxelNeck:4 ImVoxelNeck:1 ImVoxelNeck:32; do
  echo -n "cost ${n} bf16: "; timeout 200 python tools/neck_bench.py --neck ${n%%:*} --batch ${n##*:} --precision bf16 --graph 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['graph_ms'],3), round(d['ms'],3))"
done
echo -n "score Kitti 8: "; IVX_HALO_PLAN=score timeout 200 python tools/neck_bench.py --neck KittiImVoxelNeck --batch 8 --precision bf16 --graph 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['graph_ms'],3), round(d['ms'],3))"
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
