cd $GRAFT_REPO_ROOT
for plan in cost; do
 for n in FastIndoorImVoxelNeck:1 FastIndoorImVoxelNeck:2 FastIndoorImVoxelNeck:8 KittiImVoxelNeck:1 KittiImVoxelNeck:8 NuScenesImVoxelNeck:1 NuScenesImVoxelNeck:4 ImVoxelNeck:1 ImVoxelNeck:4 ImVoxelNeck:32; do
  echo -n "$plan ${n} bf16: "; IVX_HALO_PLAN=$plan timeout 200 python tools/neck_bench.py --neck ${n%%:*} --batch ${n##*:} --precision bf16 --graph 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['graph_ms'],3), round(d['ms'],3))"
 done
 for n in KittiImVoxelNeck:8 FastIndoorImVoxelNeck:8 FastIndoorImVoxelNeck:1 ImVoxelNeck:32; do
  echo -n "$plan ${n} fp32: "; IVX_HALO_PLAN=$plan timeout 200 python tools/neck_bench.py --neck ${n%%:*} --batch ${n##*:} --precision fp32 --graph 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['graph_ms'],3), round(d['ms'],3))"
 done
done
echo -n "score ImVoxelNeck:4 bf16: "; IVX_HALO_PLAN=score timeout 200 python tools/neck_bench.py --neck ImVoxelNeck --batch 4 --precision bf16 --graph 2>&1 | tail -1
echo -n "score ImVoxelNeck:32 fp32: "; IVX_HALO_PLAN=score timeout 200 python tools/neck_bench.py --neck ImVoxelNeck --batch 32 --precision fp32 --graph 2>&1 | tail -1
