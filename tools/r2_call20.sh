cd $GRAFT_REPO_ROOT
q() { python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline_backproject']; print(round(r['kernel_ms'],4), round(r['achieved'],1), round(r['frac'],3))"; }
for nar in 1 2; do for w in nuscenes nuscenes_refgrid; do
 echo -n "narrow=$nar $w: "; IVX_BP_NARROW=$nar timeout 300 python bench.py --workload $w --no-neck --no-extras --steps 20 --warmup 5 2>/dev/null | q
done; done
for nar in 0 1; do
 echo -n "narrow64=$nar scannet_fast: "; IVX_BP_NARROW64=$nar timeout 300 python bench.py --workload scannet_fast --no-neck --no-extras --steps 20 --warmup 5 2>/dev/null | q
 echo -n "narrow64=$nar sunrgbd_fast: "; IVX_BP_NARROW64=$nar timeout 300 python bench.py --workload sunrgbd_fast --no-neck --no-extras --steps 20 --warmup 5 2>/dev/null | q
done
IVX_BP_NARROW=2 IVX_BP_NARROW64=1 timeout 600 python -m pytest tests/test_backproject_gpu.py -x -q 2>&1 | tail -2
