cd $GRAFT_REPO_ROOT
for nar in 0 1; do for w in nuscenes nuscenes_refgrid; do
 echo -n "narrow=$nar $w: "; IVX_BP_NARROW=$nar timeout 300 python bench.py --workload $w --no-neck --no-extras --steps 20 --warmup 5 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline_backproject']; print(round(r['kernel_ms'],4), round(r['achieved'],1), round(r['frac'],3))"
done; done
timeout 600 python -m pytest tests/test_backproject_gpu.py -x -q 2>&1 | tail -2
