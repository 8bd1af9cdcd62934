#!/bin/bash
# Every single-GPU workload of bench.py, one JSON line each into gpurun_out/bench_<workload>.json (GPU box).
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
for w in kitti scannet_fast sunrgbd_fast total_sunrgbd nuscenes nuscenes_refgrid; do
  extra="--no-extras"; [ "$w" = kitti ] && extra=""
  timeout 600 python bench.py --workload $w --steps 20 --warmup 5 $extra > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "$w rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
r=d.get('roofline') or {}; rb=d.get('roofline_backproject') or {}
print('  step %.3f ms, bp %.4f ms frac %.3f, neck frac %s, e2e %.3f ms, scenes/s %.0f, cpu %s' % (d['ms_per_step'], rb.get('kernel_ms',0), rb.get('frac',0), round(r.get('frac',0),3) if r.get('bound')=='tensor' else None, d['e2e']['ms_per_step'], d['scenes_per_s'], (d.get('cpu_baseline') or {}).get('scenes_per_s')))
PY
done
