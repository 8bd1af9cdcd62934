#!/bin/bash
# Back-projection with fp32 and bf16 channels-last feature maps, per workload:
#   default        bench.py --no-neck (fp32 volume, the parity dtype)
#   STEP=1         the full step (bf16 volume into the bf16 neck): the stage time inside the step
# Run on the GPU box:  [STEP=1] bash tools/bp_dtype_compare.sh [workload ...]
cd "$(dirname "$0")/.."
extra="--no-neck"; [ -n "$STEP" ] && extra=""
for w in ${@:-scannet_fast kitti nuscenes_refgrid sunrgbd_fast}; do
  for d in f32 bf16; do
    python bench.py --workload $w $extra --feature-dtype $d --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); s=d['stages']['backproject']
print('$w maps $d: volume f32 %.4f ms, volume %s %.4f ms' % (s['ms'], s.get('volume_dtype_inside_step'), s.get('ms_inside_step', float('nan'))))"
  done
done
