#!/bin/bash
# Run bench.py on every workload and print a one-line digest of each (GPU box).
for w in kitti scannet_fast sunrgbd_fast nuscenes total_sunrgbd; do
  python bench.py --workload $w "$@" > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || { echo "$w FAILED"; tail -5 gpurun_out/bench_$w.err; continue; }
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_$w.json"))
r=d["roofline"]
print("$w", "ms/step %.4f"%d["ms_per_step"], "eager %.4f"%d["timing"]["eager_ms_per_step"], "value %.3g"%d["value"], "GB/s %.0f frac %.3f"%(r["achieved"], r["frac"]), "e2e %.3g"%d["e2e"]["value"], "cpu %.3g"%d["cpu_baseline"]["value"], d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
done
