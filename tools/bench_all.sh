#!/bin/bash
# Run bench.py on every workload and print a one-line digest of each (GPU box).
for w in kitti scannet_fast sunrgbd_fast nuscenes nuscenes_refgrid total_sunrgbd; do
  python bench.py --workload $w "$@" > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || { echo "$w FAILED"; tail -5 gpurun_out/bench_$w.err; continue; }
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_$w.json"))
r=d["roofline_backproject"]; st=d["stages"]
nk=st.get("neck")
print("$w", "step %.3f ms"%d["ms_per_step"], "| bp %.4f ms %.0f GB/s frac %.3f"%(st["backproject"]["ms"], r["achieved"], r["frac"]),
      "| neck " + ("%.2f ms %.0f TF/s frac %.3f"%(nk["ms"], nk["tflops"], d["roofline"]["frac"]) if nk else "-"),
      "| scenes/s %.0f e2e %.0f cpu %.3g"%(d["scenes_per_s"], d["e2e"]["scenes_per_s"], d["cpu_baseline"]["scenes_per_s"]), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
done
