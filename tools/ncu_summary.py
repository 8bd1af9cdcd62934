#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into a small text table.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--json out.json]

Prints, per captured launch, the metrics DESIGN.md and bench.py quote:
duration, DRAM bytes read/written, DRAM and L2 throughput, hit rates,
occupancy, registers, instruction counts and the top stall reasons.
"""
import csv
import json
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_warps",
    "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
]


def to_bytes(value, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value.replace(",", "")) * mult.get(unit, 1)


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(out.stdout.splitlines()))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    launches = []
    for r in rows[2:]:
        d = {"kernel": r[col["Kernel Name"]]}
        for w in WANT:
            if w in col:
                d[w] = f"{r[col[w]]} {units[col[w]]}".strip()
        stalls = []
        for h, i in col.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        d["top_stalls(warps per issue)"] = ", ".join(f"{n}={v:.1f}" for v, n in stalls[:5])
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        d["dram_bytes_per_launch"] = rd + wr
        launches.append(d)
    for d in launches:
        print("----")
        for k, v in d.items():
            print(f"{k:70s} {v}")
    if "--json" in sys.argv:
        path = sys.argv[sys.argv.index("--json") + 1]
        mean = sum(d["dram_bytes_per_launch"] for d in launches) / len(launches)
        with open(path, "w") as f:
            json.dump({"source": rep, "kernel": launches[0]["kernel"], "launches": len(launches),
                       "dram_bytes_per_launch": mean}, f, indent=1)


if __name__ == "__main__":
    main()
