"""profiles/r02_ncu_metrics.json from `ncu --page raw --csv` exports of the fused kernel: the per-launch counters that
bench.py quotes by name (roofline.traffic, roofline_issue) instead of hard-coding them.

    ncu -i b8.ncu-rep --page raw --csv > b8.csv ; ncu -i b256.ncu-rep --page raw --csv > b256.csv
    python tools/ncu_extract.py b8.csv [b256.csv] > profiles/r02_ncu_metrics.json
"""
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (for the workload string only)

WANT = {"dram__bytes_read.sum": "dram_bytes_read", "dram__bytes_write.sum": "dram_bytes_write",
        "smsp__inst_executed.sum": "warp_instructions", "gpu__time_duration.sum": "duration_under_ncu",
        "sm__cycles_elapsed.avg.per_second": "sm_clock", "launch__grid_size": "grid", "launch__cluster_size": "cluster",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
        "smsp__warps_eligible.avg.per_cycle_active": "eligible_warps_per_cycle",
        "sm__icc_request_hit_rate.pct": "icache_hit_pct", "launch__registers_per_thread": "registers"}
SCALE = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}


def read(path):
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    out = {}
    for i, name in enumerate(hdr):
        if name in WANT:
            try:
                v = float(vals[i].replace(",", ""))
            except ValueError:
                continue
            out[WANT[name]] = v * SCALE.get(units[i], 1.0)
    return out


a = read(sys.argv[1])
doc = {"workload": bench.WORKLOAD, "kernel": "proposal_fused_kernel", "source": "ncu --set full --clock-control none, one launch",
       "dram_bytes_read": a.get("dram_bytes_read"), "dram_bytes_write": a.get("dram_bytes_write"),
       "warp_instructions": a.get("warp_instructions"), "batch_8": a}
if len(sys.argv) > 2:
    b = read(sys.argv[2])
    doc["batch_256"] = b
    doc["warp_instructions_saturating"] = b.get("warp_instructions")
print(json.dumps(doc, indent=1))
