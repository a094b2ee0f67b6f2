"""Markdown summary of one `ncu --set full` capture: key raw metrics + warp-stall table (all samples) from the source page.

    ncu -i x.ncu-rep --page raw --csv > raw.csv ; ncu -i x.ncu-rep --page source --csv > src.csv
    python tools/ncu_summary.py raw.csv src.csv "title"
"""
import csv
import sys
from collections import defaultdict

raw, src, title = sys.argv[1:4]
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
m = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "duration under ncu (cold, serialised)"), ("launch__grid_size", "grid (CTAs)"),
        ("launch__cluster_size", "cluster size"), ("launch__registers_per_thread", "registers / thread"),
        ("launch__shared_mem_per_block_dynamic", "dynamic shared memory / CTA"),
        ("launch__shared_mem_per_block_static", "static shared memory / CTA"),
        ("dram__bytes_read.sum", "DRAM bytes read"), ("dram__bytes_write.sum", "DRAM bytes written"),
        ("smsp__inst_executed.sum", "warp instructions executed"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy (% of active cycles)"),
        ("smsp__warps_eligible.avg.per_cycle_active", "eligible warps / scheduler / cycle"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy (%)"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput (% of peak, elapsed)"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "FP64 pipe active (% elapsed)"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe (% of peak while active)"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe (% of peak while active)"),
        ("lts__t_sector_hit_rate.pct", "L2 hit rate (%)"), ("sm__icc_request_hit_rate.pct", "instruction cache hit rate (%)"),
        ("sm__cycles_elapsed.avg.per_second", "SM clock")]
print("## %s\n" % title)
print("| metric | value |\n|---|---|")
for k, name in want:
    if k in m:
        print("| %s | %s %s |" % (name, m[k][0], m[k][1]))
rows = list(csv.reader(open(src)))
h2 = rows[1]
ci = {n: i for i, n in enumerate(h2)}
st = defaultdict(int)
tot = 0
for r in rows[2:]:
    if len(r) < len(h2):
        continue
    for n in h2:
        if n.startswith("stall_") and "Not Issued" not in n:
            v = int(r[ci[n]] or 0)
            st[n[6:]] += v
    tot += int(r[ci["# Samples"]] or 0)
print("\nWarp-stall samples (all samples, %d): " % tot + ", ".join("%s %.1f %%" % (k, 100.0 * v / max(tot, 1))
      for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v * 200 > tot))
