"""Join an `ncu --page source --csv` export (SASS view) of one kernel with `nvdisasm -g` line info of the same
function and print warp-stall samples / executed warp instructions per source line.

    ncu -i rep.ncu-rep --page source --csv > src.csv
    cuobjdump -xelf all lib.so ; nvdisasm -g -c proposal.sm_100a.cubin > lines.txt
    python tools/ncu_by_line.py src.csv lines.txt <mangled kernel name> [top]
"""
import csv
import re
import sys
from collections import defaultdict

src_csv, lines_txt, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# address offset -> (file, line)
off2line = {}
cur = None
inside = False
for ln in open(lines_txt):
    if ln.startswith(".text."):
        inside = ln.strip() == ".text.%s:" % kname
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/", ln)
    if m:
        off2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ci = {n: i for i, n in enumerate(hdr)}
base = int(rows[2][0], 16)
agg = defaultdict(lambda: [0, 0, defaultdict(int)])
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
tot_s = tot_i = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    off = int(r[0], 16) - base
    key = off2line.get(off, ("?", 0))
    s = int(r[ci["# Samples"]] or 0)
    n = int(r[ci["Instructions Executed"]] or 0)
    agg[key][0] += s
    agg[key][1] += n
    for c in stall_cols:
        v = int(r[ci[c]] or 0)
        if v:
            agg[key][2][c] += v
    tot_s += s
    tot_i += n
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
print("%-28s %8s %6s %10s %6s  top stalls" % ("file:line", "samples", "%", "warp-inst", "%"))
for key, (s, n, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    tops = ", ".join("%s %d" % (k[6:], v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print("%-28s %8d %6.2f %10d %6.2f  %s" % ("%s:%d" % key, s, 100.0 * s / max(tot_s, 1), n, 100.0 * n / max(tot_i, 1), tops))
