"""Per-phase view of an `ncu --page source --csv` export: the SASS of one kernel is cut at its barrier instructions
(BAR.SYNC / BAR.RED / UCGABAR_WAIT) and executed warp instructions + stall samples are summed per segment, with the
source lines (nvdisasm -g) that contribute most.

    python tools/ncu_by_segment.py src.csv lines.txt <mangled kernel name>
"""
import csv
import re
import sys
from collections import defaultdict

src_csv, lines_txt, kname = sys.argv[1:4]
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
segs = []
cur_seg = {"start": 0, "inst": 0, "samples": 0, "lines": defaultdict(lambda: [0, 0]), "end_op": ""}
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    off = int(r[0], 16) - base
    sass = r[1].strip()
    s = int(r[ci["# Samples"]] or 0)
    n = int(r[ci["Instructions Executed"]] or 0)
    cur_seg["inst"] += n
    cur_seg["samples"] += s
    key = off2line.get(off, ("?", 0))
    cur_seg["lines"][key][0] += n
    cur_seg["lines"][key][1] += s
    if re.search(r"\b(BAR\.SYNC|BAR\.RED|UCGABAR_WAIT)", sass) or sass.startswith("RET") or sass.startswith("EXIT"):
        cur_seg["end_op"] = "%s @%x (%s:%d)" % (sass.split()[0], off, key[0], key[1])
        segs.append(cur_seg)
        cur_seg = {"start": off + 16, "inst": 0, "samples": 0, "lines": defaultdict(lambda: [0, 0]), "end_op": ""}
segs.append(cur_seg)
ti = sum(s["inst"] for s in segs)
ts = sum(s["samples"] for s in segs)
for s in segs:
    if s["inst"] == 0 and s["samples"] == 0:
        continue
    print("seg @%05x  inst %8d (%5.2f%%)  samples %4d (%5.2f%%)  ends: %s" % (s["start"], s["inst"], 100.0 * s["inst"] / ti,
                                                                        s["samples"], 100.0 * s["samples"] / ts, s["end_op"]))
    for k, (n, sm) in sorted(s["lines"].items(), key=lambda kv: -kv[1][0])[:4]:
        print("        %-28s inst %8d samples %4d" % ("%s:%d" % k, n, sm))
