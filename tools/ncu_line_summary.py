"""Per-CUDA-source-line totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` (stdin):
warp instructions, SIMD efficiency and stall samples per source line, hottest first.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass | python tools/ncu_line_summary.py [N]"""
import csv
import sys
from collections import defaultdict

top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = csv.reader(sys.stdin)
fname, idx, cur_line, cur_src = None, None, None, ""
acc = defaultdict(lambda: [0, 0, 0, defaultdict(int), ""])   # (file, line) -> [samples, warp inst, thread inst, stalls, text]
for r in rows:
    if not r:
        continue
    if r[0] == "File Name":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        idx = {}
        for i, h in enumerate(r):
            idx.setdefault(h, i)
        continue
    if idx is None:
        continue
    if r[0]:                      # a CUDA source line starts a group; SASS rows follow with an empty first column
        cur_line, cur_src = r[0], r[1].strip()
    addr = r[idx["Address"]] if len(r) > idx["Address"] else ""
    if not addr.startswith("0x"):
        continue
    a = acc[(fname, cur_line)]
    a[0] += int(r[idx["# Samples"]] or 0)
    a[1] += int(r[idx["Instructions Executed"]] or 0)
    a[2] += int(r[idx["Thread Instructions Executed"]] or 0)
    a[4] = cur_src
    for h, i in idx.items():
        if h.startswith("stall_") and "Not Issued" not in h and i < len(r) and r[i]:
            a[3][h[6:]] += int(r[i])
tot_s = sum(a[0] for a in acc.values()) or 1
tot_i = sum(a[1] for a in acc.values()) or 1
print(f"total stall samples {tot_s}, warp instructions {tot_i}")
print(f"{'samples':>8s} {'%':>5s} {'warp inst %':>11s} {'thr/inst':>8s}  top stalls                       source")
for (f, line), a in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
    st = " ".join(f"{k}:{v}" for k, v in sorted(a[3].items(), key=lambda kv: -kv[1])[:3])
    print(f"{a[0]:8d} {100 * a[0] / tot_s:5.1f} {100 * a[1] / tot_i:11.1f} {a[2] / max(a[1], 1):8.1f}  {st:32s} {f}:{line} {a[4][:90]}")
