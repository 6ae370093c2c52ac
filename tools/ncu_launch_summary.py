"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (stdin or file argument):
launch count, total / average duration and share of the listed launches."""
import csv
import re
import sys
from collections import defaultdict

src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
rows = [r for r in csv.reader(l for l in src if l.startswith('"'))]
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if r[0] == "ID" or len(r) <= iv:
        continue
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}.get(r[iu], 1.0)
    name = re.sub(r"\(.*", "", r[ik])[:60]
    tot[name][0] += 1
    tot[name][1] += v
all_us = sum(v[1] for v in tot.values()) or 1.0
for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:62s} n={n:5d} total={us / 1e3:9.3f} ms avg={us / n:9.2f} us share={100 * us / all_us:5.1f}%")
