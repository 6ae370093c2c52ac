"""Summarise an `ncu --page source --csv` dump: stall reasons, SIMD efficiency, hottest SASS lines.
usage: ncu -i X.ncu-rep --page source --csv [--kernel-name regex:...] | python tools/ncu_source_summary.py"""
import csv
import sys

rows = list(csv.reader(sys.stdin))
kernel, hdr, body = None, None, []
sections = []
for r in rows:
    if r and r[0] == "Kernel Name":
        if hdr is not None:
            sections.append((kernel, hdr, body))
        kernel, hdr, body = r[1], None, []
    elif r and r[0] == "Address":
        hdr = r
    elif r and r[0].startswith("0x") and hdr is not None:
        body.append(r)
if hdr is not None:
    sections.append((kernel, hdr, body))
for kernel, hdr, body in sections[:int(sys.argv[1]) if len(sys.argv) > 1 else 1]:
    idx = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = {s: sum(int(r[idx[s]] or 0) for r in body) for s in stalls}
    n = sum(int(r[idx["# Samples"]] or 0) for r in body)
    ie = sum(int(r[idx["Instructions Executed"]] or 0) for r in body)
    te = sum(int(r[idx["Thread Instructions Executed"]] or 0) for r in body)
    print(f"kernel: {kernel}")
    print(f"  SASS lines {len(body)}, warp instructions {ie}, thread instructions {te}, "
          f"avg active threads/instruction {te / max(ie, 1):.2f} of 32")
    print(f"  stall samples {n}:")
    for s in sorted(tot, key=tot.get, reverse=True)[:8]:
        print(f"    {s:26s}{tot[s]:9d} {tot[s] / max(n, 1) * 100:5.1f} %")
    print("  hottest SASS (by samples):")
    for r in sorted(body, key=lambda r: -int(r[idx["# Samples"]] or 0))[:14]:
        print(f"    {int(r[idx['# Samples']]):7d}  thr/inst {r[idx['Avg. Threads Executed']]:>5s}  {r[idx['Source']].strip()[:90]}")
