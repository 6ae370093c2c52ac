"""Per-kernel evidence table (north_star): achieved HBM and L2 GB/s, FP32 pipe utilisation, warp execution efficiency,
issue utilisation and occupancy, from `ncu --page raw --csv` dumps.  usage: ncu_kernel_table.py label=file.csv ..."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
HBM = peaks.get("hbm_gbs", 6650.0)
L2 = 16050.0      # tools/peaks.cu on this pool's B200 (profiles/r01c_peaks.json)
FP32 = 72.4


def num(s):
    return float(s.replace(",", "")) if s not in ("", "n/a") else float("nan")


print("| workload | kernel | µs | HBM GB/s (of %.0f) | L2 GB/s (of %.0f) | FP32 pipe %% | threads/inst (of 32) | issue-active %% | warps active %% | regs |" % (HBM, L2))
print("|---|---|---|---|---|---|---|---|---|---|")
for arg in sys.argv[1:]:
    label, path = arg.split("=")
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, name, want_unit=None):
        v, u = num(r[idx[name]]), units[idx[name]]
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(u, 1.0)
        return v * scale

    for r in rows[2:]:
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        t = val(r, "gpu__time_duration.sum")
        dram = (val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")) / t / 1e9
        # L2 traffic: sectors x 32 B (lts__t_sectors.sum), else derived from the L1 miss sectors
        if "lts__t_sectors.sum" in idx:
            l2 = num(r[idx["lts__t_sectors.sum"]]) * 32 / t / 1e9
        else:
            l2 = num(r[idx["SM_B.TriageCompute.l1tex__t_sectors_lookup_miss.sum"]]) * 32 / t / 1e9
        print(f"| {label} | `{name}` | {t * 1e6:.0f} | {dram:.0f} ({dram / HBM * 100:.0f} %) | {l2:.0f} ({l2 / L2 * 100:.0f} %) | "
              f"{num(r[idx['sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active']]):.0f} | "
              f"{num(r[idx['smsp__thread_inst_executed_per_inst_executed.ratio']]):.1f} | "
              f"{num(r[idx['smsp__issue_active.avg.pct_of_peak_sustained_active']]):.0f} | "
              f"{num(r[idx['sm__warps_active.avg.pct_of_peak_sustained_active']]):.0f} | {r[idx['launch__registers_per_thread']]} |")
