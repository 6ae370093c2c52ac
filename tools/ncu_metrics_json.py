"""profiles/r02_ncu_metrics.json from `ncu --page raw --csv` dumps of `--set full` captures: per (workload, kernel) the
issue-slot utilisation, threads per instruction and DRAM bytes per ray that bench.py puts beside its rooflines.
usage: ncu_metrics_json.py out.json workload=file.csv:rays_per_launch ...   (rays_per_launch: queue length of the captured
launch, from RtStats - 16777216 for a full-pool iteration; 0 = unknown -> no dram_bytes_per_ray)
   or: ncu_metrics_json.py out.json workload=file.csv:stats.json ...     (a capture of EVERY launch of one render,
tools/ncu_capture_all.sh: DRAM bytes summed over the launches of a kernel / the segments the render put through it, from the
instrumented pass of the same render; the other metrics are duration-weighted means over the launches)"""
import csv, json, re, sys

out_path, args = sys.argv[1], sys.argv[2:]
res = {}
for a in args:
    wl, rest = a.split("=")
    path, rays = rest.split(":")
    whole = None                       # whole-render capture: segment counts of that render
    if rays.endswith(".json"):
        whole = json.load(open(rays))
        rays = 0.0
    rays = float(rays)
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        if name not in idx or r[idx[name]] in ("", "n/a"):
            return None
        v = float(r[idx[name]].replace(",", ""))
        return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9}.get(units[idx[name]], 1.0)

    per = {}
    for r in rows[2:]:
        name = re.sub(r"<.*", "", r[idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("rtb::", ""))
        per.setdefault(name, []).append(r)
    res[wl] = {}
    for name, rs in per.items():
        dur = [val(r, "gpu__time_duration.sum") or 0.0 for r in rs]
        def mean(k):                   # duration-weighted over the captured launches (one launch: the value itself)
            pairs = [(val(r, k), d) for r, d in zip(rs, dur) if val(r, k) is not None]
            if not pairs:
                return None
            if k == "gpu__time_duration.sum" or sum(d for _, d in pairs) <= 0:
                return sum(v for v, _ in pairs) / len(pairs)
            return sum(v * d for v, d in pairs) / sum(d for _, d in pairs)
        dram = [(val(r, "dram__bytes_read.sum") or 0) + (val(r, "dram__bytes_write.sum") or 0) for r in rs]
        seg = None
        if whole:                      # segments that went through this kernel in the captured render
            it_rays = whole["rays"] - whole["rays_finish"]
            seg = {"k_traverse_p": whole["rays_counted"] or None, "k_shade": whole["rays_counted"] or None, "k_trace_flat": it_rays,
                   "k_finish": whole["rays_finish"] or None, "k_finish_flat": whole["rays_finish"] or None}.get(name)
        res[wl][name] = {
            "issue_active_pct": mean("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "threads_per_inst": mean("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "fp32_pipe_pct": mean("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
            "l1_hit_pct": mean("l1tex__t_sector_hit_rate.pct"), "l1_throughput_pct": mean("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
            "l2_bytes": mean("lts__t_bytes.sum"), "duration_us": (mean("gpu__time_duration.sum") or 0) * 1e6,
            "registers": mean("launch__registers_per_thread"), "warps_active_pct": mean("sm__warps_active.avg.pct_of_peak_sustained_active"),
            "dram_bytes_per_launch": sum(dram) / len(dram),
            "dram_bytes_per_ray": (sum(dram) / seg) if seg else ((sum(dram) / len(dram) / rays) if rays > 0 else None),
            "dram_read_bytes_per_ray": (sum((val(r, "dram__bytes_read.sum") or 0) for r in rs) / seg) if seg else None,
            "dram_write_bytes_per_ray": (sum((val(r, "dram__bytes_write.sum") or 0) for r in rs) / seg) if seg else None,
            "segments": seg, "total_us": sum(dur) * 1e6,
            "launches_captured": len(rs), "source": path.split("/")[-1],
        }
json.dump(res, open(out_path, "w"), indent=1)
print(json.dumps(res, indent=1))
