#!/usr/bin/env python
"""bench.py — headline benchmark of the path-tracing hot path (BASELINE.json metric: Msamples/s).

A "step" = one complete render of the workload (default: BASELINE configs[1], the Cornell box 600x600 at
200 spp, depth 50, with cs_redirectTargets) through librt_b200.so.  With N GPUs the samples per pixel are
split across ranks (strong scaling; scene replicated) and one NCCL reduce assembles the image.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scene NAME] [--impl reference]
  torchrun ... bench.py --gpus N ...            (one rank per GPU)

Prints ONE JSON line on rank 0.  `value` = whole-job Msamples/s with the scene resident in HBM;
`e2e` = the same metric through the public API with host buffers (scene commit + upload, render, image
read-back inside the timed region); `roofline` = the dominant kernel of the workload against the measured HBM peak
with its algorithmic bytes, `roofline_fp32` = the same kernel's SURVEY §8(d) flops against the measured
FP32 peak (the limiter of this path is instruction issue, not bandwidth); `cpu_baseline` = the CPU oracle (a C++ port of the reference's
algorithm; the Haskell reference cannot be built here) on all host cores, on a bounded sample.
`--impl reference` times that CPU oracle as the reference arm.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene key in raytrace_b200.scenes.CONFIGS, ctor kwargs, description)
    "cornell": ("cornell", dict(), "cornell_box_600x600_200spp_depth50_redirect"),
    "cornell_noredirect": ("cornell_noredirect", dict(), "cornell_box_600x600_200spp_depth50_noredirect"),
    "readme": ("readme", dict(), "readme_example_600x338_50spp_depth10"),
    "demo1": ("demo1", dict(), "demo1_1200x675_500spp_depth50"),
    "demo2": ("demo2", dict(), "demo2_800x800_250spp_depth50"),
    "pawn": ("pawn", dict(width=1920, aspect=16 / 9, spp=1024), "pawn_1920x1080_1024spp_depth20"),
    "bunny_volume": ("bunny_volume", dict(), "bunny_volume_1920x1080_1024spp_depth20"),
}
# spp of the timed CPU sample (full resolution): a few seconds on 16 host cores = tens of core-seconds per workload
CPU_SAMPLE_SPP = {"cornell": 200, "cornell_noredirect": 64, "readme": 50, "demo1": 64, "demo2": 32, "pawn": 16, "bunny_volume": 16}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def measure_aux_peaks():
    """FP32 FMA and L2 read peaks of THIS GPU (tools/peaks.cu): denominators MEASURED_PEAKS.json does not hold."""
    exe = os.path.join(ROOT, "tools", "peaks")
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout.strip().splitlines()[-1]
        return json.loads(out)
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)}


# F_shade of SURVEY §8(d): flops per shaded hit are small next to traversal; the per-ray traversal flops below use the
# reference's own test counts (oracle counters): 24 per box test, 24 per sphere test, 64 per planar test.
def flops_per_ray(n_node, n_sph, n_pl):
    return 24 * n_node + 24 * n_sph + 64 * n_pl


# DRAM bytes per k_traverse launch from the committed `ncu --set full` captures (profiles/, per workload; None = not captured)
NCU_TRAFFIC = {
    # (workload, kernel): DRAM bytes PER RAY = (dram__bytes_read.sum + dram__bytes_write.sum) / rays of the captured launch
    # (ncu --set full, full 16 Mi-path launches; profiles/r01h_kernel_table.md, r01h_ncu_full_summary.txt).  The bench line
    # reports traffic = this x the measured mean rays per launch of the run, so it stays "per launch like achieved".
    ("pawn", "k_shade"): (1146.6e6 + 342.0e6) / 16777216, ("pawn", "k_traverse"): (633.9e6 + 106.5e6) / 16777216,
    ("demo1", "k_shade"): (1114.9e6 + 547.8e6) / 16777216, ("demo1", "k_traverse"): (536.9e6 + 107.9e6) / 16777216,
    ("demo2", "k_shade"): (1131.2e6 + 590.7e6) / 16777216, ("demo2", "k_traverse"): (940.2e6 + 112.2e6) / 16777216,
    # fused flat-scene kernel (raygen + closest hit + shading + compaction in one launch): 660.9 MB read + 609.6 MB written
    ("cornell", "k_trace_flat"): (660.9e6 + 609.6e6) / 16777216,
}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


from raytrace_b200.multigpu import split_samples  # noqa: E402  (sample-range sharding)


def cpu_oracle_rate(world_desc, settings, seed, spp, threads=0):
    from oracle.harness import OracleScene
    ora = OracleScene(world_desc)
    _, cnt = ora.render(settings, seed, spp=spp, threads=threads)
    return cnt


def run_reference(args, rank, world_size):
    """Reference arm: the reference's algorithm on the host cores (CPU oracle port; no GHC here)."""
    if rank != 0:
        return
    from raytrace_b200 import scenes
    key, kw, desc = WORKLOADS[args.scene]
    ctor, seed = scenes.CONFIGS[key]
    settings, world = ctor(**kw)
    spp = CPU_SAMPLE_SPP[args.scene]
    from oracle.harness import OracleScene
    ora = OracleScene(world)
    for _ in range(args.warmup):
        ora.render(settings, seed, spp=max(1, spp // 8))
    secs, samples, cnt = 0.0, 0, None
    for k in range(args.steps):
        _, cnt = ora.render(settings, seed + k, spp=spp)
        secs += cnt["seconds"]
        samples += cnt["samples"]
    value = samples / secs / 1e6
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "note": "CPU oracle (C++ port of the Haskell reference; GHC absent), all host threads"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cnt["threads"], "kind": "port",
                         "sample": f"{spp} of {settings.cs_samplesPerPixel} spp per step, full resolution"},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mrays_per_s": cnt["rays"] / cnt["seconds"] / 1e6,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--scene", default="cornell", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world_size)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from raytrace_b200 import scenes, mkStdGen
    from raytrace_b200.ray import CompiledScene, imageHeight
    from raytrace_b200.multigpu import ImageReducer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world_size > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    key, kw, desc = WORKLOADS[args.scene]
    ctor, seed = scenes.CONFIGS[key]
    settings, world = ctor(**kw)
    W, H, spp = settings.cs_imageWidth, imageHeight(settings), settings.cs_samplesPerPixel
    first, count = split_samples(spp, world_size, rank)
    total_samples = W * H * spp

    cs = CompiledScene(world, device=local_rank)           # commit + upload once: scene resident in HBM
    stream = torch.cuda.current_stream(dev)
    acc = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
    out_host = torch.empty(H * W * 3, dtype=torch.float32).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # sample-sharded runs: accumulators in symmetric memory, image reduce by the root's own NVLS / peer-load kernel
    # (raytrace_b200.multigpu.ImageReducer -> rt_reduce_image); RT_B200_REDUCE=nccl selects dist.reduce instead
    reducer = ImageReducer(H * W * 3, dev) if distributed else None
    result = {"img": acc}

    def one_step(step_seed, want_stats=True):
        """device-resident step: render this rank's sample range, reduce to rank 0, normalise."""
        if distributed:
            a = reducer.acc()
            a.zero_()
            st = cs.render_device(settings, mkStdGen(step_seed), first, count, a.data_ptr(), 0, stream.cuda_stream, want_stats=want_stats)
            img = reducer.reduce(1.0 / spp)
            if rank == 0:
                result["img"] = img
            return st
        acc.zero_()
        st = cs.render_device(settings, mkStdGen(step_seed), first, count, acc.data_ptr(), 0, stream.cuda_stream)
        acc.mul_(1.0 / spp)
        return st

    for k in range(args.warmup):
        one_step(seed + 1000 + k)
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    step_ms, stats = [], []
    barrier()
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        flush.fill_(k & 0xFF)                                # flush L2 between timed iterations (outside the events)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if distributed:
            dist.barrier()
        e0.record(stream)
        # sharded runs ask for no statistics inside the timed step: nothing is read back, and a share that fits the
        # pool returns without a host synchronisation, so the reduce queues right behind the render
        st = one_step(seed + k, want_stats=not distributed)
        e1.record(stream)
        e1.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        stats.append(st)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    if distributed:   # ray counts and per-kernel times of the same steps, from an instrumented (untimed) re-run
        stats = [one_step(seed + k, want_stats=True) for k in range(args.steps)]
        barrier()

    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if distributed:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)      # max over ranks
    total_ms = float(total_ms.item())
    lum = float(result["img"].mean().item()) if rank == 0 else 0.0

    # ---- e2e: public API with HOST buffers; scene commit + upload + render + image read-back per step ----
    e2e_ms = []
    h2d = 0
    n_e2e_warm = 2                                           # first calls allocate the workspace's image buffers and the page-locked result
    for k in range(n_e2e_warm + max(1, min(args.steps, 5))):
        barrier()
        t0 = time.perf_counter()
        cs2 = CompiledScene(world, device=local_rank)       # host description -> flatten -> H2D
        h2d = cs2.info()["device_bytes"]
        if not distributed:
            img = cs2.render(settings, mkStdGen(seed + 77 + k))   # rt_render: D2H of the normalised image inside
        else:
            a = reducer.acc()
            a.zero_()
            cs2.render_device(settings, mkStdGen(seed + 77 + k), first, count, a.data_ptr(), 0, stream.cuda_stream, want_stats=False)
            img = reducer.reduce(1.0 / spp)
            if rank == 0:
                out_host.copy_(img, non_blocking=True)
            torch.cuda.synchronize(dev)
        dt = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
        if distributed:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if k >= n_e2e_warm:
            e2e_ms.append(float(dt.item()))
        del cs2

    # ---- the collective alone (sample-sharded runs): one reduce(SUM) of the fp32 image to rank 0, device-timed ----
    reduce_ms = reduce_nccl_ms = None
    if distributed:
        def timed(fn):
            ts = []
            for k in range(7):
                dist.barrier()
                r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                r0.record(stream)
                fn()
                r1.record(stream)
                r1.synchronize()
                ts.append(r0.elapsed_time(r1))
            t = torch.tensor([statistics.median(ts[2:])], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        reduce_ms = timed(lambda: reducer.reduce(1.0 / spp))                       # barrier + rt_reduce_image (or dist.reduce)
        reduce_nccl_ms = timed(lambda: dist.reduce(acc, dst=0, op=dist.ReduceOp.SUM))  # the ncclReduce it replaces

    if rank != 0:
        if distributed:
            dist.destroy_process_group()
        return

    ms_per_step = total_ms / args.steps
    value = total_samples / (ms_per_step * 1e-3) / 1e6
    rays = sum(s["rays"] for s in stats)
    trav_ms = sum(s["ms_traverse"] for s in stats)
    shade_ms = sum(s["ms_shade"] for s in stats)
    iters = sum(s["iterations"] for s in stats)
    fin_ms = sum(s["ms_finish"] for s in stats)
    fin_rays = sum(s["rays_finish"] for s in stats)
    fin_paths = sum(s["paths_finish"] for s in stats)
    iter_launches = iters - sum(1 for s in stats if s["paths_finish"])   # the tail kernel counts as one iteration
    iter_rays = rays - fin_rays                                           # segments traced by the iteration kernels
    launches = sum(s["launches"] for s in stats)

    # ---- CPU baseline + the oracle's traversal counters (they define the algorithmic bytes per ray) ----
    cpu = None
    bytes_per_ray = None
    if not args.no_cpu_baseline:
        cspp = CPU_SAMPLE_SPP[args.scene]
        cnt = cpu_oracle_rate(world, settings, seed, cspp)
        n_node, n_sph, n_pl = (cnt[k] / cnt["rays"] for k in ("box_tests", "sphere_tests", "planar_tests"))
        bytes_per_ray = 32 * n_node + 16 * n_sph + 48 * n_pl + 128
        cpu = {"value": cnt["samples"] / cnt["seconds"] / 1e6, "unit": "Msamples/s", "cores": cnt["threads"], "kind": "port",
               "sample": f"{cspp} of {spp} spp, full resolution ({cnt['samples']} samples, {cnt['seconds']:.2f} s)",
               "mrays_per_s": cnt["rays"] / cnt["seconds"] / 1e6, "rays_per_sample": cnt["rays"] / cnt["samples"],
               "tests_per_ray": {"box": n_node, "sphere": n_sph, "planar": n_pl}}
    peak, peak_kind = load_peaks()
    aux = measure_aux_peaks()
    roof = None
    roof_fp32 = None
    kernels = {}
    samples_rendered = count * W * H * args.steps
    # flat scenes run ONE fused kernel per iteration (k_trace_flat: raygen + closest hit + shading + compaction); the
    # library then reports its time as ms_shade and ms_traverse ~ 0
    fused = shade_ms > 0 and trav_ms < 0.05 * shade_ms
    if fused:
        # compulsory HBM traffic of the fused kernel: every segment except a path's first reads its 56-byte record, and
        # every path that continues writes one (new camera paths are generated in registers; no hit buffer)
        # (the paths handed to the tail kernel were written by the last iteration and are read by k_finish_flat)
        fz_bytes = 112.0 * max(iter_rays - samples_rendered, 0) + 56.0 * fin_paths
        ach = fz_bytes / (shade_ms * 1e-3) / 1e9
        kernels["k_trace_flat"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                   "traffic": (NCU_TRAFFIC[(args.scene, "k_trace_flat")] * iter_rays / max(iter_launches, 1)) if (args.scene, "k_trace_flat") in NCU_TRAFFIC else None,
                                   "bytes_per_ray_algorithmic": fz_bytes / max(iter_rays, 1),
                                   "avg_launch_ms": shade_ms / max(iter_launches, 1), "ms_per_step": shade_ms / args.steps,
                                   "algorithmic_bytes_per_launch": fz_bytes / max(iter_launches, 1),
                                   "rays_per_launch": iter_rays / max(iter_launches, 1),
                                   "note": "queue traffic only: the scene list is constant-bank resident.  The kernel's limiter is "
                                           "instruction issue (ncu: 73 % issue-active, 27 of 32 threads per instruction) - see roofline_fp32"}
        if bytes_per_ray is not None and aux.get("fp32_fma_tflops"):
            # traversal flops from the reference's test counts + F_shade (lambertian + cosine sample ~ 40,
            # each redirect target one planar test + 10)
            f_shade = 40 + 74 * len(settings.cs_redirectTargets)
            fl = flops_per_ray(n_node, n_sph, n_pl) + f_shade
            ach = iter_rays * fl / (shade_ms * 1e-3) / 1e12
            roof_fp32 = {"kernel": "k_trace_flat", "achieved": ach, "peak": aux["fp32_fma_tflops"], "unit": "TFLOP/s",
                         "frac": ach / aux["fp32_fma_tflops"], "flops_per_ray_algorithmic": fl,
                         "peak_source": "tools/peaks.cu, measured in this run (FFMA chains)",
                         "l2_read_gbs_measured": aux.get("l2_read_gbs")}
    elif shade_ms > 0:
        # k_shade streams the wavefront state (SURVEY 8(d)'s constant term): per ray segment it reads the 56-byte path
        # record + the 8-byte hit, and writes a 56-byte record for every path that continues (rays - samples of them:
        # each sample ends exactly once).  This is compulsory HBM traffic (the queues exceed L2).
        sh_bytes = 64.0 * iter_rays + 56.0 * (max(iter_rays - samples_rendered, 0) + fin_paths)
        ach = sh_bytes / (shade_ms * 1e-3) / 1e9
        kernels["k_shade"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                              "traffic": (NCU_TRAFFIC[(args.scene, "k_shade")] * iter_rays / max(iter_launches, 1)) if (args.scene, "k_shade") in NCU_TRAFFIC else None, "bytes_per_ray_algorithmic": sh_bytes / max(iter_rays, 1),
                              "avg_launch_ms": shade_ms / max(iter_launches, 1), "ms_per_step": shade_ms / args.steps,
                              "algorithmic_bytes_per_launch": sh_bytes / max(iter_launches, 1)}
    if bytes_per_ray is not None and trav_ms > 0 and not fused:
        achieved = iter_rays * bytes_per_ray / (trav_ms * 1e-3) / 1e9
        kernels["k_traverse"] = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                 "traffic": (NCU_TRAFFIC[(args.scene, "k_traverse")] * iter_rays / max(iter_launches, 1)) if (args.scene, "k_traverse") in NCU_TRAFFIC else None, "bytes_per_ray_algorithmic": bytes_per_ray,
                                 "avg_launch_ms": trav_ms / max(iter_launches, 1), "ms_per_step": trav_ms / args.steps,
                                 "frac_of_measured_l2": achieved / aux["l2_read_gbs"] if aux.get("l2_read_gbs") else None,
                                 "note": "SURVEY 8(d) algorithmic bytes (32*N_node + 16*N_sphere + 48*N_planar + 128, the REFERENCE's "
                                         "per-ray test counts from the oracle).  The scene records are cache / constant-bank resident "
                                         "on the GPU, so this rate of algorithmic work exceeds the HBM and L2 peaks by construction; the "
                                         "kernel's real limiter is instruction issue — see roofline_fp32 and profiles/"}
        if aux.get("fp32_fma_tflops"):
            fl = flops_per_ray(n_node, n_sph, n_pl)
            ach = iter_rays * fl / (trav_ms * 1e-3) / 1e12
            roof_fp32 = {"kernel": "k_traverse", "achieved": ach, "peak": aux["fp32_fma_tflops"], "unit": "TFLOP/s",
                         "frac": ach / aux["fp32_fma_tflops"], "flops_per_ray_algorithmic": fl,
                         "peak_source": "tools/peaks.cu, measured in this run (FFMA chains)",
                         "l2_read_gbs_measured": aux.get("l2_read_gbs")}
    if fin_paths and fin_ms > 0:
        # the tail kernel: each lane carries a path to its end; HBM traffic is one 56-byte record per path it takes over
        # (plus the accumulator atomics), so it is latency / issue bound by construction - reported for its share of the step
        kernels["k_finish"] = {"bound": "hbm", "achieved": 56.0 * fin_paths / (fin_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                               "frac": 56.0 * fin_paths / (fin_ms * 1e-3) / 1e9 / peak, "traffic": None,
                               "bytes_per_ray_algorithmic": 56.0 * fin_paths / max(fin_rays, 1), "avg_launch_ms": fin_ms / args.steps,
                               "ms_per_step": fin_ms / args.steps, "rays_per_launch": fin_rays / args.steps,
                               "mrays_per_s": fin_rays / (fin_ms * 1e-3) / 1e6,
                               "note": "tail kernel (k_finish_flat / k_finish): no queue traffic after the first read; issue / latency bound"}
    if kernels:
        # the dominant kernel = the one with the larger share of the step
        dom = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
        roof = dict(kernels[dom], kernel=dom, peak_source=f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})",
                    other_kernels={k: v for k, v in kernels.items() if k != dom})
        roof.setdefault("rays_per_launch", iter_rays / max(iter_launches, 1))
    e2e_val = total_samples / (statistics.median(e2e_ms) * 1e-3) / 1e6
    line = {
        "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world_size, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "image": [W, H], "spp": spp, "max_depth": settings.cs_maxRecursionDepth,
                   "parallelism": f"sample-sharded x{world_size}" if distributed else "single GPU",
                   "l2": "256 MiB flush between timed steps", "pool": stats[-1]["pool"]},
        "mrays_per_s": rays / (total_ms * 1e-3) / 1e6 * (world_size if distributed else 1),
        "rays_per_sample": rays / (count * W * H * args.steps),
        "kernel_ms_per_step": {"traverse": trav_ms / args.steps, "shade": shade_ms / args.steps, "finish": fin_ms / args.steps,
                               "iterations": iters / args.steps, "reduce": reduce_ms, "reduce_nccl": reduce_nccl_ms,
                               "reduce_mode": (reducer.mode + (" (" + reducer.why + ")" if reducer.why else "")) if reducer else None},
        "gpu_launches": int(launches),
        "e2e": {"value": e2e_val, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": W * H * 3 * 4,
                "ms_per_step": statistics.median(e2e_ms)},
        "roofline": roof, "roofline_fp32": roof_fp32, "cpu_baseline": cpu, "clocks": clocks,
        "mean_radiance": lum, "wall_s": t_wall,
    }
    print(json.dumps(line), flush=True)
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
