#!/usr/bin/env python
"""bench.py — benchmark of the path-tracing hot path (BASELINE.json metric: Msamples/s and Mrays/s, per scene).

A "step" = one complete render of a workload through librt_b200.so.  The headline workload (`config.workload`) is
BASELINE configs[4], the pawn mesh with its subsurface volume at 1920x1080 / 1024 spp / depth 20 - the largest
single-GPU configuration and the one north_star names for the 8-GPU split; the SAME JSON line carries a `per_scene`
block with every other BASELINE workload (README, Cornell with / without redirect, demo1, demo2, bunny + volume), each
with its own value / e2e / Mrays/s / roofline / cpu_baseline.  With N GPUs (torchrun, one rank per GPU) the samples per
pixel are split across ranks (strong scaling; scene replicated) and one reduce over NVLink assembles the image.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scene NAME] [--impl reference] [--no-per-scene]
  torchrun ... bench.py --gpus N ...            (one rank per GPU)

Prints ONE JSON line on rank 0.
  value        whole-job Msamples/s, scene resident in HBM, device-timed (CUDA events on the launching stream, max over ranks)
  e2e          the same metric through the public API with HOST buffers: scene commit + upload, render, image read-back
  roofline     the dominant kernel of the workload.  Scene data (<= 0.5 MB of nodes + primitives) is cache resident, so the
               HBM roofline counts the wavefront-queue bytes the kernel must stream; `l2` / `fp32` are the bytes and flops the
               GPU traversal really EXECUTES (device counters, RtOptions.count_work) against the L2 and FFMA peaks measured in
               this run (tools/peaks.cu); `issue` is the ncu issue-slot utilisation (profiles/r02_ncu_metrics.json)
  cpu_baseline the CPU oracle (C++ port of the reference's algorithm; the Haskell reference cannot be built here) on all
               host cores, on a bounded sample; its per-ray test counts are the "algorithmic work avoided"
`--impl reference` times that CPU oracle as the reference arm (same default workload).
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
    "pawn": ("pawn", dict(width=1920, aspect=16 / 9, spp=1024), "pawn_1920x1080_1024spp_depth20"),
    "bunny_volume": ("bunny_volume", dict(), "bunny_volume_1920x1080_1024spp_depth20"),
    "demo1": ("demo1", dict(), "demo1_1200x675_500spp_depth50"),
    "demo2": ("demo2", dict(), "demo2_800x800_250spp_depth50"),
    "cornell": ("cornell", dict(), "cornell_box_600x600_200spp_depth50_redirect"),
    "cornell_noredirect": ("cornell_noredirect", dict(), "cornell_box_600x600_200spp_depth50_noredirect"),
    "readme": ("readme", dict(), "readme_example_600x338_50spp_depth10"),
}
DEFAULT_SCENE = "pawn"
# spp of the timed CPU sample (full resolution): a few seconds on 16 host cores = tens of core-seconds per workload
CPU_SAMPLE_SPP = {"cornell": 200, "cornell_noredirect": 64, "readme": 50, "demo1": 64, "demo2": 32, "pawn": 16, "bunny_volume": 16}

# flops of one executed test: node pair = two slab tests (6 FMA + 12 min/max each); sphere = half-b quadratic;
# planar = plane distance + two planar coordinates with the precomputed v x nS / nS x u ("~39"); cuboid slab test
FLOPS = {"pair": 48.0, "sphere": 24.0, "planar": 39.0, "box": 75.0}
# bytes of one executed fetch: node pair 64; primitive record 64; aggregated cuboid = its head record + the 64-byte box record
BYTES = {"pair": 64.0, "sphere": 64.0, "planar": 64.0, "box": 128.0}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def load_ncu_metrics():
    """per (workload, kernel): issue-slot utilisation, threads per instruction, DRAM bytes per ray from the committed ncu
    captures (tools/ncu_metrics_json.py writes the file from `ncu --page raw --csv` dumps; absent entries -> None)"""
    p = os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")      # one full-pool launch per kernel: steady-state issue metrics
    res = {}
    if os.path.exists(p):
        with open(p) as f:
            res = json.load(f)
    # DRAM bytes per segment: `--set full` over EVERY launch of one render per workload (tools/ncu_capture_all.sh), summed and
    # divided by the segments the instrumented pass of the same render counted
    p2 = os.path.join(ROOT, "profiles", "r02z_ncu_metrics.json")
    if os.path.exists(p2):
        with open(p2) as f:
            whole = json.load(f)
        for wl, ks in whole.items():
            for k, m in ks.items():
                e = res.setdefault(wl, {}).setdefault(k, {"issue_active_pct": m.get("issue_active_pct"), "threads_per_inst": m.get("threads_per_inst"),
                                                          "source": m.get("source")})
                if m.get("dram_bytes_per_ray") is not None:
                    e["dram_bytes_per_ray"] = m["dram_bytes_per_ray"]
                    e["traffic_source"] = m.get("source")
    return res


def measure_aux_peaks():
    """FP32 FMA and L2 read peaks of THIS GPU (tools/peaks.cu): denominators MEASURED_PEAKS.json does not hold."""
    exe = os.path.join(ROOT, "tools", "peaks")
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout.strip().splitlines()[-1]
        return json.loads(out)
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)}


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


def build_workload(name):
    from raytrace_b200 import scenes
    key, kw, desc = WORKLOADS[name]
    ctor, seed = scenes.CONFIGS[key]
    settings, world = ctor(**kw)
    return settings, world, seed, desc


def cpu_oracle_sample(world, settings, seed, spp, threads=0):
    from oracle.harness import OracleScene
    ora = OracleScene(world)
    _, cnt = ora.render(settings, seed, spp=spp, threads=threads)
    return cnt


def cpu_baseline_block(name, world, settings, seed):
    cspp = CPU_SAMPLE_SPP[name]
    cnt = cpu_oracle_sample(world, settings, seed, cspp)
    n_node, n_sph, n_pl = (cnt[k] / cnt["rays"] for k in ("box_tests", "sphere_tests", "planar_tests"))
    return {"value": cnt["samples"] / cnt["seconds"] / 1e6, "unit": "Msamples/s", "cores": cnt["threads"], "kind": "port",
            "sample": f"{cspp} of {settings.cs_samplesPerPixel} spp, full resolution ({cnt['samples']} samples, {cnt['seconds']:.2f} s)",
            "mrays_per_s": cnt["rays"] / cnt["seconds"] / 1e6, "rays_per_sample": cnt["rays"] / cnt["samples"],
            # the REFERENCE's per-ray test counts (left-first traversal, `group` scans everything, media re-traverse):
            # SURVEY 8(d)'s algorithmic work, most of which the GPU's SAH / ordered / bounded traversal never executes
            "tests_per_ray": {"box": n_node, "sphere": n_sph, "planar": n_pl},
            "bytes_per_ray_survey_8d": 32 * n_node + 16 * n_sph + 48 * n_pl + 128,
            "flops_per_ray_survey_8d": 24 * n_node + 24 * n_sph + 64 * n_pl}


def arm_config(desc, settings, n_gpus, pool=None):
    """`config` of the JSON line - the same dict in both arms (the reference arm runs ON this arm's config): the workload, and how
    the GPU arm is measured (sample sharding, L2 policy, live-path pool of one GPU's share)."""
    from raytrace_b200.ray import imageHeight
    W, H, spp = settings.cs_imageWidth, imageHeight(settings), settings.cs_samplesPerPixel
    if pool is None:
        items = W * H * ((spp + n_gpus - 1) // n_gpus)
        pool = min(1 << 25, (max(items, 1024) + 1023) // 1024 * 1024)
    return {"workload": desc, "image": [W, H], "spp": spp, "max_depth": settings.cs_maxRecursionDepth,
            "parallelism": f"sample-sharded x{n_gpus}" if n_gpus > 1 else "single GPU",
            "l2": "256 MiB flush between timed steps", "pool": int(pool),
            "why_this_workload": "largest single-GPU configuration of BASELINE.json (configs[4]); the other configs are in per_scene"}


def run_reference(args, rank):
    """Reference arm: the reference's algorithm on the host cores (CPU oracle port; no GHC here)."""
    if rank != 0:
        return
    from oracle.harness import OracleScene
    from raytrace_b200.ray import imageHeight
    settings, world, seed, desc = build_workload(args.scene)
    spp = CPU_SAMPLE_SPP[args.scene]
    ora = OracleScene(world)
    for _ in range(args.warmup):
        ora.render(settings, seed, spp=max(1, spp // 8))
    secs, samples, cnt = 0.0, 0, None
    for k in range(args.steps):
        _, cnt = ora.render(settings, seed + k, spp=spp)
        secs += cnt["seconds"]
        samples += cnt["samples"]
    value = samples / secs / 1e6
    per_scene = {}
    if not args.no_per_scene:
        for name in WORKLOADS:
            if name == args.scene:
                continue
            s2, w2, seed2, desc2 = build_workload(name)
            c2 = cpu_oracle_sample(w2, s2, seed2, CPU_SAMPLE_SPP[name])
            per_scene[name] = {"workload": desc2, "value": c2["samples"] / c2["seconds"] / 1e6, "unit": "Msamples/s",
                               "mrays_per_s": c2["rays"] / c2["seconds"] / 1e6, "rays_per_sample": c2["rays"] / c2["samples"],
                               "sample": f"{CPU_SAMPLE_SPP[name]} of {s2.cs_samplesPerPixel} spp, full resolution, one step"}
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": arm_config(desc, settings, args.gpus),
        "reference_note": "CPU oracle (C++ port of the Haskell reference; GHC absent), all host threads; `config` is the GPU arm's",

        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cnt["threads"], "kind": "port",
                         "sample": f"{spp} of {settings.cs_samplesPerPixel} spp per step, full resolution"},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mrays_per_s": cnt["rays"] / cnt["seconds"] / 1e6,
        "per_scene": per_scene,
    }
    print(json.dumps(line), flush=True)


class Bench:
    """Shared state of one bench process (one rank)."""

    def __init__(self, rank, local_rank, world_size):
        import torch
        self.torch = torch
        self.rank, self.local_rank, self.world = rank, local_rank, world_size
        self.distributed = world_size > 1
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        if self.distributed:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.stream = torch.cuda.current_stream(self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)   # > 126 MB L2
        self.reducers = {}
        self.peak, self.peak_kind = load_peaks()
        self.aux = measure_aux_peaks() if rank == 0 else {}
        self.ncu = load_ncu_metrics()

    def barrier(self):
        if self.distributed:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def reducer(self, n_floats):
        from raytrace_b200.multigpu import ImageReducer
        if n_floats not in self.reducers:
            self.reducers[n_floats] = ImageReducer(n_floats, self.dev)
        return self.reducers[n_floats]

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.distributed:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    # ------------------------------------------------------------------------------------------------------------
    def measure(self, name, steps, warmup, e2e_steps, want_cpu, clocks=False):
        """One workload: device-resident throughput, end-to-end throughput, executed-work counters, rooflines."""
        torch = self.torch
        from raytrace_b200 import mkStdGen
        from raytrace_b200.ray import CompiledScene, imageHeight
        settings, world, seed, desc = build_workload(name)
        W, H, spp = settings.cs_imageWidth, imageHeight(settings), settings.cs_samplesPerPixel
        first, count = split_samples(spp, self.world, self.rank)
        total_samples = W * H * spp
        dev, stream, distributed = self.dev, self.stream, self.distributed

        cs = CompiledScene(world, device=self.local_rank)           # commit + upload once: scene resident in HBM
        acc = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
        out_host = torch.empty(H * W * 3, dtype=torch.float32).pin_memory()
        reducer = self.reducer(H * W * 3) if distributed else None
        result = {"img": acc}

        def one_step(step_seed, want_stats=True):
            """device-resident step: render this rank's sample range, reduce to rank 0, normalise."""
            if distributed:
                a = reducer.acc()
                a.zero_()
                st = cs.render_device(settings, mkStdGen(step_seed), first, count, a.data_ptr(), 0, stream.cuda_stream, want_stats=want_stats)
                img = reducer.reduce(1.0 / spp)
                if self.rank == 0:
                    result["img"] = img
                return st
            acc.zero_()
            st = cs.render_device(settings, mkStdGen(step_seed), first, count, acc.data_ptr(), 0, stream.cuda_stream)
            acc.mul_(1.0 / spp)
            return st

        for k in range(warmup):
            one_step(seed + 1000 + k)
        self.barrier()
        sampler = ClockSampler(self.local_rank) if (clocks and self.rank == 0) else None
        if sampler:
            sampler.start()
        step_ms, stats = [], []
        self.barrier()
        t_wall0 = time.perf_counter()
        for k in range(steps):
            self.flush.fill_(k & 0xFF)                                # flush L2 between timed iterations (outside the events)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if distributed:
                self.dist.barrier()
            e0.record(stream)
            # sharded runs ask for no statistics inside the timed step: nothing is read back, and a share that fits the
            # pool returns without a host synchronisation, so the reduce queues right behind the render
            st = one_step(seed + k, want_stats=not distributed)
            e1.record(stream)
            e1.synchronize()
            step_ms.append(e0.elapsed_time(e1))
            stats.append(st)
        self.barrier()
        t_wall = time.perf_counter() - t_wall0
        clk = sampler.stop() if sampler else None
        if distributed:   # ray counts and per-kernel times of the same steps, from an instrumented (untimed) re-run
            stats = [one_step(seed + k, want_stats=True) for k in range(steps)]
            self.barrier()
        # executed-work counters of the traversal (instrumented kernel variant, untimed; same seeds -> same rays)
        cs.handle.set_options(count_work=1)
        work = one_step(seed, want_stats=True)
        cs.handle.set_options(count_work=None)
        self.barrier()

        total_ms = self.max_over_ranks(sum(step_ms))
        lum = float(result["img"].mean().item()) if self.rank == 0 else 0.0

        # ---- e2e: public API with HOST buffers; scene commit + upload + render + image read-back per step ----
        e2e_ms, h2d = [], 0
        n_e2e_warm = 2                                           # first calls allocate the workspace's image buffers and the page-locked result
        for k in range(n_e2e_warm + e2e_steps):
            self.barrier()
            t0 = time.perf_counter()
            cs2 = CompiledScene(world, device=self.local_rank)       # host description -> flatten -> H2D
            h2d = cs2.info()["device_bytes"]
            if not distributed:
                img = cs2.render(settings, mkStdGen(seed + 77 + k))   # rt_render: D2H of the normalised image inside
            else:
                a = reducer.acc()
                a.zero_()
                cs2.render_device(settings, mkStdGen(seed + 77 + k), first, count, a.data_ptr(), 0, stream.cuda_stream, want_stats=False)
                img = reducer.reduce(1.0 / spp)
                if self.rank == 0:
                    out_host.copy_(img, non_blocking=True)
                torch.cuda.synchronize(dev)
            dt = self.max_over_ranks((time.perf_counter() - t0) * 1e3)
            if k >= n_e2e_warm:
                e2e_ms.append(dt)
            del cs2

        # ---- the collective alone (sample-sharded runs): one reduce(SUM) of the fp32 image to rank 0, device-timed ----
        reduce_ms = reduce_nccl_ms = None
        if distributed:
            def timed(fn):
                ts = []
                for k in range(7):
                    self.dist.barrier()
                    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    r0.record(stream)
                    fn()
                    r1.record(stream)
                    r1.synchronize()
                    ts.append(r0.elapsed_time(r1))
                return self.max_over_ranks(statistics.median(ts[2:]))
            reduce_ms = timed(lambda: reducer.reduce(1.0 / spp))                       # barrier + rt_reduce_image (or dist.reduce)
            reduce_nccl_ms = timed(lambda: self.dist.reduce(acc, dst=0, op=self.dist.ReduceOp.SUM))  # the ncclReduce it replaces
        if self.rank != 0:
            return None

        ms_per_step = total_ms / steps
        value = total_samples / (ms_per_step * 1e-3) / 1e6
        rays = sum(s["rays"] for s in stats)
        trav_ms = sum(s["ms_traverse"] for s in stats)
        shade_ms = sum(s["ms_shade"] for s in stats)
        gen_ms = sum(s["ms_raygen"] for s in stats)
        iters = sum(s["iterations"] for s in stats)
        fin_ms = sum(s["ms_finish"] for s in stats)
        fin_rays = sum(s["rays_finish"] for s in stats)
        fin_paths = sum(s["paths_finish"] for s in stats)
        iter_launches = max(iters - sum(1 for s in stats if s["paths_finish"]), 1)   # the tail kernel counts as one iteration
        iter_rays = rays - fin_rays                                                   # segments traced by the iteration kernels
        launches = sum(s["launches"] for s in stats)
        samples_rendered = count * W * H * steps
        info = cs.info()

        cpu = cpu_baseline_block(name, world, settings, seed) if want_cpu else None
        peak, aux = self.peak, self.aux
        ncu = self.ncu.get(name, {})

        def with_ncu(k, entry):
            m = ncu.get(k)
            if m:
                entry["issue"] = {"issue_active_pct": m.get("issue_active_pct"), "threads_per_inst": m.get("threads_per_inst"),
                                  "lane_issue_frac": (m["issue_active_pct"] / 100.0 * m["threads_per_inst"] / 32.0)
                                  if m.get("issue_active_pct") is not None and m.get("threads_per_inst") is not None else None,
                                  "source": m.get("source")}
                if m.get("dram_bytes_per_ray") is not None:
                    entry["traffic"] = m["dram_bytes_per_ray"] * entry["rays_per_launch"]
                    entry["traffic_source"] = "ncu --set full over every launch of one render: %.1f B per segment (%s)" % (
                        m["dram_bytes_per_ray"], m.get("traffic_source"))
            return entry

        kernels = {}
        fused = shade_ms > 0 and trav_ms < 0.05 * shade_ms           # flat scenes: ONE fused kernel per iteration, reported as ms_shade
        # Camera rays that miss the world box end inside k_raygen: they are path segments (Mrays/s counts them, Ray.hs:178-179)
        # but never reach k_traverse_p / k_shade.  The kernels' bytes and rates are computed from the segments that DO reach
        # them: the instrumented pass of the same step counts those (RtStats.rays_counted).
        q_rays = iter_rays
        if work and work.get("rays_counted") and work["rays"] > work["rays_finish"]:
            q_rays = iter_rays * min(1.0, work["rays_counted"] / float(work["rays"] - work["rays_finish"]))
        if fused:
            # compulsory HBM traffic of the fused kernel: every segment except a path's first reads its 56-byte record, and
            # every path that continues writes one (new camera paths are generated in registers; no hit buffer)
            fz_bytes = 112.0 * max(iter_rays - samples_rendered, 0) + 56.0 * fin_paths
            ach = fz_bytes / (shade_ms * 1e-3) / 1e9
            k = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                 "bytes_per_ray_algorithmic": fz_bytes / max(iter_rays, 1), "avg_launch_ms": shade_ms / iter_launches,
                 "ms_per_step": shade_ms / steps, "algorithmic_bytes_per_launch": fz_bytes / iter_launches,
                 "rays_per_launch": iter_rays / iter_launches,
                 "note": "queue traffic only: the primitive list is constant-bank resident; the limiter is instruction issue"}
            if cpu and aux.get("fp32_fma_tflops"):
                # flops the kernel executes per ray: the list scan (planar tests + cuboid slab tests + sphere tests) + shading
                # (lambertian + cosine sample ~ 40, each redirect target one planar test + 10)
                n_box = 2 if name.startswith("cornell") else 0
                n_planar = info["n_planars"] - 6 * n_box
                fl = FLOPS["planar"] * n_planar + FLOPS["box"] * n_box + FLOPS["sphere"] * info["n_spheres"] + 40 + 49 * len(settings.cs_redirectTargets)
                a32 = iter_rays * fl / (shade_ms * 1e-3) / 1e12
                k["fp32"] = {"achieved": a32, "peak": aux["fp32_fma_tflops"], "unit": "TFLOP/s", "frac": a32 / aux["fp32_fma_tflops"],
                             "flops_per_ray_executed": fl, "peak_source": "tools/peaks.cu, measured in this run (FFMA chains)"}
            kernels["k_trace_flat"] = with_ncu("k_trace_flat", k)
        else:
            if shade_ms > 0:
                # k_shade streams the wavefront state: per ray segment it reads the 56-byte path record + the 8-byte hit and
                # writes a 56-byte record for every path that continues - compulsory HBM traffic (the queues exceed L2)
                # (a culled camera ray is the first segment of its path: the number of continuing paths is iter_rays - samples either way)
                sh_bytes = 64.0 * q_rays + 56.0 * (max(iter_rays - samples_rendered, 0) + fin_paths)
                ach = sh_bytes / (shade_ms * 1e-3) / 1e9
                kernels["k_shade"] = with_ncu("k_shade", {
                    "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                    "bytes_per_ray_algorithmic": sh_bytes / max(q_rays, 1), "avg_launch_ms": shade_ms / iter_launches,
                    "ms_per_step": shade_ms / steps, "algorithmic_bytes_per_launch": sh_bytes / iter_launches,
                    "rays_per_launch": q_rays / iter_launches})
            if trav_ms > 0:
                # k_traverse_p: the HBM bytes it must move are the queue records (32 B origin + direction read, 8 B hit
                # written; media scenes read pixel + sample index as well: + 24 B); nodes and primitives come from L1 / L2
                q_bytes = (40.0 + (24.0 if info["n_media"] else 0.0)) * q_rays
                ach = q_bytes / (trav_ms * 1e-3) / 1e9
                k = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                     "bytes_per_ray_algorithmic": q_bytes / max(q_rays, 1), "avg_launch_ms": trav_ms / iter_launches,
                     "ms_per_step": trav_ms / steps, "algorithmic_bytes_per_launch": q_bytes / iter_launches,
                     "rays_per_launch": q_rays / iter_launches, "mrays_per_s_through_kernel": q_rays / (trav_ms * 1e-3) / 1e6,
                     "note": "HBM roofline = wavefront-queue bytes only; scene data is cache resident, see l2 / fp32 / issue for the binding limits"}
                rc = max(work.get("rays_counted", 0), 1) if work else 1
                if work and work.get("rays_counted"):
                    per_ray = {t: work[f] / rc for t, f in (("pair", "node_pairs"), ("sphere", "sphere_tests"), ("planar", "planar_tests"), ("box", "box_tests"))}
                    ex_bytes = sum(BYTES[t] * per_ray[t] for t in per_ray)
                    ex_flops = sum(FLOPS[t] * per_ray[t] for t in per_ray)
                    k["executed_per_ray"] = dict(per_ray, bytes=ex_bytes, flops=ex_flops)
                    if aux.get("l2_read_gbs"):
                        a2 = q_rays * ex_bytes / (trav_ms * 1e-3) / 1e9
                        k["l2"] = {"achieved": a2, "peak": aux["l2_read_gbs"], "unit": "GB/s", "frac": a2 / aux["l2_read_gbs"],
                                   "note": "bytes of node pairs + primitive records the traversal really fetched (device counters); an "
                                           "upper bound on L2 traffic - ncu L1 hit rate 70-80 %", "peak_source": "tools/peaks.cu (L2-resident float4 reads)"}
                    if aux.get("fp32_fma_tflops"):
                        a32 = q_rays * ex_flops / (trav_ms * 1e-3) / 1e12
                        k["fp32"] = {"achieved": a32, "peak": aux["fp32_fma_tflops"], "unit": "TFLOP/s", "frac": a32 / aux["fp32_fma_tflops"],
                                     "peak_source": "tools/peaks.cu, measured in this run (FFMA chains)"}
                    if cpu:
                        k["algorithmic_work_avoided"] = {
                            "reference_bytes_per_ray": cpu["bytes_per_ray_survey_8d"], "executed_bytes_per_ray": ex_bytes + 40.0,
                            "reference_flops_per_ray": cpu["flops_per_ray_survey_8d"], "executed_flops_per_ray": ex_flops}
                kernels["k_traverse_p"] = with_ncu("k_traverse_p", k)
            if gen_ms > 0:
                kernels["k_raygen+gaps"] = {"ms_per_step": gen_ms / steps,
                                            "note": "k_raygen, k_begin and launch gaps: ms_total minus the event-bracketed kernels"}
        if fin_paths and fin_ms > 0:
            # the tail kernel: each lane carries a path to its end; HBM traffic is one 56-byte record per path it takes over
            kernels["k_finish"] = {"bound": "hbm", "achieved": 56.0 * fin_paths / (fin_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": 56.0 * fin_paths / (fin_ms * 1e-3) / 1e9 / peak, "traffic": None,
                                   "bytes_per_ray_algorithmic": 56.0 * fin_paths / max(fin_rays, 1), "avg_launch_ms": fin_ms / steps,
                                   "ms_per_step": fin_ms / steps, "rays_per_launch": fin_rays / steps,
                                   "mrays_per_s": fin_rays / (fin_ms * 1e-3) / 1e6,
                                   "note": "tail kernel (k_finish_flat / k_finish): no queue traffic after the first read; issue / latency bound"}
        roof = None
        timed_kernels = {k: v for k, v in kernels.items() if "achieved" in v}
        if timed_kernels:
            dom = max(timed_kernels, key=lambda k: timed_kernels[k]["ms_per_step"])     # dominant = largest share of the step
            roof = dict(kernels[dom], kernel=dom, peak_source=f"MEASURED_PEAKS.json hbm_gbs ({self.peak_kind})",
                        other_kernels={k: v for k, v in kernels.items() if k != dom})
        e2e_med = statistics.median(e2e_ms)
        res = {
            "workload": desc, "value": value, "unit": "Msamples/s", "ms_per_step": ms_per_step, "steps": steps, "warmup": warmup,
            "image": [W, H], "spp": spp, "max_depth": settings.cs_maxRecursionDepth, "pool": stats[-1]["pool"],
            "mrays_per_s": rays / (total_ms * 1e-3) / 1e6 * (self.world if distributed else 1),
            "rays_per_sample": rays / (count * W * H * steps),
            "kernel_ms_per_step": {"traverse": trav_ms / steps, "shade": shade_ms / steps, "finish": fin_ms / steps, "raygen_and_gaps": gen_ms / steps,
                                   "iterations": iters / steps, "reduce": reduce_ms, "reduce_nccl": reduce_nccl_ms,
                                   "reduce_mode": (reducer.mode + (" (" + reducer.why + ")" if reducer.why else "")) if reducer else None},
            "gpu_launches": int(launches),
            "e2e": {"value": total_samples / (e2e_med * 1e-3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": W * H * 3 * 4, "ms_per_step": e2e_med, "steps": len(e2e_ms)},
            "roofline": roof, "cpu_baseline": cpu, "clocks": clk, "mean_radiance": lum, "wall_s": t_wall,
            "scene": {"prims": info["n_prims"], "bvh_nodes": info["n_bvh_nodes"], "media": info["n_media"], "device_bytes": info["device_bytes"]},
        }
        if cpu:
            res["vs_cpu_baseline"] = {"value_ratio": value / cpu["value"], "e2e_ratio": res["e2e"]["value"] / cpu["value"]}
        return res


def single_process_arm(scene, n_gpus, steps):
    """rank 0 only: the same workload through ONE rt_render_multi call per step (one process driving all N GPUs, host
    buffers) in a separate process, so that a fault there cannot take the bench line down."""
    try:
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gpu_multi_check.py"), "--scene", scene, "--gpus", str(n_gpus),
                              "--steps", str(steps), "--json"], capture_output=True, text=True, timeout=420)
        for line in reversed(out.stdout.strip().splitlines()):
            if line.startswith("{"):
                return json.loads(line)
        return {"error": (out.stderr or out.stdout)[-400:]}
    except Exception as e:  # noqa: BLE001
        return {"error": f"{type(e).__name__}: {e}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--scene", default=DEFAULT_SCENE, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-per-scene", action="store_true")
    ap.add_argument("--no-single-process", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    args.steps = max(args.steps, 1)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback; use --impl reference for the CPU arm)")
    done_flag = f"/tmp/rt_b200_bench_{os.environ.get('MASTER_PORT', 'local')}.done"
    if rank == 0 and os.path.exists(done_flag):
        os.remove(done_flag)
    b = Bench(rank, local_rank, world_size)
    want_cpu = (not args.no_cpu_baseline) and world_size == 1
    main_res = b.measure(args.scene, args.steps, max(args.warmup, 3), max(1, min(args.steps, 5)), want_cpu, clocks=True)
    main_settings = build_workload(args.scene)[0]
    per_scene = {}
    if not args.no_per_scene:
        for name in WORKLOADS:
            if name == args.scene:
                continue
            r = b.measure(name, max(1, min(args.steps, 3)), 3, 3, want_cpu)
            if rank == 0:
                for drop in ("clocks", "wall_s"):
                    r.pop(drop, None)
                per_scene[name] = r
    b.barrier()
    single = None
    if b.distributed and not args.no_single_process:
        # the other ranks idle on the HOST (no NCCL kernel spinning on their GPUs) while rank 0's child process drives all N GPUs
        if rank == 0:
            single = single_process_arm(args.scene, world_size, max(1, min(args.steps, 5)))
            open(done_flag, "w").close()
        else:
            t_end = time.time() + 480
            while not os.path.exists(done_flag) and time.time() < t_end:
                time.sleep(0.05)
        b.barrier()
    if rank != 0:
        if b.distributed:
            b.dist.destroy_process_group()
        return
    if os.path.exists(done_flag):
        os.remove(done_flag)

    r = main_res
    line = {
        "metric": "Msamples/s", "value": r["value"], "unit": "Msamples/s", "n_gpus": world_size, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": arm_config(r["workload"], main_settings, world_size, pool=r["pool"]),
        "mrays_per_s": r["mrays_per_s"], "rays_per_sample": r["rays_per_sample"], "kernel_ms_per_step": r["kernel_ms_per_step"],
        "gpu_launches": r["gpu_launches"], "e2e": r["e2e"], "roofline": r["roofline"], "cpu_baseline": r["cpu_baseline"],
        "vs_cpu_baseline": r.get("vs_cpu_baseline"), "clocks": r["clocks"], "mean_radiance": r["mean_radiance"], "wall_s": r["wall_s"],
        "scene": r["scene"], "peaks_measured_in_run": b.aux, "per_scene": per_scene, "single_process": single,
    }
    print(json.dumps(line), flush=True)
    if b.distributed:
        b.dist.destroy_process_group()


if __name__ == "__main__":
    main()
