"""Latency of ONE GPU's share of a sample-sharded render (GPU only): renders spp/G samples of a scene on one device under
different tail settings (finish_max = path count at which k_finish takes over, tail_batch = iterations per host
round trip in the tail, finish_refill) and prints the device-timed total, the time inside the iteration kernels and
the rest (k_finish, host round trips, launch gaps).  The knobs are read per render, so one process sweeps them all.

usage: python tools/gpu_share.py [scene] [shares, e.g. 1,2,4,8]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene

name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
shares = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "8,1").split(",")]
KW = {"pawn": dict(width=1920, aspect=16 / 9, spp=64), "bunny_volume": dict(spp=64), "demo1": dict(spp=100), "demo2": dict(spp=100)}
ctor, seed = scenes.CONFIGS[name]
s, w = ctor(**KW.get(name, {}))
cs = CompiledScene(w)
KNOBS = ("finish_max", "tail_batch", "finish_refill")
sweeps = [{}]
if os.environ.get("SHARE_SWEEP", "full") == "quick":
    sweeps += [{"finish_max": v} for v in (0, 262144, 1048576, 4194304)]
else:
    sweeps += [{"finish_max": v} for v in (0, 262144, 1048576, 2097152, 8388608, 16777216)]
    sweeps += [{"tail_batch": v} for v in (2, 3, 6, 8)]
    sweeps += [{"finish_refill": v} for v in (1, 12, 32)]
for g in shares:
    n = max(1, s.cs_samplesPerPixel // g)
    for cfg in sweeps:
        cs.handle.set_options(**{k: cfg.get(k) for k in KNOBS})
        best = None
        for rep in range(6):
            img, st = cs.render(s, mkStdGen(seed + rep), n_samples=n, want_stats=True)
            if rep and (best is None or st["ms_total"] < best["ms_total"]):
                best = st
        print(json.dumps(dict(scene=name, share=g, spp=n, cfg=cfg, ms=round(best["ms_total"], 4), kernels=round(best["ms_traverse"] + best["ms_shade"], 4),
                              other=round(best["ms_raygen"], 4), iters=best["iterations"], launches=best["launches"], lum=float(img.mean()) * s.cs_samplesPerPixel / n)), flush=True)
