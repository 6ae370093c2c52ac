"""Full-size throughput runs (GPU only, no oracle).
usage: python tools/gpu_full.py [--opts k=v,k=v] [scene ...]      (--opts: RtOptions fields, include/rt_b200.h)"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
CASES = {
 "readme": dict(), "cornell": dict(), "cornell_noredirect": dict(), "demo1": dict(spp=100),
 "demo2": dict(spp=100), "pawn": dict(width=1920, aspect=16/9, spp=64), "bunny_volume": dict(spp=64)}
argv = sys.argv[1:]
OPTS = {}
if argv and argv[0] == "--opts":
    OPTS = {k: float(v) if "." in v else int(v) for k, v in (kv.split("=") for kv in argv[1].split(",") if kv)}
    argv = argv[2:]
names = argv or list(CASES)
for name in names:
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**CASES[name])
    cs = CompiledScene(w, options=OPTS)
    for rep in range(3):
        img, st = cs.render(s, mkStdGen(seed), want_stats=True)
    print(json.dumps(dict(scene=name, W=s.cs_imageWidth, spp=s.cs_samplesPerPixel, ms=st["ms_total"], trav=st["ms_traverse"], shade=st["ms_shade"],
        other=st["ms_raygen"], iters=st["iterations"], msamples_s=st["samples"]/st["ms_total"]/1e3, mrays_s=st["rays"]/st["ms_total"]/1e3,
        rays_per_sample=st["rays"]/st["samples"], lum=float(img.mean()), nan=int(np.isnan(img).sum()))), flush=True)
