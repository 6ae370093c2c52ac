"""Two (or more) independent wavefronts on ONE GPU: a context that lists device 0 several times renders disjoint sample ranges on
separate streams (traversal is latency-bound, shading HBM-bound: do they fill each other's idle slots?).  GPU box only.
usage: python tools/gpu_overlap.py [scene ...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene, RenderContext
CASES = {"cornell": dict(), "demo1": dict(spp=100), "demo2": dict(spp=100), "pawn": dict(width=1920, aspect=16 / 9, spp=64), "bunny_volume": dict(spp=64)}
for name in (sys.argv[1:] or list(CASES)):
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**CASES[name])
    cs = CompiledScene(w)
    row = {}
    for shares, pool in ((1, None), (2, None), (2, 1 << 23), (3, 1 << 23), (4, 1 << 22)):
        cs.handle.set_options(pool=pool)
        ctx = RenderContext([0] * shares)
        best = None
        for rep in range(4):
            t0 = time.perf_counter()
            img = ctx.render(cs, s, mkStdGen(seed + rep))
            dt = (time.perf_counter() - t0) * 1e3
            if rep and (best is None or dt < best):
                best = dt
        row[f"{shares}x{'16Mi' if pool is None else str(pool >> 20) + 'Mi'}"] = round(best, 2)
        ctx.close()
    print(json.dumps(dict(scene=name, wall_ms=row, mean=float(img.mean()))), flush=True)
