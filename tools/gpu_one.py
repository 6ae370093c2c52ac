"""Render one small scene (GPU only; for compute-sanitizer runs): python tools/gpu_one.py scene width spp"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
name, width, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
OPTS = {k: int(v) for k, v in (kv.split("=") for kv in (sys.argv[4] if len(sys.argv) > 4 else "").split(",") if kv)}
ctor, seed = scenes.CONFIGS[name]
s, w = ctor(width=width, spp=spp)
img, st = CompiledScene(w, options=OPTS).render(s, mkStdGen(seed), want_stats=True)
print(name, img.shape, float(img.mean()), st["rays"], st["samples"], st["iterations"])
