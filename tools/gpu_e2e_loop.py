import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
import raytrace_b200 as rt
name = sys.argv[1]
ctor, seed = scenes.CONFIGS[name]
s, w = ctor()
s = s.replace(cs_samplesPerPixel=8)
CompiledScene(w).render(s, mkStdGen(1))
for rep in range(10):
    t0 = time.perf_counter(); cs = CompiledScene(w, upload=False)
    t1 = time.perf_counter(); cs.handle.upload(0)
    t2 = time.perf_counter(); img = cs.render(s, mkStdGen(seed + rep))
    t3 = time.perf_counter(); del cs
    t4 = time.perf_counter()
    print(f"{name}: commit {1e3*(t1-t0):.2f}  upload {1e3*(t2-t1):.2f}  render {1e3*(t3-t2):.2f} destroy {1e3*(t4-t3):.2f} ms", flush=True)
t0 = time.perf_counter()
for rep in range(10):
    img = rt.raytrace(s, w, mkStdGen(rep))
print("raytrace() x10 mean ms", (time.perf_counter() - t0) * 100)
