"""Where the end-to-end time of one raytrace() call goes (GPU only)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
name = sys.argv[1] if len(sys.argv) > 1 else "cornell"
ctor, seed = scenes.CONFIGS[name]
s, w = ctor()
CompiledScene(w).render(s, mkStdGen(1))          # warm everything
for rep in range(4):
    t0 = time.perf_counter(); cs = CompiledScene(w, upload=False)
    t1 = time.perf_counter(); cs.handle.upload(0)
    t2 = time.perf_counter(); img, st = cs.render(s, mkStdGen(seed + rep), want_stats=True)
    t3 = time.perf_counter(); del cs
    t4 = time.perf_counter()
    print(f"{name}: commit {1e3*(t1-t0):.2f}  upload {1e3*(t2-t1):.2f}  render call {1e3*(t3-t2):.2f} (device {st['ms_total']:.2f}, copy {st['ms_copy']:.2f}, iters {st['iterations']}, launches {st['launches']})  destroy {1e3*(t4-t3):.2f} ms")
