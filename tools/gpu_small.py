"""Fixed overhead of a render: Cornell at 200/N spp for N = 1, 2, 4, 8 on one GPU (what each rank of an N-GPU run does)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
s, w = scenes.cornell_box()
cs = CompiledScene(w)
for n in (1, 2, 4, 8):
    spp = 200 // n
    for rep in range(3):
        t0 = time.perf_counter()
        img, st = cs.render(s, mkStdGen(234 + rep), first_sample=0, n_samples=spp, want_stats=True)
        wall = (time.perf_counter() - t0) * 1e3
    print(json.dumps(dict(share=n, spp=spp, device_ms=st["ms_total"], wall_ms=wall, trav=st["ms_traverse"], shade=st["ms_shade"], iters=st["iterations"], launches=st["launches"], ideal_ms=None)))
