"""One render of a workload for an `ncu --set full` pass over ALL its launches (GPU only).
usage: gpu_traffic_pass.py <scene> <spp> stats|plain
stats: instrumented render (RtOptions.count_work), prints the segment counts the captured DRAM bytes are divided by;
plain: the render ncu profiles (same seed, same launches)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
CASES = {"readme": dict(), "cornell": dict(), "cornell_noredirect": dict(), "demo1": dict(), "demo2": dict(),
         "pawn": dict(width=1920, aspect=16 / 9), "bunny_volume": dict()}
name, spp, mode = sys.argv[1], int(sys.argv[2]), sys.argv[3]
ctor, seed = scenes.CONFIGS[name]
s, w = ctor(spp=spp, **CASES[name])
cs = CompiledScene(w, options=dict(count_work=1) if mode == "stats" else {})
img, st = cs.render(s, mkStdGen(seed), want_stats=True)
if mode == "stats":
    print(json.dumps(dict(scene=name, spp=spp, samples=st["samples"], rays=st["rays"], rays_finish=st["rays_finish"], rays_counted=st["rays_counted"],
                          iterations=st["iterations"], launches=st["launches"], pool=st["pool"])))
