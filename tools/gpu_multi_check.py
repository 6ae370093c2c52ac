"""ONE process driving N GPUs through rt_context_create / rt_render_multi (include/rt_b200.h): correctness against the
one-GPU render and throughput with host buffers.  GPU box only.
usage: python tools/gpu_multi_check.py [--scene pawn] [--gpus N] [--steps K] [--json]
  check        small render on all N GPUs vs the same render on GPU 0 (same sample set: equal up to summation order)
  render_call  Msamples/s of rt_render_multi alone (scene committed, resident on every device; host image written by N DMA copies)
  e2e          Msamples/s including the scene commit (description tree -> flatten -> BVH) of every step"""
import argparse, json, os, statistics, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene, RenderContext, imageHeight

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="pawn")
ap.add_argument("--gpus", type=int, default=0)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--json", action="store_true")
a = ap.parse_args()
sys.path.insert(0, ROOT)
import bench as B
key, kw, desc = B.WORKLOADS[a.scene]
ctor, seed = scenes.CONFIGS[key]
ctx = RenderContext(list(range(a.gpus)) if a.gpus else None)
res = {"workload": desc}
# ---- correctness at a small size ----
small_kw = dict(kw)
small_kw.update(width=97, spp=13)
s0, w0 = ctor(**small_kw)
cs0 = CompiledScene(w0, upload=False)
multi, st_m = ctx.render(cs0, s0, mkStdGen(seed), want_stats=True)
one, st_1 = CompiledScene(w0, device=0).render(s0, mkStdGen(seed), want_stats=True)
res["check"] = {"max_abs_diff": float(np.abs(multi - one).max()), "allclose": bool(np.allclose(multi, one, rtol=2e-4, atol=2e-5)),
                "rays_equal": bool(st_m["rays"] == st_1["rays"]), "image": list(multi.shape[:2]), "spp": 13}
info = ctx.info()
res.update(n_gpus=info["n_devices"], reduce_mode=info["reduce_mode"], reduce_note=info["reduce_note"])
# ---- throughput at the full size ----
s, w = ctor(**kw)
W, H, spp = s.cs_imageWidth, imageHeight(s), s.cs_samplesPerPixel
total = W * H * spp
cs = CompiledScene(w, upload=False)
for k in range(3):
    ctx.render(cs, s, mkStdGen(seed + 100 + k))
call_ms, dev_ms, rays = [], [], 0
for k in range(a.steps):
    t0 = time.perf_counter()
    img, st = ctx.render(cs, s, mkStdGen(seed + k), want_stats=True)
    call_ms.append((time.perf_counter() - t0) * 1e3)
    dev_ms.append(st["ms_total"])
    rays = st["rays"]
e2e_ms = []
for k in range(a.steps):
    t0 = time.perf_counter()
    cs2 = CompiledScene(w, upload=False)
    img = ctx.render(cs2, s, mkStdGen(seed + 50 + k))
    e2e_ms.append((time.perf_counter() - t0) * 1e3)
    del cs2
m = statistics.median
res["render_call"] = {"value": total / (m(call_ms) * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": m(call_ms), "device_ms_max_over_gpus": m(dev_ms),
                      "mrays_per_s": rays / (m(call_ms) * 1e-3) / 1e6}
res["e2e"] = {"value": total / (m(e2e_ms) * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": m(e2e_ms), "d2h_bytes_per_step": W * H * 12,
              "note": "scene commit + upload to every GPU + rt_render_multi with a host image, per step"}
res["mean_radiance"] = float(img.mean())
print(json.dumps(res) if a.json else json.dumps(res, indent=1), flush=True)
