"""Device-side tree build vs the host builder (GPU only): commit + first upload, and how fast the resulting tree renders.

    python tools/gpu_build_bench.py [pawn bunny_volume demo2 mesh:20000 mesh:200000 mesh:1000000] [--render]

`mesh:N` is a synthetic closed surface of about N triangles (a sphere displaced by a few sine waves: uneven triangle sizes,
lambertian, seen from outside).  Prints one JSON line per (scene, builder)."""
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                                          # noqa: E402
import raytrace_b200 as rt                                  # noqa: E402
from raytrace_b200 import scenes, mkStdGen                  # noqa: E402
from raytrace_b200.geometry import Mesh                     # noqa: E402
from raytrace_b200.ray import CompiledScene                 # noqa: E402


def bumpy_sphere(n_tris: int):
    rows = max(4, int(round(math.sqrt(n_tris / 4.0))))
    cols = 2 * rows
    th = np.linspace(0.02, math.pi - 0.02, rows + 1)[:, None]
    ph = np.linspace(0, 2 * math.pi, cols, endpoint=False)[None, :]
    r = 1.0 + 0.15 * np.sin(5 * th) * np.cos(7 * ph) + 0.05 * np.sin(23 * ph + 3 * th) + 0.3 * np.exp(-8 * (th - 1.0) ** 2) * np.cos(ph) ** 8
    x, y, z = r * np.sin(th) * np.cos(ph), r * np.cos(th) + 0 * ph, r * np.sin(th) * np.sin(ph)
    verts = np.stack([x, y, z], axis=-1).reshape(-1, 3)
    i, j = np.meshgrid(np.arange(rows), np.arange(cols), indexing="ij")
    a, b = i * cols + j, i * cols + (j + 1) % cols
    c, d = (i + 1) * cols + j, (i + 1) * cols + (j + 1) % cols
    tri = np.concatenate([np.stack([a, c, b], -1).reshape(-1, 3), np.stack([b, c, d], -1).reshape(-1, 3)])
    faces = np.full((len(tri), 6), -1, np.int32)
    faces[:, 0::2] = tri
    world = rt.lambertian(rt.constantTexture((0.7, 0.6, 0.5))) << rt.triangleMesh(Mesh(verts, np.zeros((0, 2)), faces))
    s = rt.defaultCameraSettings.replace(cs_center=(0, 0.6, 3.2), cs_lookAt=(0, 0, 0), cs_imageWidth=960, cs_aspectRatio=16 / 9,
                                         cs_vfov=math.radians(45), cs_samplesPerPixel=16, cs_maxRecursionDepth=8, cs_background=scenes.sky)
    return s, world


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")] or ["pawn", "bunny_volume", "demo2", "mesh:20000", "mesh:200000", "mesh:1000000"]
    render = "--render" in sys.argv
    trace = "--trace" in sys.argv
    s0, w0 = scenes.cornell_box(spp=1, width=32)
    CompiledScene(w0).render(s0, mkStdGen(1))               # CUDA context, workspace
    CompiledScene(scenes.pawn(width=32, spp=1)[1]).probe_primary(scenes.pawn(width=32, spp=1)[0])     # builder scratch, kernels loaded
    for name in args:
        if name.startswith("mesh:"):
            s, w = bumpy_sphere(int(name[5:]))
        else:
            s, w = scenes.CONFIGS[name][0]()
            s = s.replace(cs_imageWidth=960, cs_samplesPerPixel=32)
        ref = None
        for builder, opts in (("host", dict(gpu_build=0)), ("device", dict(gpu_build=1))):
            best = None
            for rep in range(4):
                t0 = time.perf_counter()
                cs = CompiledScene(w, upload=False, options=dict(opts, trace_commit=int(trace and rep == 3)))
                t1 = time.perf_counter()
                cs.handle.upload(0)
                t2 = time.perf_counter()
                cur = (t2 - t0, t1 - t0, t2 - t1)
                best = cur if best is None or cur[0] < best[0] else best
                if rep < 3:
                    del cs
            info = cs.info()
            line = {"scene": name, "builder": builder, "n_prims": info["n_prims"], "node_pairs": info["n_bvh_nodes"] // 2, "depth": info["max_depth"],
                    "commit_ms": round(best[1] * 1e3, 3), "first_upload_ms": round(best[2] * 1e3, 3), "commit_plus_upload_ms": round(best[0] * 1e3, 3)}
            ids, ts = cs.probe_primary(s)
            if ref is None:
                ref = (ids, ts)
            else:
                line["same_hits_as_host_tree"] = bool(np.array_equal(ids, ref[0]) and np.array_equal(ts.view(np.uint32), ref[1].view(np.uint32)))
            if render:
                cs.render(s, mkStdGen(3))
                ms = []
                for rep in range(3):
                    _, st = cs.render(s, mkStdGen(4 + rep), want_stats=True)
                    ms.append(st["ms_total"])
                line["render_ms"] = round(min(ms), 3)
                line["mrays_per_s"] = round(st["rays"] / (min(ms) * 1e-3) / 1e6, 1)
            print(json.dumps(line), flush=True)
            del cs


if __name__ == "__main__":
    main()
