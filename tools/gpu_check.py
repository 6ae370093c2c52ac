"""Exploratory GPU check (not a test): probe parity + luminance + throughput for every config."""
import os, sys, time, json, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene

CASES = [
    ("readme", dict(width=600, spp=50), 50),
    ("cornell", dict(width=200, spp=64), 64),
    ("cornell_noredirect", dict(width=200, spp=64), 64),
    ("demo1", dict(width=400, spp=32), 32),
    ("demo2", dict(width=200, spp=32, depth=50), 32),
    ("pawn", dict(width=200, spp=32), 32),
    ("bunny_volume", dict(width=320, spp=16), 16),
]
only = sys.argv[1:] 
for name, kw, spp in CASES:
    if only and name not in only: continue
    try:
        ctor, seed = scenes.CONFIGS[name]
        s, w = ctor(**kw)
        cs = CompiledScene(w)
        ora = OracleScene(w)
        t0 = time.time(); ids, ts = cs.probe_primary(s); tp = time.time() - t0
        oids, ots = ora.probe_primary(s)
        match = (ids == oids).mean()
        both = (ids == oids) & (oids != 0xFFFFFFFF)
        rel = np.abs(ts[both] - ots[both]) / ots[both]
        bad = np.argwhere(ids != oids)
        for (j, i) in bad[:8]:
            print('   mismatch px', i, j, 'gpu', ids[j, i], ts[j, i], 'oracle', oids[j, i], ots[j, i])
        if rel.size and rel.max() > 1e-5:
            k = np.argmax(np.where(both, np.abs(ts - ots) / np.where(both, ots, 1), 0)); j, i = divmod(k, ids.shape[1])
            print('   worst dt px', i, j, 'prim', ids[j, i], ts[j, i], ots[j, i])
        t0 = time.time(); img, st = cs.render(s, mkStdGen(seed), want_stats=True); tr = time.time() - t0
        t0 = time.time(); ref, cnt = ora.render(s, seed); to = time.time() - t0
        print(json.dumps(dict(scene=name, id_match=float(match), max_rel_dt=float(rel.max()) if rel.size else 0,
              lum_gpu=float(img.mean()), lum_oracle=float(ref.mean()), nan=int(np.isnan(img).sum()),
              gpu_ms=st["ms_total"], trav_ms=st["ms_traverse"], shade_ms=st["ms_shade"], iters=st["iterations"],
              msamples_s=st["samples"] / st["ms_total"] / 1e3, mrays_s=st["rays"] / st["ms_total"] / 1e3,
              rays_per_sample=st["rays"] / st["samples"], oracle_rays_per_sample=cnt["rays"] / cnt["samples"],
              oracle_msamples_s=cnt["samples"] / cnt["seconds"] / 1e6, wall_render=tr)), flush=True)
    except Exception as e:
        traceback.print_exc()
        print("FAILED", name, e, flush=True)
