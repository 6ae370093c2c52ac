"""north_star gate 3 on every BASELINE config at 4096 spp (reduced resolution so the CPU oracle finishes in a minute):
mean-luminance agreement within 1 % and the per-pixel 4-sigma test with the independent re-seed confirmation, for the
default device tree (binned SAH) AND the reference-shaped tree (RtOptions.bvh_reference = 1); the oracle render of a scene is
shared by the two.  GPU box only.  usage: python tools/gpu_parity_4096.py [--head GITHASH] [scene ...]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from raytrace_b200 import scenes, mkStdGen
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene
from util import luminance, z_scores

argv = sys.argv[1:]
head = ""
if argv and argv[0] == "--head":
    head, argv = argv[1], argv[2:]
CASES = [("readme", dict(width=160)), ("cornell", dict(width=128)), ("cornell_noredirect", dict(width=96)),
         ("demo1", dict(width=160)), ("demo2", dict(width=112, depth=50)), ("pawn", dict(width=128)), ("bunny_volume", dict(width=128))]
SPP = 4096
for name, kw in CASES:
    if argv and name not in argv:
        continue
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(spp=SPP, **kw)
    ora = OracleScene(w)
    t0 = time.time()
    ref, rsq, cnt = ora.render(s, seed, want_sumsq=True)
    to = time.time() - t0
    ref2 = rsq2 = None
    for tree in ("sah", "reference"):
        cs = CompiledScene(w, options=dict(bvh_reference=int(tree == "reference")))
        t0 = time.time()
        img, sq, st = cs.render(s, mkStdGen(seed), want_sumsq=True, want_stats=True)
        tg = time.time() - t0
        img, sq = img.astype(np.float64), sq.astype(np.float64)
        z1 = z_scores(img, sq, SPP, ref, rsq, SPP)
        fail1 = np.abs(z1) > 4
        confirmed = 0
        if fail1.any():
            if ref2 is None:
                ref2, rsq2, _ = ora.render(s, seed + 101, want_sumsq=True)
            img2, sq2 = cs.render(s, mkStdGen(seed + 101), want_sumsq=True)
            z2 = z_scores(img2.astype(np.float64), sq2.astype(np.float64), SPP, ref2, rsq2, SPP)
            confirmed = int((fail1 & (np.abs(z2) > 4) & (np.sign(z1) == np.sign(z2))).sum())
        lg, lo = luminance(img).mean(), luminance(ref).mean()
        print(json.dumps(dict(scene=name, tree=tree, head=head, size=list(img.shape[:2]), spp=SPP, lum_gpu=lg, lum_oracle=lo, rel_lum_diff=lg / lo - 1,
                              nonfinite=int((~np.isfinite(img)).sum()), rays_per_sample_gpu=st["rays"] / st["samples"], rays_per_sample_oracle=cnt["rays"] / cnt["samples"],
                              first_pass_4sigma=int(fail1.sum()), pixel_channels=int(z1.size), confirmed_failures=confirmed,
                              max_abs_z=float(np.abs(z1).max()), gpu_s=round(tg, 3), oracle_s=round(to, 2),
                              gate="PASS" if abs(lg / lo - 1) <= 0.01 and confirmed == 0 else "FAIL")), flush=True)
