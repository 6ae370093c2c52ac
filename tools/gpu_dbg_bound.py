"""Debug: bounded vs unbounded interior queries (GPU only)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import raytrace_b200 as rt
from raytrace_b200 import scenes, V3, constantTexture, dielectric, isotropic, group, pureGeometry, constantMedium, sphere, lambertian
from raytrace_b200.ray import CompiledScene
ball = sphere((0, 0, -3), 1.5)
fog_ball = group([pureGeometry(dielectric(1.5) << ball), isotropic(constantTexture(V3(0.9, 0.6, 0.3))) << constantMedium(3.0, ball),
                  pureGeometry(lambertian(constantTexture(V3(0.5, 0.5, 0.5))) << sphere((0, -101.5, -3), 100))])
s_ball = rt.defaultCameraSettings.replace(cs_imageWidth=72, cs_samplesPerPixel=32, cs_maxRecursionDepth=30, cs_background=scenes.sky)
s_pawn, w_pawn = scenes.pawn(width=56, spp=24)
for name, s, w, seed in (("ball", s_ball, fog_ball, 7), ("pawn", s_pawn, w_pawn, 55)):
    for depth in (2, 3, 4, 30):
        s2 = s.replace(cs_maxRecursionDepth=depth)
        cs = CompiledScene(w)
        os.environ["RT_B200_NO_INTERIOR_BOUND"] = "1"
        cs_ref = CompiledScene(w)
        del os.environ["RT_B200_NO_INTERIOR_BOUND"]
        a, sa = cs.render(s2, rt.mkStdGen(seed), want_stats=True)
        b, sb = cs_ref.render(s2, rt.mkStdGen(seed), want_stats=True)
        b2, sb2 = cs_ref.render(s2, rt.mkStdGen(seed), want_stats=True)
        close = np.isclose(a, b, rtol=2e-4, atol=2e-5).all(axis=2)
        print(name, depth, cs.info()["n_closed_media"], cs_ref.info()["n_closed_media"], "rays", sa["rays"], sb["rays"], sb2["rays"], "close", close.mean(), "means", a.mean(), b.mean(), flush=True)
