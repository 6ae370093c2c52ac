"""Material-supplied direction generator (RtOptions.material_sampling) against the reference estimator, both on the GPU:
block z-scores of 4x4 pixel blocks, hook at N spp vs reference estimator at 32 N, with the reference estimator at N as control."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytrace_b200 as rt
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene

def lum(x): return 0.2126 * x[..., 0] + 0.7152 * x[..., 1] + 0.0722 * x[..., 2]
def block_z(a, asq, na, b, bsq_, nb, B=4):
    def blocks(x):
        h, w, c = x.shape
        return x[: h // B * B, : w // B * B].reshape(h // B, B, w // B, B, c)
    va = np.maximum(asq / na - a ** 2, 0.0) / na
    vb = np.maximum(bsq_ / nb - b ** 2, 0.0) / nb
    ma, mb = blocks(a).mean(axis=(1, 3)), blocks(b).mean(axis=(1, 3))
    v = (blocks(va).sum(axis=(1, 3)) + blocks(vb).sum(axis=(1, 3))) / B ** 4
    return (ma - mb) / np.sqrt(v + (1e-3 * np.maximum(np.abs(mb), 1e-2)) ** 2)

for g, redirect, depth in ((0.7, True, 10), (0.7, True, 2), (0.7, True, 3), (0.3, True, 10), (0.7, False, 10)):
    fog = rt.anisotropic(g, rt.constantTexture((0.9, 0.8, 0.3))) << rt.constantMedium(2.5, rt.sphere((0, 0.2, -2.6), 1.0))
    lamp = rt.lightSource(rt.constantTexture(6.0)) << rt.parallelogram((-0.5, 2.4, -3.1), (1, 0, 0), (0, 0, 1))
    world = rt.group([fog, rt.pureGeometry(lamp), rt.pureGeometry(rt.lambertian(rt.constantTexture(0.5)) << rt.sphere((0, -100.8, -2.6), 100))])
    s = rt.defaultCameraSettings.replace(cs_imageWidth=64, cs_samplesPerPixel=1024, cs_background=scenes.sky, cs_maxRecursionDepth=depth)
    if redirect:
        s = s.replace(cs_redirectTargets=[(0.4, (-0.5, 2.4, -3.1), (1, 0, 0), (0, 0, 1))])
    cs = CompiledScene(world)
    f64 = lambda t: tuple(x.astype(np.float64) for x in t)
    hi, hsq = f64(cs.render(s.replace(cs_samplesPerPixel=32768), rt.mkStdGen(10), want_sumsq=True))
    ctl, csq = f64(cs.render(s, rt.mkStdGen(11), want_sumsq=True))
    cs.handle.set_options(material_sampling=1)
    on, osq = f64(cs.render(s, rt.mkStdGen(11), want_sumsq=True))
    hi2, hsq2 = f64(cs.render(s.replace(cs_samplesPerPixel=32768), rt.mkStdGen(12), want_sumsq=True))
    cs.handle.set_options(material_sampling=None)
    zb, zc, zh = block_z(on, osq, 1024, hi, hsq, 32768), block_z(ctl, csq, 1024, hi, hsq, 32768), block_z(hi2, hsq2, 32768, hi, hsq, 32768)
    print(f"g={g} redirect={redirect} depth={depth}: lum hi {lum(hi).mean():.5f} hook-hi {lum(hi2).mean():.5f} ctl {lum(ctl).mean():.5f} hook {lum(on).mean():.5f} | "
          f"hook z rms {np.sqrt((zb**2).mean()):.3f} max {np.abs(zb).max():.2f} | control z rms {np.sqrt((zc**2).mean()):.3f} max {np.abs(zc).max():.2f} | "
          f"hook32k vs ref32k z rms {np.sqrt((zh**2).mean()):.3f} max {np.abs(zh).max():.2f}", flush=True)
    bad = np.argwhere(np.abs(zh) > 4)
    for b in bad[:6]:
        print("   block", b, "z", zh[tuple(b)], flush=True)
