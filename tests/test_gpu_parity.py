"""GPU parity tests proper (-m gpu): the CUDA path, called through the C ABI (ctypes → librt_b200.so),
against the CPU oracle on the same scenes.

Gates (BASELINE.json north_star):
  1. primary-ray primitive id matches on >= 99.99 % of pixels;
  2. hit distance t within 1e-5 relative;
  3. images agree statistically: mean luminance within 1 %, and no pixel fails a 4-sigma test
     (a pixel fails only if it fails again under an independent re-seed: SURVEY §7.3 item 4).
Gates 1–2 count ALL pixels (oracle.harness.primary_gate): a pixel whose id differs from the oracle's centre answer
passes only if the GPU's id is one the oracle itself returns at a ±2e-6 rad tilt of the same ray and the GPU's t is
within 1e-5 of the oracle's t for that id (the pixel-centre ray runs through a primitive edge: README scene, the
mirror's left edge projects exactly onto pixel column 46, a = 2.2e-16, and which side wins is fp64 rounding noise in
the reference itself); raw, tie-explained and unexplained counts are printed."""
import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene
from util import MISS, luminance, z_scores

pytestmark = pytest.mark.gpu

PROBE_CASES = [
    ("readme", dict()),
    ("cornell", dict(spp=1)),
    ("demo1", dict(spp=1)),
    ("demo2", dict(width=800, spp=1)),
    ("pawn", dict(width=1920, aspect=16 / 9, spp=1)),
    ("bunny_volume", dict(spp=1)),
]


_ORACLE_PROBES = {}


def _oracle_probe(name, kw):
    """centre + four tilted oracle probes of a full-size config (cached: the SAH and reference-tree runs share it)"""
    if name not in _ORACLE_PROBES:
        ctor, _ = scenes.CONFIGS[name]
        s, w = ctor(**kw)
        _ORACLE_PROBES[name] = OracleScene(w).probe_tilted(s)
    return _ORACLE_PROBES[name]


@pytest.mark.parametrize("tree", ["sah", "reference"])
@pytest.mark.parametrize("name,kw", PROBE_CASES, ids=[c[0] for c in PROBE_CASES])
def test_primary_ray_id_and_distance(built, name, kw, tree):
    """Gates 1 and 2 at BASELINE.json's full image sizes, over ALL pixels (nothing masked): a pixel passes when the
    GPU's id is one the oracle returns at the centre ray or at one of the four ±2e-6 rad tilts AND its t is within 1e-5
    of the oracle's t for that id (oracle.harness.primary_gate); >= 99.99 % of all pixels must pass.  Run for the
    default device tree (binned SAH, near-first traversal) and for the reference-shaped tree (RtOptions.bvh_reference:
    the split policy north_star names, Geometry.hs:369-377)."""
    from oracle.harness import primary_gate
    ctor, _ = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    ids, ts = CompiledScene(w, options=dict(bvh_reference=int(tree == "reference"))).probe_primary(s)
    ids5, ts5 = _oracle_probe(name, kw)
    g = primary_gate(ids, ts, ids5, ts5)
    _, _, stable = OracleScene.probe_stable(None, s, tilted=(ids5, ts5))
    both = stable & (ids == ids5[0]) & (ids5[0] != MISS)
    rel = np.abs(ts[both] - ts5[0][both]) / ts5[0][both]
    print(f"{name}/{tree}: {g['pixels']} px, raw id match {g['raw_frac']:.6f} ({g['pixels'] - g['raw_match']} differ: "
          f"{g['tie_explained']} tie-explained, {g['id_unexplained']} unexplained id, {g['t_failed']} t beyond 1e-5), "
          f"pass {g['pass_frac']:.6f}; max rel dt on well-conditioned px {rel.max():.2e}")
    assert g["pass_frac"] >= 0.9999, g
    assert g["raw_frac"] >= 0.999
    assert rel.max() <= 1e-5
    assert np.isinf(ts[ids == MISS]).all()


EQUAL_SEED_TREES = [("demo1", dict(width=96, spp=8)), ("demo2", dict(width=64, spp=8, depth=12)), ("pawn", dict(width=64, spp=8)),
                    ("bunny_volume", dict(width=64, spp=4)), ("cornell", dict(width=48, spp=8))]


@pytest.mark.parametrize("name,kw", EQUAL_SEED_TREES, ids=[c[0] for c in EQUAL_SEED_TREES])
def test_reference_shaped_tree_renders_the_same_samples(built, name, kw):
    """The device tree only changes HOW the closest hit is searched: the reference-shaped tree (bvhTree's object-median
    policy, left-first evaluation order = storage order) and the default SAH tree give the same path segments and,
    at equal seeds, the same image up to fp32 summation order (a knife-edge tie may flip a handful of paths)."""
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    a, sa = CompiledScene(w).render(s, rt.mkStdGen(seed), want_stats=True)
    b, sb = CompiledScene(w, options=dict(bvh_reference=1)).render(s, rt.mkStdGen(seed), want_stats=True)
    assert sa["samples"] == sb["samples"]
    assert abs(int(sa["rays"]) - int(sb["rays"])) <= 2e-4 * sb["rays"] + 2
    close = np.isclose(a, b, rtol=2e-4, atol=2e-5).all(axis=2)
    assert close.mean() >= 0.998, close.mean()
    assert abs(a.mean() - b.mean()) <= 2e-3 * b.mean()


STAT_CASES = [
    # name, ctor kwargs, spp per run
    ("readme", dict(width=192), 256),
    ("cornell", dict(width=96), 256),
    ("cornell_noredirect", dict(width=96), 256),
    ("demo1", dict(width=160), 128),
    ("demo2", dict(width=96, depth=50), 128),
    ("pawn", dict(width=96), 128),
    ("bunny_volume", dict(width=128), 64),
]


def _render_both(cs, ora, s, seed):
    img, sq, st = cs.render(s, rt.mkStdGen(seed), want_sumsq=True, want_stats=True)
    ref, rsq, cnt = ora.render(s, seed, want_sumsq=True)
    return img.astype(np.float64), sq.astype(np.float64), st, ref, rsq, cnt


@pytest.mark.parametrize("name,kw,spp", STAT_CASES, ids=[c[0] for c in STAT_CASES])
def test_statistical_image_parity(built, name, kw, spp):
    """Gate 3 at sizes the oracle renders in seconds (same camera, fewer pixels)."""
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(spp=spp, **kw)
    cs, ora = CompiledScene(w), OracleScene(w)
    img, sq, st, ref, rsq, cnt = _render_both(cs, ora, s, seed)
    assert np.isfinite(img).all()
    n_px = img.shape[0] * img.shape[1]
    assert st["samples"] == n_px * spp == cnt["samples"]
    # same path-length statistics (rays = rayColor segments that reach hitWorld, Ray.hs:178)
    rps_g, rps_o = st["rays"] / st["samples"], cnt["rays"] / cnt["samples"]
    assert abs(rps_g - rps_o) <= 0.01 * rps_o, (rps_g, rps_o)
    lum_g, lum_o = luminance(img).mean(), luminance(ref).mean()
    print(f"{name}: mean luminance gpu {lum_g:.5f} oracle {lum_o:.5f} ({(lum_g / lum_o - 1) * 100:+.3f} %), rays/sample {rps_g:.3f} vs {rps_o:.3f}")
    assert abs(lum_g - lum_o) <= 0.01 * lum_o
    for c in range(3):
        mg, mo = img[..., c].mean(), ref[..., c].mean()
        assert abs(mg - mo) <= 0.015 * max(mo, 1e-3), (c, mg, mo)
    z1 = z_scores(img, sq, spp, ref, rsq, spp)
    fail1 = np.abs(z1) > 4
    print(f"   first pass: {int(fail1.sum())} of {z1.size} pixel-channels beyond 4 sigma (max |z| {np.abs(z1).max():.2f})")
    assert fail1.mean() < 2e-3
    if fail1.any():   # confirm on an independent re-seed
        img2, sq2, _, ref2, rsq2, _ = _render_both(cs, ora, s, seed + 7919)
        z2 = z_scores(img2, sq2, spp, ref2, rsq2, spp)
        confirmed = fail1 & (np.abs(z2) > 4) & (np.sign(z1) == np.sign(z2))
        assert not confirmed.any(), np.argwhere(confirmed)[:10]
    # block-level agreement (tighter than per-pixel: averages 64 pixels)
    b = 8
    h, wd = img.shape[0] // b * b, img.shape[1] // b * b
    bg = luminance(img)[:h, :wd].reshape(h // b, b, wd // b, b).mean(axis=(1, 3))
    bo = luminance(ref)[:h, :wd].reshape(h // b, b, wd // b, b).mean(axis=(1, 3))
    assert np.abs(bg - bo).mean() <= 0.05 * max(bo.mean(), 1e-3)


FEATURE_CASES = [
    ("metal_test", scenes.metal_test, dict(width=128, spp=128)),          # dielectric bubble, metal 1.0, defocus
    ("noise_test", scenes.noise_test, dict(width=128, spp=64)),           # noiseTexture + marbleTexture (Perlin in smem)
    ("quad_test", scenes.quad_test, dict(width=96, spp=64)),              # parallelograms
    ("cuboid_test", scenes.cuboid_test, dict(width=96, spp=64)),          # cuboid + imageTexture + transform
    ("cornell_smoke", scenes.cornell_smoke, dict(width=80, spp=128)),     # pitchBlack / isotropic media in transformed cuboids
    ("lommel_seeliger", scenes.lommel_seeliger_test, dict(width=96, spp=128)),  # lommelSeeliger + redirect p = 0.5
    ("bunny_test", scenes.bunny_test, dict(width=128, spp=32)),           # 4968-triangle BVH, lambertian
]


@pytest.mark.parametrize("name,ctor,kw", FEATURE_CASES, ids=[c[0] for c in FEATURE_CASES])
def test_reference_feature_scenes(built, name, ctor, kw):
    """The reference's other scene programs (test/Main.hs): each exercises one feature family."""
    s, w = ctor(**kw)
    spp = s.cs_samplesPerPixel
    cs, ora = CompiledScene(w), OracleScene(w)
    ids, ts = cs.probe_primary(s)
    oids, ots, stable = ora.probe_stable(s)
    assert (ids == oids)[stable].mean() >= 0.9999
    both = stable & (ids == oids) & (oids != MISS)
    if both.any():
        assert (np.abs(ts[both] - ots[both]) / ots[both]).max() <= 1e-5
    img, sq, st, ref, rsq, cnt = _render_both(cs, ora, s, 55)
    lum_g, lum_o = luminance(img).mean(), luminance(ref).mean()
    print(f"{name}: luminance gpu {lum_g:.5f} oracle {lum_o:.5f}")
    assert abs(lum_g - lum_o) <= 0.01 * lum_o + 1e-4
    assert abs(st["rays"] / st["samples"] - cnt["rays"] / cnt["samples"]) <= 0.01 * cnt["rays"] / cnt["samples"]
    z = z_scores(img, sq, spp, ref, rsq, spp)
    assert (np.abs(z) > 4).mean() < 2e-3
    for c in range(3):
        assert abs(img[..., c].mean() - ref[..., c].mean()) <= 0.015 * max(ref[..., c].mean(), 1e-3) + 1e-4


def test_moving_and_anisotropic(built):
    """moving under transform (chain collapse), anisotropic + transparent materials, bvhNode, triangle with uv."""
    tex = rt.checkerTexture(4, 4, (0.9, 0.2, 0.2), (0.2, 0.2, 0.9))
    tri = rt.triangle(((-2, -1, -4), (0, 0)), ((2, -1, -4), (1, 0)), ((0, 2.5, -4.5), (0.5, 1)))
    mover = rt.transform(rt.translate((0.5, 0, 0)) @ rt.rotateZ(0.3),
                         rt.moving((0, 0, 0), (0, 0.6, 0), rt.transform(rt.rotateY(0.7), rt.cuboid(rt.fromCorners((-0.4, -0.4, -3), (0.4, 0.4, -2.2))))))
    world = rt.group([
        rt.pureGeometry(rt.bvhNode(rt.lambertian(tex) << tri, rt.metal(0.2, rt.constantTexture(0.8)) << mover)),
        rt.pureGeometry(rt.transparent(rt.constantTexture((0.6, 0.9, 0.6))) << rt.sphere((-1.2, 0.2, -2.5), 0.4)),
        rt.anisotropic(0.6, rt.constantTexture((0.9, 0.8, 0.3))) << rt.constantMedium(1.5, rt.sphere((1.3, 0.3, -2.6), 0.6)),
    ])
    s = rt.defaultCameraSettings.replace(cs_imageWidth=96, cs_samplesPerPixel=128, cs_background=scenes.sky, cs_maxRecursionDepth=12)
    cs, ora = CompiledScene(world), OracleScene(world)
    for time in (0.0, 0.35, 1.0):
        ids, ts = cs.probe_primary(s, time=time)
        oids, ots, stable = ora.probe_stable(s, time=time)
        assert (ids == oids)[stable].mean() >= 0.9995
        both = stable & (ids == oids) & (oids != MISS)
        assert (np.abs(ts[both] - ots[both]) / ots[both]).max() <= 1e-5
    img, sq, st, ref, rsq, cnt = _render_both(cs, ora, s, 3)
    assert abs(luminance(img).mean() - luminance(ref).mean()) <= 0.01 * luminance(ref).mean()
    z = z_scores(img, sq, 128, ref, rsq, 128)
    assert (np.abs(z) > 4).mean() < 2e-3


@pytest.mark.parametrize("g,redirect", [(0.6, False), (-0.5, False), (0.4, True)], ids=["forward", "backward", "with_redirect"])
def test_material_supplied_direction_generator(built, g, redirect):
    """RtOptions.material_sampling (the generalisation sketched in Material.hs:36-38): `anisotropic g` draws the
    no-redirection direction from its own Henyey-Greenstein phase function, with the mixture pdf of Ray.hs:187-224 built on that
    generator's pdf.  Off by default.  On, the image converges to the same values as the reference's estimator (oracle, uniform
    directions weighted by the phase function) - and, without redirect targets, with less variance: the weight is the albedo."""
    fog = rt.anisotropic(g, rt.constantTexture((0.9, 0.8, 0.3))) << rt.constantMedium(2.5, rt.sphere((0, 0.2, -2.6), 1.0))
    lamp = rt.lightSource(rt.constantTexture(6.0)) << rt.parallelogram((-0.5, 2.4, -3.1), (1, 0, 0), (0, 0, 1))
    world = rt.group([fog, rt.pureGeometry(lamp),
                      rt.pureGeometry(rt.lambertian(rt.constantTexture(0.5)) << rt.sphere((0, -100.8, -2.6), 100))])
    s = rt.defaultCameraSettings.replace(cs_imageWidth=64, cs_samplesPerPixel=256, cs_background=scenes.sky, cs_maxRecursionDepth=10)
    if redirect:
        s = s.replace(cs_redirectTargets=[(0.4, (-0.5, 2.4, -3.1), (1, 0, 0), (0, 0, 1))])
    cs, ora = CompiledScene(world), OracleScene(world)
    base, bsq = cs.render(s, rt.mkStdGen(9), want_sumsq=True)
    cs.handle.set_options(material_sampling=1)
    img, sq = cs.render(s, rt.mkStdGen(9), want_sumsq=True)
    cs.handle.set_options(material_sampling=None)
    again = cs.render(s, rt.mkStdGen(9))
    assert np.allclose(again, base, rtol=2e-4, atol=2e-5)          # the option leaves nothing behind
    assert not np.allclose(img, base, rtol=2e-4, atol=2e-5)         # other samples ...
    ref, rsq, cnt = ora.render(s, 9, want_sumsq=True)
    img, sq = img.astype(np.float64), sq.astype(np.float64)
    lg, lo = luminance(img).mean(), luminance(ref).mean()
    assert abs(lg - lo) <= 0.01 * lo, (lg, lo)                      # ... same expectation
    # Unbiasedness, pixel block by pixel block.  The REFERENCE estimator is heavy-tailed here (its phase-function weight reaches
    # (1 + |g|) / (1 - |g|)^2 = 10 at g = 0.6, 19 at 0.7), so at 256 spp neither its means nor its variance estimates are
    # Gaussian enough to gate on; the yardstick is the GPU's own reference-estimator render (the code path the oracle gates
    # above and test_moving_and_anisotropic pin) at 32 x the samples, against 4 x 4 block means of a 1024-spp hook render.
    # (At g = 0.7 and depth 10 even 32768 spp of the reference estimator sit 4 - 6 sigma BELOW the hook's value in the middle of
    # the fog ball, and 4.9 sigma from its own 1024-spp render: profiles/r02y_material_sampling_check.txt, tools/gpu_hook_check.py.)
    def block_z(a, asq, na, b, bsq_, nb, B=4):
        def blocks(x):
            h, w, c = x.shape
            return x[: h // B * B, : w // B * B].reshape(h // B, B, w // B, B, c)
        va = np.maximum(asq / na - a ** 2, 0.0) / na
        vb = np.maximum(bsq_ / nb - b ** 2, 0.0) / nb
        ma, mb = blocks(a).mean(axis=(1, 3)), blocks(b).mean(axis=(1, 3))
        v = (blocks(va).sum(axis=(1, 3)) + blocks(vb).sum(axis=(1, 3))) / B ** 4
        return (ma - mb) / np.sqrt(v + (1e-3 * np.maximum(np.abs(mb), 1e-2)) ** 2)
    hi, hsq = (x.astype(np.float64) for x in cs.render(s.replace(cs_samplesPerPixel=32768), rt.mkStdGen(10), want_sumsq=True))
    cs.handle.set_options(material_sampling=1)
    on, osq = (x.astype(np.float64) for x in cs.render(s.replace(cs_samplesPerPixel=1024), rt.mkStdGen(11), want_sumsq=True))
    cs.handle.set_options(material_sampling=None)
    zb = block_z(on, osq, 1024, hi, hsq, 32768)
    print(f"block z-scores, hook at 1024 spp vs reference estimator at 32768 spp: rms {np.sqrt((zb ** 2).mean()):.3f} max {np.abs(zb).max():.2f}")
    assert (np.abs(zb) > 4).mean() < 5e-3 and np.sqrt((zb ** 2).mean()) < 1.3
    assert abs(luminance(on).mean() - luminance(hi).mean()) <= 0.005 * luminance(hi).mean()
    var_on = np.maximum(sq / 256 - img ** 2, 0).sum()
    var_off = np.maximum(bsq.astype(np.float64) / 256 - base.astype(np.float64) ** 2, 0).sum()
    print(f"material_sampling g={g} redirect={redirect}: summed pixel variance {var_off:.4f} -> {var_on:.4f}")
    if not redirect:
        assert var_on < var_off


def test_material_sampling_leaves_other_materials_alone(built):
    """With RtOptions.material_sampling on, materials without a generator of their own draw exactly the reference's samples: a
    flat scene (which then runs the wavefront kernels instead of the fused one) and a BVH scene render the same images."""
    for ctor, kw, seed in ((scenes.cornell_box, dict(spp=16, width=64), 3), (scenes.demo1, dict(spp=8, width=64), 1)):
        s, w = ctor(**kw)
        cs = CompiledScene(w)
        base, st0 = cs.render(s, rt.mkStdGen(seed), want_stats=True)
        cs.handle.set_options(material_sampling=1)
        img, st = cs.render(s, rt.mkStdGen(seed), want_stats=True)
        assert st["rays"] == st0["rays"]
        assert np.allclose(img, base, rtol=2e-4, atol=2e-5)


def test_redirect_probability_one(built):
    """cs_redirectTargets whose probabilities sum to 1 (rem_prob = 0): the mixture pdf is the light pdf alone, so a
    sampled light point that fp32 would place a few ulp outside the target must still get its analytic pdf - an
    unguarded 0 makes the weight infinite and one NaN owns the pixel for the whole render.  Every pixel stays finite and
    the image agrees with the oracle (same estimator in fp64, Ray.hs:187-206)."""
    s, w = scenes.cornell_box(spp=512, width=64)
    (p, q, s0, s1), = s.cs_redirectTargets
    s = s.replace(cs_redirectTargets=[(1.0, q, s0, s1)], cs_maxRecursionDepth=8)
    cs, ora = CompiledScene(w), OracleScene(w)
    img, sq, st, ref, rsq, cnt = _render_both(cs, ora, s, 77)
    assert np.isfinite(img).all() and np.isfinite(sq).all()
    assert img.max() < 1e3
    lg, lo = luminance(img).mean(), luminance(ref).mean()
    print(f"redirect p=1: luminance gpu {lg:.5f} oracle {lo:.5f}")
    assert abs(lg - lo) <= 0.02 * lo
    z = z_scores(img, sq, 512, ref, rsq, 512)
    assert (np.abs(z) > 4).mean() < 5e-3
    # a long render: 2^22 redirect samples per pixel-row would have met the 1e-6 edge case many times over
    big = cs.render(s.replace(cs_samplesPerPixel=4096, cs_imageWidth=48), rt.mkStdGen(3))
    assert np.isfinite(big).all() and big.max() < 1e3


def test_sphere_uv_known_answer_gpu(built):
    """sphereUVTest (test/Main.hs:128-134): fully deterministic — the centre texel of earthmap."""
    s, w = scenes.sphere_uv_test()
    cs = CompiledScene(w)
    ids, ts = cs.probe_primary(s)
    assert ids[0, 0] == 1 and abs(ts[0, 0] - 0.6) < 1e-5
    img = cs.render(s.replace(cs_samplesPerPixel=256), rt.mkStdGen(12))
    texel = scenes.load_earthmap()[256, 512]
    assert np.allclose(img[0, 0], texel, atol=1e-2)


def test_sample_ranges_compose_and_are_deterministic(built):
    """Sample-range sharding (the multi-GPU partition): rendering [0,a) and [a,n) separately gives the same
    image as [0,n) up to fp32 summation order; the sample set does not depend on the pool size."""
    s, w = scenes.cornell_box(spp=32, width=64)
    cs = CompiledScene(w)
    full = cs.render(s, rt.mkStdGen(9))
    a = cs.render(s, rt.mkStdGen(9), first_sample=0, n_samples=12)
    b = cs.render(s, rt.mkStdGen(9), first_sample=12, n_samples=20)
    assert np.allclose(a + b, full, rtol=1e-4, atol=1e-5)
    cs.handle.set_pool_size(4096)
    small_pool = cs.render(s, rt.mkStdGen(9))
    assert np.allclose(small_pool, full, rtol=1e-4, atol=1e-5)
    other = cs.render(s, rt.mkStdGen(10))
    assert not np.allclose(other, full, rtol=1e-3)


def test_image_output_step(built, tmp_path):
    """writeImage / writeImageSqrt (Ray.hs:248-260): OETF pin 0.74 -> 224, clamp 15 -> 255."""
    m = np.zeros((2, 2, 3), np.float32)
    m[0, 0] = 0.74
    m[0, 1] = 15.0
    m[1, 0] = 0.25
    m[1, 1] = (-1.0, 0.0031308, 1.0)
    from raytrace_b200.ray import _tonemap
    o = _tonemap(m, 0)
    assert o[0, 0, 0] == 224 and (o[0, 1] == 255).all() and o[1, 0, 0] == 137 and tuple(o[1, 1]) == (0, 10, 255)
    q = _tonemap(m, 1)
    assert q[1, 0, 0] == 128 and q[0, 0, 0] == int(np.sqrt(0.74) * 256)
    rt.writeImage(str(tmp_path / "a.png"), m)
    back = rt.readImage(str(tmp_path / "a.png"))
    assert abs(back[0, 0, 0] - 0.74) < 0.01 and abs(back[1, 0, 0] - 0.25) < 0.005
