"""Pins the CPU oracle against the reference's own golden renders (tests/golden/png/*.npz, made from
/root/reference/*.png by tests/golden/make_fixtures.py).  The reference has no numeric known-answer
tests (test/Main.hs:7); its shipped PNGs are the only fixtures, so the pins are:
  * exact geometry/camera conventions: full-miss mask and light mask, pixel for pixel;
  * converged appearance: block-averaged linear luminance within Monte Carlo noise.
The RNG stream cannot be pinned (third-party `random`, not in the tree) — images are never compared
bit for bit."""
import os

import numpy as np
import pytest

from raytrace_b200 import scenes
from oracle.harness import OracleScene
from util import block_mean, quantise8, srgb_eotf, srgb_oetf

G = os.path.join(os.path.dirname(__file__), "golden", "png")


def _golden(name):
    return np.load(os.path.join(G, name + ".npz"))["rgb8"]


def test_readme_example_matches_golden(built):
    """README.md:41-55, seed 100, 50 spp, writeImage (sRGB OETF)."""
    g = _golden("example_image")
    s, w = scenes.readme_example()
    ora = OracleScene(w)
    img, cnt = ora.render(s, 100)
    assert img.shape == g.shape == (338, 600, 3)          # round-half-even: 337.5 -> 338 (Ray.hs:123)
    # pixels no sample of which can hit anything are exactly white in both
    ids, _, stable = ora.probe_stable(s, delta=0.5, t_tol=float("inf"))        # same id at centre and the four edge midpoints
    o8 = quantise8(srgb_oetf(img))
    white_g, white_o = (g == 255).all(axis=2), (o8 == 255).all(axis=2)
    sure_miss = (ids == 0xFFFFFFFF) & stable
    sure_hit = (ids != 0xFFFFFFFF) & stable
    assert sure_miss.sum() > 70000
    assert white_g[sure_miss].mean() > 0.9995 and white_o[sure_miss].mean() > 0.9995
    assert white_g[sure_hit].sum() == 0 and white_o[sure_hit].sum() == 0
    assert (white_g ^ white_o).sum() < 200                 # only stochastic edge pixels differ
    # converged appearance
    gl = srgb_eotf(g.astype(np.float64) / 255)
    d = block_mean(gl, 8) - block_mean(np.clip(img, 0, 1), 8)
    assert abs(d.mean()) < 2e-3
    assert np.abs(d).mean() < 4e-3
    assert abs(gl.mean() - np.clip(img, 0, 1).mean()) < 0.005 * gl.mean()
    assert 1.9 < cnt["rays"] / cnt["samples"] < 2.0


@pytest.mark.parametrize("redirect,name", [(True, "cornell_box_redirect"), (False, "cornell_box_noisy")])
def test_cornell_matches_golden(built, redirect, name):
    """cornellBox 200 50 (test/Main.hs:188-218), mkStdGen 234, writeImageSqrt; oracle at 24 spp."""
    g = _golden(name)
    s, w = scenes.cornell_box(spp=24, depth=50, redirect=redirect)
    ora = OracleScene(w)
    img, _ = ora.render(s, 234)
    assert img.shape == g.shape == (600, 600, 3)
    # the light (15,15,15) clamps to 255: its mask is geometry only
    ids, _, stable = ora.probe_stable(s, delta=0.5, t_tol=float("inf"))
    light = (ids == 2) & stable
    assert light.sum() > 1900
    assert (g[light] == 255).all()
    rows, cols = np.where(ids == 2)
    grow, gcol = np.where((g == 255).all(axis=2))
    assert (grow.min(), grow.max(), gcol.min(), gcol.max()) == (77, 99, 247, 352)
    # the saturated mask = pixel-centre light mask grown by the partially covered rim (<= 1 px)
    for a, b in zip((rows.min(), rows.max(), cols.min(), cols.max()), (77, 99, 247, 352)):
        assert abs(int(a) - b) <= 1
    o8 = quantise8(np.sqrt(np.clip(img, 0, 1)))
    assert (o8[light] == 255).all()
    # block luminance (undo the sqrt 'gamma'); clamp both at 1 as the PNG does
    gl = (g.astype(np.float64) / 255) ** 2
    b = 30
    d = block_mean(gl, b) - block_mean(np.clip(img, 0, 1), b)
    assert abs(d.mean()) < 4e-3, d.mean()
    assert np.abs(d).mean() < (0.012 if redirect else 0.03), np.abs(d).mean()


def test_pawn_matches_golden(built):
    """pawnTest (test/Main.hs:323-344): 500x500, depth 20, writeImage; oracle at 12 spp."""
    g = _golden("pawn_demo")
    s, w = scenes.pawn(spp=12)
    ora = OracleScene(w)
    img, _ = ora.render(s, 55)
    assert img.shape == g.shape == (500, 500, 3)
    ids, _, stable = ora.probe_stable(s, delta=0.5, t_tol=float("inf"))
    bg = (ids == 0xFFFFFFFF) & stable
    # background = grayFade, deterministic up to the sub-pixel jitter: silhouette + gradient pin the camera
    o8 = quantise8(srgb_oetf(img)).astype(np.int32)
    assert bg.sum() > 100000
    diff = np.abs(g[bg].astype(np.int32) - o8[bg]).max(axis=1)
    assert (diff <= 1).mean() > 0.999, (diff <= 1).mean()     # the rest: triangle tips inside pixel corners
    # quantiser pin: floor(256 y) reproduces the golden exactly on most background pixels even at 12 spp of
    # jitter noise; round(255 y) does not (81 % when converged)
    assert (diff == 0).mean() > 0.9, (diff == 0).mean()
    r255 = np.rint(srgb_oetf(img) * 255).astype(np.int32)
    assert (np.abs(g[bg].astype(np.int32) - r255[bg]).max(axis=1) == 0).mean() < 0.87
    gl = srgb_eotf(g.astype(np.float64) / 255)
    obj = ~bg
    # the red subsurface glow: mean colour over the pawn within noise
    mg, mo = gl[obj].mean(axis=0), np.clip(img, 0, 1)[obj].mean(axis=0)
    assert np.abs(mg - mo).max() < 0.02, (mg, mo)
    d = block_mean(gl, 20) - block_mean(np.clip(img, 0, 1), 20)
    assert np.abs(d).mean() < 0.01


def test_sphere_uv_known_answer(built):
    """sphereUVTest (test/Main.hs:128-134), the one fully deterministic reference program: the nearest sphere is
    hit at outward normal (0,0,1) -> (u,v) = (0.5,0.5) -> centre texel of earthmap."""
    s, w = scenes.sphere_uv_test()
    ora = OracleScene(w)
    h = ora.hit((0, 0, 0), (0, 0, -1))
    assert h["prim"] == 1 and abs(h["t"] - 0.6) < 1e-12
    assert abs(h["uv"][0] - 0.5) < 1e-12 and abs(h["uv"][1] - 0.5) < 1e-12
    earth = scenes.load_earthmap()
    img, _ = ora.render(s, 12, spp=64)
    texel = earth[256, 512].astype(np.float64)
    # one lambertian bounce under a white background: pixel = texel colour * 1
    assert np.allclose(img[0, 0], texel, atol=2e-2)


def _high_pass(a, mask, r=12):
    """a divided by its box-blurred self inside the mask: removes the smooth shading, keeps the texture detail"""
    from scipy.ndimage import uniform_filter
    w = uniform_filter(mask.astype(np.float64), 2 * r + 1)
    sm = uniform_filter(np.where(mask, a, 0.0), 2 * r + 1) / np.maximum(w, 1e-9)
    return a / np.maximum(sm, 1e-6)


def test_demo2_matches_golden(built):
    """demo2 (test/Main.hs:259-321; README's demo2.png: 800x800, writeImageSqrt).  The 400 boxes and 1000 small balls come
    from the reference's unreproducible RNG stream, but the camera, the light, and the seven large objects are deterministic,
    and they pin `scenes.demo2` (so far checked only against an oracle reading the same scenes.py):
      * the light parallelogram: every pixel whose centre sees it is saturated in the golden (exact, 36 k pixels);
      * the earth sphere (imageTexture under transform (translate . rotateY (pi/2)), sphereUV, Texture.hs:31-41): the
        golden's detail correlates with the deterministic texture lookup at the primary hit, and NOT with a mirrored one;
      * every large object's mean linear colour agrees with a low-spp oracle render.
    What the golden does NOT pin, measured the same way: the marble sphere's veins.  Their high-pass correlation with the
    library's default gradient table (SplitMix restatement of mkStdGen 666, Noise.hs:88-92) is ~0.17 - the same as with a
    random table (~0.14; what is left is the table-independent sine stripe).  The Perlin gradient table therefore stays
    "parity unpinned" and an ABI input (rt_scene_set_noise_tables)."""
    from scipy.ndimage import binary_erosion
    g = _golden("demo2")
    assert g.shape == (800, 800, 3)
    gl = (g.astype(np.float64) / 255) ** 2                     # undo writeImageSqrt
    lum = lambda a: 0.2126 * a[..., 0] + 0.7152 * a[..., 1] + 0.0722 * a[..., 2]   # noqa: E731
    s, w = scenes.demo2(width=800, spp=1, depth=50)
    ora = OracleScene(w)
    ids, _, stable = ora.probe_stable(s, delta=0.5, t_tol=float("inf"))
    LIGHT, MOVING, GLASS, BOUNDARY, METAL, EARTH, MARBLE = range(3400, 3407)      # canonical ids of largeObjects, in list order
    light = (ids == LIGHT) & stable
    assert light.sum() > 36000 and (g[light] == 255).all()
    rows, cols = np.where(light)
    assert (rows.min(), cols.min()) == (0, 100) and abs(int(rows.max()) - 116) <= 1 and abs(int(cols.max()) - 532) <= 1
    # the saturated region of the golden's top rows IS the light (plus its partially covered rim)
    sat_top = (g[:110] == 255).all(axis=2)
    grown = np.zeros_like(sat_top)
    lm = (ids[:110] == LIGHT)
    for dj in (-1, 0, 1):
        for di in (-1, 0, 1):
            grown |= np.roll(np.roll(lm, dj, 0), di, 1)
    assert (sat_top & ~grown).sum() <= 40
    # earth: texture orientation / instance transform / sphereUV
    surf = ora.probe_surface(s)
    earth = (ids == EARTH) & stable
    core = binary_erosion(earth, iterations=14)
    assert core.sum() > 15000
    tex_l = lum(surf[..., 6:9])
    a = _high_pass(lum(gl), earth)[core]
    c_true = np.corrcoef(a, _high_pass(tex_l, earth)[core])[0, 1]
    c_mirror = np.corrcoef(a, _high_pass(tex_l[:, ::-1], earth)[core])[0, 1]
    assert c_true > 0.6 and abs(c_mirror) < 0.2, (c_true, c_mirror)
    # mean colours of the deterministic objects against a low-spp oracle render (half resolution, golden 2x2-averaged)
    s2, w2 = scenes.demo2(width=400, spp=16, depth=50)
    img, _ = OracleScene(w2).render(s2, 1234)
    oc = np.clip(img, 0, 1)
    g2 = gl.reshape(400, 2, 400, 2, 3).mean(axis=(1, 3))
    for pid, tol in ((MOVING, 0.01), (GLASS, 0.015), (BOUNDARY, 0.01), (METAL, 0.02), (EARTH, 0.01), (MARBLE, 0.02)):
        m = ((ids == pid) & stable).reshape(400, 2, 400, 2).all(axis=(1, 3))
        assert m.sum() > 2000
        assert np.abs(g2[m].mean(axis=0) - oc[m].mean(axis=0)).max() < tol, (pid, g2[m].mean(axis=0), oc[m].mean(axis=0))


def test_demo1_matches_golden(built):
    """demo1 (test/Main.hs:136-186; README's demo1.png: 1200x675, writeImageSqrt).  The ~480 small spheres come from
    `newStdGen`, but the camera (vfov 20, defocus 0.6 deg at focus distance 10), the sky, the ground and the three big
    spheres - glass, lambertian (0.4, 0.2, 0.1), mirror (0.7, 0.6, 0.5) - are deterministic, and they pin `scenes.demo1`:
      * the big spheres' silhouettes: where the oracle's primary rays see them, the golden shows them (edge positions);
      * the horizon row (camera pitch, ground sphere of radius 1000);
      * mean linear colours of sky, brown, mirror and glass sphere against a 16-spp oracle render (the small spheres only
        enter through reflections and bounce light: 1 - 3 %)."""
    from scipy.ndimage import binary_erosion
    g = _golden("demo1")
    assert g.shape == (675, 1200, 3)
    gl = (g.astype(np.float64) / 255) ** 2                     # undo writeImageSqrt
    lum = lambda a: 0.2126 * a[..., 0] + 0.7152 * a[..., 1] + 0.0722 * a[..., 2]   # noqa: E731
    s, w = scenes.demo1(width=1200, spp=1)
    ids, _, stable = OracleScene(w).probe_stable(s, delta=0.5, t_tol=float("inf"))
    GROUND, GLASS, BROWN, MIRROR, MISS = 0, 1, 2, 3, 0xFFFFFFFF           # bigSpheres come first in the bvhTree list
    # silhouettes (pixel-centre rays; the lens blurs the golden's edges by a few pixels)
    L = lum(gl)
    for x in (50, 150, 250, 1100):                                         # horizon: first non-sky row of the golden / first hit row
        assert abs(int(np.argmax(L[:, x] < 0.55)) - int(np.argmax(ids[:, x] != MISS))) <= 3
    assert abs(int(np.argmax(gl[120, :, 2] < 0.5)) - int(np.argmax(ids[120] == BROWN))) <= 3    # left edge of the brown sphere against the sky
    top = int(np.argmax((ids == MIRROR).any(axis=1)))
    x_top = int(np.median(np.where(ids[top] == MIRROR)[0]))
    assert abs(int(np.argmax(L[:, x_top] < 0.8)) - top) <= 4                # top of the mirror sphere against the sky
    # colours: half resolution, golden 2x2-averaged in linear space
    s2, w2 = scenes.demo1(width=600, spp=16)
    img, _ = OracleScene(w2).render(s2, 1)
    g2 = gl[:674].reshape(337, 2, 600, 2, 3).mean(axis=(1, 3))
    ids2, img = ids[:674:2, ::2], img[:337]
    upper = np.arange(337)[:, None] < 60
    for name, k, it, tol, extra in (("sky", MISS, 6, 0.02, upper), ("brown", BROWN, 8, 0.06, True), ("mirror", MIRROR, 8, 0.04, True), ("glass", GLASS, 8, 0.06, True)):
        m = binary_erosion(ids2 == k, iterations=it) & extra
        assert m.sum() > 1000, name
        a, b = g2[m].mean(axis=0), img[m].mean(axis=0)
        assert (np.abs(a - b) <= tol * a).all(), (name, a, b)
