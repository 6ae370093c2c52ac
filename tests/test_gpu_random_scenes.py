"""Randomised differential test of the flattener + traversal + hit reconstruction against the CPU oracle (-m gpu).

Every scene is a random nesting of the reference's combinators - group / bvhTree / bvhNode over spheres, parallelograms,
triangles and cuboids, under chains of Euclidean transforms and `moving`, with materials attached at random levels (the
outermost `<$` wins, Geometry.hs:44-47) - built from a fixed seed, so a failure reproduces.  Checked deterministically (no
Monte Carlo): the all-pixel primitive-id + distance gate at three shutter times, for the SAH tree and the reference-shaped
tree, and the shading inputs (normal, uv, front flag, material, texture colour) at the primary hit; then one low-spp render
for path-length statistics and luminance."""
import math

import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene, primary_gate
from util import MISS, luminance

pytestmark = pytest.mark.gpu


def _random_world(seed: int):
    rng = np.random.default_rng(seed)
    U = lambda lo, hi: float(rng.uniform(lo, hi))                       # noqa: E731
    P = lambda r=3.0: (U(-r, r), U(-1.5, 2.0), U(-7.0, -3.0))           # noqa: E731  (in front of the default camera)

    def texture():
        k = int(rng.integers(0, 4))
        if k == 0:
            return rt.constantTexture((U(0.1, 0.9), U(0.1, 0.9), U(0.1, 0.9)))
        if k == 1:
            return rt.checkerTexture(int(rng.integers(2, 9)), int(rng.integers(2, 9)), (U(0, 0.4),) * 3, (U(0.6, 1),) * 3)
        if k == 2:
            return rt.noiseTexture(int(rng.integers(1, 5)), U(0.5, 3), (U(-1, 1), U(-1, 1), U(-1, 1)), (0.1, 0.1, 0.3), (0.9, 0.8, 0.6))
        return rt.marbleTexture((0, 0, U(0.5, 2)), U(1, 4), (U(-1, 1), 0, 0))

    def material():
        k = int(rng.integers(0, 6))
        return [lambda: rt.lambertian(texture()), lambda: rt.metal(U(0, 0.6), texture()), lambda: rt.mirror(texture()),
                lambda: rt.dielectric(U(1.2, 1.8)), lambda: rt.lommelSeeliger(texture()), lambda: rt.lightSource(texture())][k]()

    def primitive():
        k = int(rng.integers(0, 4))
        if k == 0:
            return rt.sphere(P(), U(0.2, 0.9))
        if k == 1:
            return rt.parallelogram(P(), (U(-1.5, 1.5), U(-0.5, 0.5), U(-0.5, 0.5)), (U(-0.5, 0.5), U(0.5, 1.5), U(-0.5, 0.5)))
        if k == 2:
            a = P()
            return rt.triangle((a, (0, 0)), ((a[0] + U(0.5, 2), a[1] + U(-0.5, 0.5), a[2] + U(-0.5, 0.5)), (1, 0)),
                               ((a[0] + U(-0.5, 0.5), a[1] + U(0.5, 2), a[2] + U(-0.5, 0.5)), (0.5, 1)))
        lo = P()
        return rt.cuboid(rt.fromCorners(lo, (lo[0] + U(0.3, 1.2), lo[1] + U(0.3, 1.2), lo[2] + U(0.3, 1.2))))

    def euclid():
        m = rt.translate((U(-0.6, 0.6), U(-0.4, 0.4), U(-0.5, 0.5)))
        for rot in (rt.rotateX, rt.rotateY, rt.rotateZ):
            if rng.random() < 0.6:
                m = m @ rot(U(-0.5, 0.5))
        return m

    def subtree(depth, have_material):
        """geometry with a material on EVERY leaf path by the time the root is reached"""
        r = rng.random()
        if depth == 0 or r < 0.35:
            g = primitive()
        elif r < 0.55:
            g = rt.group([subtree(depth - 1, have_material) for _ in range(int(rng.integers(2, 5)))])
        elif r < 0.75:
            g = rt.bvhTree([subtree(depth - 1, have_material) for _ in range(int(rng.integers(2, 7)))])
        elif r < 0.85:
            g = rt.bvhNode(subtree(depth - 1, have_material), subtree(depth - 1, have_material))
        elif r < 0.95:
            g = rt.transform(euclid(), subtree(depth - 1, have_material))
        else:
            g = rt.moving((0, 0, 0), (U(-0.5, 0.5), U(-0.3, 0.3), 0), subtree(depth - 1, have_material))
        return g

    def with_materials(depth):
        """a material is attached at a random level of every branch; a few branches get a second, outer one (which wins)"""
        r = rng.random()
        if depth == 0 or r < 0.4:
            return material() << subtree(depth, True)
        kids = [with_materials(depth - 1) for _ in range(int(rng.integers(2, 5)))]
        g = rt.group(kids) if rng.random() < 0.5 else rt.bvhTree(kids)
        if rng.random() < 0.25:
            g = material() << g                      # outermost <$ wins over everything below
        if rng.random() < 0.3:
            g = rt.transform(euclid(), g)
        return g

    # the root is always a group of several branches (a lone leaf would be a trivial scene)
    return rt.group([with_materials(2) for _ in range(int(rng.integers(4, 8)))])


SEEDS = [11, 23, 37, 58, 71, 96]


@pytest.mark.parametrize("seed", SEEDS)
def test_random_scene_against_the_oracle(built, seed):
    world = _random_world(seed)
    s = rt.defaultCameraSettings.replace(cs_imageWidth=240, cs_aspectRatio=1.5, cs_vfov=math.radians(70), cs_samplesPerPixel=32,
                                         cs_maxRecursionDepth=10, cs_background=scenes.sky)
    ora = OracleScene(world)
    trees = [CompiledScene(world), CompiledScene(world, options=dict(bvh_reference=1)), CompiledScene(world, options=dict(split_moving=1, aggregate_boxes=0))]
    n_prims = trees[0].info()["n_prims"]
    assert n_prims == ora.num_prims() and n_prims >= 2
    assert np.array_equal(trees[0].handle.get_eval_order()[: len(ora.leaf_order())], ora.leaf_order())
    hit_frac = 0.0
    for time in (0.0, 0.4, 1.0):
        ids5, ts5 = ora.probe_tilted(s, time=time)
        for cs in trees:
            ids, ts = cs.probe_primary(s, time=time)
            g = primary_gate(ids, ts, ids5, ts5)
            assert g["pass_frac"] >= 0.9995, (seed, time, {k: v for k, v in g.items() if k != "passed_mask"})
            assert g["id_unexplained"] <= 3
        hit_frac = max(hit_frac, float((ids5[0] != MISS).mean()))
    assert hit_frac > 0.05                                   # the camera does see the scene
    # shading inputs at the primary hit
    cs = trees[0]
    ids, _ = cs.probe_primary(s)
    oids, _, stable = ora.probe_stable(s)
    gsurf, osurf = cs.probe_surface(s).astype(np.float64), ora.probe_surface(s)
    ok = stable & (ids == oids) & (oids != MISS)
    if ok.sum() > 50:
        assert (gsurf[ok][:, 9] == osurf[ok][:, 9]).all()                                   # material kind: outermost <$ wins
        assert (gsurf[ok][:, 5] == osurf[ok][:, 5]).mean() > 0.999                          # front flag
        assert np.abs(gsurf[ok][:, 0:3] - osurf[ok][:, 0:3]).max() < 5e-4                   # normal
        du = np.abs(gsurf[ok][:, 3] - osurf[ok][:, 3])
        assert np.minimum(du, 1 - du).max() < 5e-4 and np.abs(gsurf[ok][:, 4] - osurf[ok][:, 4]).max() < 5e-4
        dc = np.abs(gsurf[ok][:, 6:9] - osurf[ok][:, 6:9]).max(axis=1)
        assert (dc > 5e-3).mean() < 5e-3                                                    # checker cells may flip on a border
    # one low-spp render: same path-length statistics, close luminance, finite
    img, st = cs.render(s, rt.mkStdGen(seed), want_stats=True)
    ref, cnt = ora.render(s, seed)
    assert np.isfinite(img).all()
    assert abs(st["rays"] / st["samples"] - cnt["rays"] / cnt["samples"]) <= 0.02 * cnt["rays"] / cnt["samples"]
    lg, lo = luminance(img.astype(np.float64)).mean(), luminance(ref).mean()
    assert abs(lg - lo) <= 0.03 * lo + 1e-3, (lg, lo)
