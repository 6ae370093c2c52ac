"""More -m gpu parity cases: the converged-image gate of the north_star (4096 spp, 1 % mean luminance, 4-sigma pixels)
at resolutions the oracle renders in seconds, and the edge cases of the path (degenerate sizes, depth limits,
empty sample ranges, tiny pools, concurrent renders)."""
import threading

import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene
from util import luminance, z_scores

pytestmark = pytest.mark.gpu

CONVERGED = [
    ("readme", dict(width=64)),
    ("cornell", dict(width=40)),
    ("cornell_noredirect", dict(width=32)),
    ("pawn", dict(width=40)),
]


@pytest.mark.parametrize("name,kw", CONVERGED, ids=[c[0] for c in CONVERGED])
def test_converged_4096spp_image(built, name, kw):
    """north_star gate 3: 4096-spp images agree within 1 % relative mean luminance; no pixel fails the 4-sigma
    Monte Carlo variance test (confirmed on an independent re-seed)."""
    spp = 4096
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(spp=spp, **kw)
    cs, ora = CompiledScene(w), OracleScene(w)
    img, sq = cs.render(s, rt.mkStdGen(seed), want_sumsq=True)
    ref, rsq, _ = ora.render(s, seed, want_sumsq=True)
    img, sq = img.astype(np.float64), sq.astype(np.float64)
    lg, lo = luminance(img).mean(), luminance(ref).mean()
    print(f"{name}: 4096 spp mean luminance gpu {lg:.5f} oracle {lo:.5f} ({(lg / lo - 1) * 100:+.3f} %)")
    assert abs(lg - lo) <= 0.01 * lo
    z1 = z_scores(img, sq, spp, ref, rsq, spp)
    fail1 = np.abs(z1) > 4
    assert fail1.mean() < 2e-3
    if fail1.any():
        img2, sq2 = cs.render(s, rt.mkStdGen(seed + 101), want_sumsq=True)
        ref2, rsq2, _ = ora.render(s, seed + 101, want_sumsq=True)
        z2 = z_scores(img2.astype(np.float64), sq2.astype(np.float64), spp, ref2, rsq2, spp)
        confirmed = fail1 & (np.abs(z2) > 4) & (np.sign(z1) == np.sign(z2))
        assert not confirmed.any(), np.argwhere(confirmed)[:10]
    # block means (8x8 pixels) agree to a few standard errors of their own Monte Carlo noise
    b = 8
    h, wd = img.shape[0] // b * b, img.shape[1] // b * b
    blk = lambda a: a[:h, :wd].reshape(h // b, b, wd // b, b, -1).mean(axis=(1, 3))   # noqa: E731
    var_px = np.maximum(sq / spp - img ** 2, 0) / spp + np.maximum(rsq / spp - ref ** 2, 0) / spp
    se_blk = np.sqrt(blk(var_px) / (b * b)) + 1e-3 * np.maximum(blk(ref), 1e-2)
    assert (np.abs(blk(img) - blk(ref)) / se_blk).max() < 6.0


def test_degenerate_sizes_and_depths(built):
    s, w = scenes.readme_example(width=1, spp=3)
    s = s.replace(cs_aspectRatio=1.0)
    img = CompiledScene(w).render(s, rt.mkStdGen(1))
    assert img.shape == (1, 1, 3) and np.isfinite(img).all()
    # depth 1: every path is one segment; hits on non-emitters contribute nothing, misses see the background
    s, w = scenes.readme_example(width=64, spp=16)
    cs, ora = CompiledScene(w), OracleScene(w)
    d1 = s.replace(cs_maxRecursionDepth=1)
    img, st = cs.render(d1, rt.mkStdGen(5), want_stats=True)
    ref, cnt = ora.render(d1, 5)
    assert st["rays"] == st["samples"] == cnt["rays"]
    assert abs(img.mean() - ref.mean()) < 0.01 * ref.mean()
    # depth 0: rayColor 0 = 0 (Ray.hs:176) — black image, no rays
    d0 = s.replace(cs_maxRecursionDepth=0)
    ref0, cnt0 = ora.render(d0, 5)
    img0, st0 = cs.render(d0, rt.mkStdGen(5), want_stats=True)
    assert ref0.max() == 0.0 and img0.max() == 0.0
    # empty sample range: zeros, nothing launched beyond the normalisation
    z = cs.render(s, rt.mkStdGen(5), first_sample=7, n_samples=0)
    assert z.shape == img.shape and (z == 0).all()


def test_tiny_pool_and_ragged_ranges(built):
    """Pool far smaller than the image (many iterations, constant refill) and sample ranges that do not divide spp."""
    s, w = scenes.cornell_box(spp=30, width=50)
    cs = CompiledScene(w)
    full = cs.render(s, rt.mkStdGen(3))
    cs.handle.set_pool_size(1024)
    small = cs.render(s, rt.mkStdGen(3))
    assert np.allclose(small, full, rtol=2e-4, atol=2e-5)
    cs.handle.set_pool_size(0)
    parts = sum(cs.render(s, rt.mkStdGen(3), first_sample=f, n_samples=n) for f, n in [(0, 7), (7, 1), (8, 22)])
    assert np.allclose(parts, full, rtol=2e-4, atol=2e-5)
    # sample indices beyond 2^32 keep working (64-bit counter)
    hi = cs.render(s.replace(cs_samplesPerPixel=4), rt.mkStdGen(3), first_sample=(1 << 33) + 5, n_samples=4)
    assert np.isfinite(hi).all() and abs(hi.mean() - full.mean()) < 0.25 * full.mean()
    # a range that crosses a multiple of 2^32 equals its two halves rendered separately
    s8 = s.replace(cs_samplesPerPixel=8)
    cross = cs.render(s8, rt.mkStdGen(3), first_sample=(1 << 32) - 3, n_samples=8)
    lo = cs.render(s8, rt.mkStdGen(3), first_sample=(1 << 32) - 3, n_samples=3)
    hi2 = cs.render(s8, rt.mkStdGen(3), first_sample=(1 << 32), n_samples=5)
    assert np.allclose(cross, lo + hi2, rtol=2e-4, atol=2e-5)


def test_concurrent_renders_from_threads(built):
    """A committed scene is immutable and may be rendered from several OS threads at once (GHC safe calls migrate
    threads): every call borrows its own workspace; results equal the serial ones."""
    s, w = scenes.readme_example(width=96, spp=32)
    cs = CompiledScene(w)
    serial = [cs.render(s, rt.mkStdGen(k)) for k in range(4)]
    out = [None] * 4

    def work(k):
        out[k] = cs.render(s, rt.mkStdGen(k))

    ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    for a, b in zip(serial, out):
        assert np.allclose(a, b, rtol=2e-4, atol=2e-5)


def test_media_only_and_surface_only_shortcuts(built):
    """constantMedium whose boundary is ALSO a surface (shared-boundary shortcut) against the same medium with a
    separate, geometrically identical boundary object: identical distributions."""
    glass, fog = rt.dielectric(1.5), rt.isotropic(rt.constantTexture((0.9, 0.3, 0.2)))
    ball = rt.sphere((0, 0, -2), 0.8)
    shared = rt.group([rt.pureGeometry(glass << ball), fog << rt.constantMedium(2.0, ball)])
    s = rt.defaultCameraSettings.replace(cs_imageWidth=64, cs_samplesPerPixel=256, cs_background=scenes.sky, cs_maxRecursionDepth=12)
    cs, ora = CompiledScene(shared), OracleScene(shared)
    img, sq, st = cs.render(s, rt.mkStdGen(2), want_sumsq=True, want_stats=True)
    ref, rsq, cnt = ora.render(s, 2, want_sumsq=True)
    assert abs(luminance(img).mean() - luminance(ref).mean()) < 0.01 * luminance(ref).mean()
    assert abs(st["rays"] / st["samples"] - cnt["rays"] / cnt["samples"]) < 0.01 * cnt["rays"] / cnt["samples"]
    z = z_scores(img.astype(np.float64), sq.astype(np.float64), 256, ref, rsq, 256)
    assert (np.abs(z) > 4).mean() < 2e-3


SURFACE_CASES = [
    ("readme", scenes.readme_example, dict()),                        # checker on a sphere (sphereUV), mirror parallelogram
    ("noise_test", scenes.noise_test, dict(width=200, spp=1)),        # noiseTexture + marbleTexture: Perlin lattice, 7 octaves
    ("cuboid_test", scenes.cuboid_test, dict(width=200, spp=1)),      # image texture on cuboid faces under a transform
    ("demo2", scenes.demo2, dict(width=400, spp=1)),                  # earth (image, rotated sphere), marble, instanced boxes
    ("bunny_test", scenes.bunny_test, dict(width=200, spp=1)),        # triangle normals / front flags
]


@pytest.mark.parametrize("name,ctor,kw", SURFACE_CASES, ids=[c[0] for c in SURFACE_CASES])
def test_surface_probe_pointwise(built, name, ctor, kw):
    """Deterministic parity of the hit record and the textures (no Monte Carlo): per pixel-centre primary hit the
    GPU's normal, uv, front flag, material kind and texture colour against the oracle's, on well-conditioned pixels."""
    s, w = ctor(**kw)
    cs, ora = CompiledScene(w), OracleScene(w)
    ids, _ = cs.probe_primary(s)
    oids, _, stable = ora.probe_stable(s)
    g, o = cs.probe_surface(s).astype(np.float64), ora.probe_surface(s)
    ok = stable & (ids == oids) & (oids != 0xFFFFFFFF)
    assert ok.mean() > 0.05
    assert (g[ok][:, 9] == o[ok][:, 9]).all()                                    # material kind
    assert (g[ok][:, 5] == o[ok][:, 5]).mean() > 0.9999                          # front flag
    assert np.abs(g[ok][:, 0:3] - o[ok][:, 0:3]).max() < 2e-4                    # unit normal
    du = np.abs(g[ok][:, 3] - o[ok][:, 3])
    du = np.minimum(du, 1.0 - du)                                                # sphere u wraps at the seam
    assert du.max() < 2e-4 and np.abs(g[ok][:, 4] - o[ok][:, 4]).max() < 2e-4   # uv
    # textures: continuous ones (noise, marble, constant) agree closely everywhere; piecewise-constant ones (checker,
    # image texels) may flip on a cell border within fp32 uv error, on a vanishing fraction of pixels
    dc = np.abs(g[ok][:, 6:9] - o[ok][:, 6:9]).max(axis=1)
    print(f"{name}: {int(ok.sum())} px, texture colour max diff {dc.max():.2e}, > 1e-3 on {(dc > 1e-3).mean() * 100:.4f} %")
    assert (dc > 2e-3).mean() < 2e-3
    assert np.median(dc) < 1e-4


def _cuboid_zoo(tree):
    """Cuboids in every role the reference allows: glass (rays run inside and leave through the far face), rotated,
    moving, boundary of a constantMedium, with and without a bvhTree above them."""
    V3 = rt.V3
    glass, fog = rt.dielectric(1.5), rt.isotropic(rt.constantTexture((0.2, 0.4, 0.9)))
    white = rt.lambertian(rt.constantTexture((0.73, 0.73, 0.73)))
    red = rt.lambertian(rt.constantTexture((0.65, 0.05, 0.05)))
    light = rt.lightSource(rt.constantTexture((7, 7, 7)))
    box = lambda a, b: rt.cuboid(rt.fromCorners(V3(*a), V3(*b)))
    parts = [
        rt.pureGeometry(white << rt.parallelogram(V3(-6, -1, -9), V3(12, 0, 0), V3(0, 0, 12))),              # floor
        rt.pureGeometry(light << rt.parallelogram(V3(-1.5, 3.5, -5), V3(3, 0, 0), V3(0, 0, 2))),
        rt.pureGeometry(glass << box((-2.6, -1, -4.2), (-1.4, 0.4, -3.0))),
        rt.pureGeometry(rt.transform(rt.translate(V3(0.2, -1, -4.6)) @ rt.rotateY(0.5), red << box((0, 0, 0), (1.0, 1.8, 1.0)))),
        rt.pureGeometry(rt.moving(V3(0, 0, 0), V3(0.4, 0.3, 0), white << box((1.8, -1, -3.6), (2.6, -0.2, -2.8)))),
        fog << rt.constantMedium(1.5, box((-0.8, -1, -2.6), (0.6, 0.2, -1.6))),
    ]
    world = rt.bvhTree(parts) if tree else rt.group(parts)
    s = rt.defaultCameraSettings.replace(cs_imageWidth=96, cs_aspectRatio=1.5, cs_samplesPerPixel=512, cs_maxRecursionDepth=12,
                                         cs_background=scenes.sky, cs_center=V3(0, 1.0, 2.5), cs_lookAt=V3(0, 0, -3.5), cs_vfov=0.9)
    return s, world


@pytest.mark.parametrize("tree", [False, True], ids=["group", "bvhTree"])
def test_aggregated_cuboids(built, tree):
    """One slab test per cuboid (PRIM_FLAG_BOXHEAD) must pick the same face at the same distance as six independent
    parallelogram tests (RtOptions.aggregate_boxes = 0, the reference's formulation) and converge to the oracle's image."""
    s, world = _cuboid_zoo(tree)
    agg = CompiledScene(world)
    faces = CompiledScene(world, options=dict(aggregate_boxes=0))
    probe_s = s.replace(cs_imageWidth=480)
    for t in (0.0, 0.5, 1.0):
        ia, ta = agg.probe_primary(probe_s, time=t)
        ib, tb = faces.probe_primary(probe_s, time=t)
        same = ia == ib
        assert same.mean() > 0.999, f"face ids differ on {(~same).sum()} pixels at time {t}"      # edges only
        hit = same & (ia != 0xFFFFFFFF)
        assert hit.any() and np.all(np.abs(ta[hit] - tb[hit]) <= 1e-5 * np.abs(tb[hit]))
    ora = OracleScene(world)
    img, sq, st = agg.render(s, rt.mkStdGen(5), want_sumsq=True, want_stats=True)
    ref, rsq, cnt = ora.render(s, 5, want_sumsq=True)
    assert abs(luminance(img).mean() - luminance(ref).mean()) < 0.01 * luminance(ref).mean()
    assert abs(st["rays"] / st["samples"] - cnt["rays"] / cnt["samples"]) < 0.01 * cnt["rays"] / cnt["samples"]
    z = z_scores(img.astype(np.float64), sq.astype(np.float64), 512, ref, rsq, 512)
    assert (np.abs(z) > 4).mean() < 2e-3


TAIL_CASES = [
    ("cornell", dict(width=37, spp=24)),               # flat scene: k_finish_flat takes the whole tail
    ("demo1", dict(width=45, spp=8)),                  # BVH, depth 50: generic k_finish below 1 Mi paths
    ("pawn", dict(width=35, spp=8)),                   # BVH + subsurface medium, depth 20: no tail kernel by default
]


@pytest.mark.parametrize("name,kw", TAIL_CASES, ids=[c[0] for c in TAIL_CASES])
def test_tail_kernel_and_tiled_start_order_do_not_change_the_samples(built, name, kw):
    """The tail kernel (lanes carry a path to its end and claim the next one) and the 8x4-tiled order in which camera
    paths start only reschedule work: the RNG is keyed by (pixel, sample, bounce), so every variant renders the same
    sample set and images differ by fp32 summation order alone.  Image sizes are deliberately not multiples of the
    tile (ragged last band / last tile)."""
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    cs = CompiledScene(w)
    base, st0 = cs.render(s, rt.mkStdGen(seed), want_stats=True)
    assert np.isfinite(base).all()
    # primary_cull: camera rays that miss the world box end on the background inside k_raygen (pawn: the camera
    # looks at the object from outside) or go through the queue like every other ray - the same paths either way
    ALL = 4294967295
    variants = [dict(finish_max=0), dict(finish_max=ALL), dict(finish_max=300, tail_batch=1),
                dict(tile=0), dict(tile=1), dict(finish_refill=32, finish_max=100000),
                dict(primary_cull=0), dict(primary_cull=1), dict(primary_cull=1, finish_max=ALL),
                dict(shade_sort=0), dict(shade_sort=1), dict(count_work=1),
                dict(primary_cull=1, raygen_block=1), dict(primary_cull=1, raygen_block=3), dict(primary_cull=1, raygen_block=5, pool=2048)]
    for env in variants:
        cs.handle.set_options(**env)
        img, st = cs.render(s, rt.mkStdGen(seed), want_stats=True)
        cs.handle.set_options(**{k: None for k in env})
        assert st["samples"] == base.shape[0] * base.shape[1] * s.cs_samplesPerPixel
        assert st["rays"] == st0["rays"], env          # the same path segments, however they are scheduled
        assert np.allclose(img, base, rtol=2e-4, atol=2e-5), env


def test_page_locked_output_buffers(built):
    """rt_render writes page-locked caller buffers (rt_host_alloc) directly and pageable ones through its staging
    buffer; both hold the same image.  Freed buffers are recycled; foreign pointers are refused."""
    import ctypes as C
    from raytrace_b200 import _ffi
    s, w = scenes.readme_example(width=80, spp=8)
    cs = CompiledScene(w)
    pinned, sq = cs.render(s, rt.mkStdGen(5), want_sumsq=True)
    pageable, sq2 = cs.render(s, rt.mkStdGen(5), want_sumsq=True, pinned_out=False)
    assert np.allclose(pinned, pageable, rtol=2e-4, atol=2e-5) and np.allclose(sq, sq2, rtol=2e-4, atol=2e-5)
    keep = pinned.copy()
    del pinned, sq                     # buffers go back to the library's free list ...
    again = cs.render(s, rt.mkStdGen(5))
    assert np.allclose(again, keep, rtol=2e-4, atol=2e-5)
    L = _ffi.lib()
    p, q = C.c_void_p(), C.c_void_p()
    assert L.rt_host_alloc(7654321, C.byref(p)) == 0 and p.value
    assert L.rt_host_free(p) == 0
    assert L.rt_host_alloc(7654000, C.byref(q)) == 0 and q.value == p.value      # ... and are handed out again
    assert L.rt_host_free(q) == 0
    assert L.rt_host_free(q) == 1      # double free / foreign pointer: RT_ERR_INVALID_ARG, nothing is touched


@pytest.mark.parametrize("name,kw", [("cornell", dict(width=64, spp=16)), ("demo1", dict(width=64, spp=8))], ids=["flat", "bvh"])
def test_render_device_is_stream_ordered(built, name, kw):
    """rt_render_device accumulates into a caller-owned device buffer on the caller's stream (NULL = default stream).
    Without statistics a flat-scene render returns before the GPU has finished (single-shot launch sequence), so
    whatever the caller queues next on that stream - here the scale and the copy - must still see the finished sums;
    back-to-back renders reuse one workspace and must not disturb each other."""
    import torch
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    cs = CompiledScene(w)
    want = cs.render(s, rt.mkStdGen(seed)).astype(np.float64)
    spp = s.cs_samplesPerPixel
    dev = torch.device("cuda", 0)
    n = want.size
    side = torch.cuda.Stream(dev)
    for stream in (torch.cuda.current_stream(dev), side):
        with torch.cuda.stream(stream):
            accs = [torch.zeros(n, dtype=torch.float32, device=dev) for _ in range(3)]
            for a in accs:                                   # three renders queued back to back, nothing read back
                assert cs.render_device(s, rt.mkStdGen(seed), 0, spp, a.data_ptr(), 0, stream.cuda_stream, want_stats=False) is None
            imgs = [(a * (1.0 / spp)).cpu().numpy() for a in accs]
        for img in imgs:
            assert np.allclose(img.reshape(want.shape), want, rtol=2e-4, atol=2e-5)
    a = torch.zeros(n, dtype=torch.float32, device=dev)
    st = cs.render_device(s, rt.mkStdGen(seed), 0, spp, a.data_ptr(), 0, 0)           # with statistics: synchronous
    assert st["samples"] == want.shape[0] * want.shape[1] * spp and st["rays"] >= st["samples"]
    assert np.allclose((a / spp).cpu().numpy().reshape(want.shape), want, rtol=2e-4, atol=2e-5)


def test_interior_queries_bounded_by_free_flight_distance(built):
    """A path that has just scattered inside the scene's only medium, whose boundary is a closed surface shared with the
    surfaces (pawn; a glass ball filled with fog), draws its free-flight distance BEFORE the traversal and bounds the
    surface query with it.  The counter RNG hands out the same number either way, so the event sequence - and the image
    at equal seeds - is that of the unbounded queries (RtOptions.interior_bound = 0 at commit), up to summation order."""
    from raytrace_b200 import V3, constantTexture, dielectric, isotropic, group, pureGeometry, constantMedium, sphere, lambertian
    ball = sphere((0, 0, -3), 1.5)
    fog_ball = group([pureGeometry(dielectric(1.5) << ball), isotropic(constantTexture(V3(0.9, 0.6, 0.3))) << constantMedium(3.0, ball),
                      pureGeometry(lambertian(constantTexture(V3(0.5, 0.5, 0.5))) << sphere((0, -101.5, -3), 100))])
    s_ball = rt.defaultCameraSettings.replace(cs_imageWidth=72, cs_samplesPerPixel=32, cs_maxRecursionDepth=30, cs_background=scenes.sky)
    s_pawn, w_pawn = scenes.pawn(width=56, spp=24)
    for s, w, seed in ((s_ball, fog_ball, 7), (s_pawn, w_pawn, 55)):
        cs = CompiledScene(w)
        assert cs.info()["n_closed_media"] == 1
        cs_ref = CompiledScene(w, options=dict(interior_bound=0))
        assert cs_ref.info()["n_closed_media"] == 0
        a, st_a = cs.render(s, rt.mkStdGen(seed), want_stats=True)
        b, st_b = cs_ref.render(s, rt.mkStdGen(seed), want_stats=True)
        # identical event sequences: same ray count (a knife-edge tie may flip a handful of paths), same image
        assert abs(int(st_a["rays"]) - int(st_b["rays"])) <= 1e-4 * st_b["rays"]
        close = np.isclose(a, b, rtol=2e-4, atol=2e-5).all(axis=2)
        assert close.mean() >= 0.999, close.mean()
        assert abs(a.mean() - b.mean()) <= 1e-3 * b.mean()
    # the tail kernel (trace_segment) takes the same shortcut
    c = CompiledScene(w_pawn, options=dict(finish_max=4294967295)).render(s_pawn, rt.mkStdGen(55))
    assert np.isclose(c, a, rtol=2e-4, atol=2e-5).all(axis=2).mean() >= 0.999


def test_primary_ray_culling_with_a_lens(built):
    """Culling is on only when the whole lens lies outside the world box; a thin-lens camera outside the box (defocus
    blur, pool smaller than one sample pass so that compacted tiles cross iteration boundaries) renders the same
    image as with the camera rays sent through the queue, and statistics count the culled rays as path segments."""
    from raytrace_b200 import V3, constantTexture, lambertian, dielectric, group, sphere, bvhTree
    balls = [sphere((x * 0.8, y * 0.8, -6 - 0.2 * x), 0.35) for x in range(-4, 5) for y in range(-2, 3)]      # 45: a real BVH, not a flat list
    world = group([lambertian(constantTexture(V3(0.8, 0.3, 0.3))) << bvhTree(balls[:25]), dielectric(1.5) << bvhTree(balls[25:])])
    s = rt.defaultCameraSettings.replace(cs_imageWidth=90, cs_aspectRatio=1.5, cs_samplesPerPixel=12, cs_maxRecursionDepth=8,
                                         cs_background=scenes.sky, cs_defocusAngle=0.02, cs_focusDist=6.0, cs_vfov=1.2)
    cs = CompiledScene(world)
    assert cs.info()["n_bvh_nodes"] > 0
    cs.handle.set_pool_size(3072)
    a, sa = cs.render(s, rt.mkStdGen(11), want_stats=True)
    cs.handle.set_options(primary_cull=0)
    b, sb = cs.render(s, rt.mkStdGen(11), want_stats=True)
    cs.handle.set_options(primary_cull=None)
    assert sa["rays"] == sb["rays"] and sa["samples"] == sb["samples"]
    assert np.allclose(a, b, rtol=2e-4, atol=2e-5)
    assert sa["iterations"] <= sb["iterations"]         # culled paths never occupy pool slots
    ref, _ = OracleScene(world).render(s, 11)
    assert abs(luminance(a.astype(np.float64)).mean() - luminance(ref).mean()) < 0.03 * luminance(ref).mean()


def test_moving_primitives_in_a_list_of_their_own(built):
    """demo2's one motion-blurred sphere lives outside the tree (TF_MOVLIST: scanned once per segment) so that the 3406
    static primitives run the traversal kernel without the per-primitive transform branch (RtOptions.split_moving = 1); by
    default it sits in the tree and every primitive test carries the branch.  Same primitive ids and distances at three shutter
    times, same path segments, same image at equal seeds."""
    s, w = scenes.demo2(width=72, spp=8, depth=12)
    a, b = CompiledScene(w, options=dict(split_moving=1)), CompiledScene(w)
    assert a.info()["n_moving_list"] == 1 and b.info()["n_moving_list"] == 0
    probe = s.replace(cs_imageWidth=400)
    for t in (0.0, 0.5, 1.0):
        ia, ta = a.probe_primary(probe, time=t)
        ib, tb = b.probe_primary(probe, time=t)
        assert (ia == ib).mean() >= 0.9999
        hit = (ia == ib) & (ia != 0xFFFFFFFF)
        assert np.all(np.abs(ta[hit] - tb[hit]) <= 1e-5 * np.abs(tb[hit]))
    moving_id = 3401                                   # the lambertian moving sphere is hit somewhere in the frame
    img_a, st_a = a.render(s, rt.mkStdGen(1234), want_stats=True)
    img_b, st_b = b.render(s, rt.mkStdGen(1234), want_stats=True)
    assert abs(int(st_a["rays"]) - int(st_b["rays"])) <= 1e-4 * st_b["rays"] + 2
    assert np.isclose(img_a, img_b, rtol=2e-4, atol=2e-5).all(axis=2).mean() >= 0.999
    # the tail kernel and the instrumented traversal see the list too
    a.handle.set_options(finish_max=4294967295, count_work=1)
    img_c, st_c = a.render(s, rt.mkStdGen(1234), want_stats=True)
    assert np.isclose(img_c, img_a, rtol=2e-4, atol=2e-5).all(axis=2).mean() >= 0.999
    assert st_c["rays_counted"] > 0 and st_c["sphere_tests"] >= st_c["rays_counted"]     # every segment tests the moving sphere
    del moving_id


@pytest.mark.parametrize("name,kw", [("pawn", dict(width=64, spp=16)), ("bunny_volume", dict(width=64, spp=8))], ids=["pawn", "bunny"])
def test_boundary_queries_on_the_surface_tree(built, name, kw):
    """When a medium's boundary IS the surface geometry its boundary queries walk the surface tree (RtSceneInfo.n_shared_trees);
    with RtOptions.shared_boundary = 0 the boundary has a tree of its own and every boundary query is a traversal of it.  Same
    primitives, same records: the same events at equal seeds (the bunny has holes, so its second boundary queries are real
    traversals of the shared tree), up to a knife-edge tie."""
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    a, b = CompiledScene(w), CompiledScene(w, options=dict(shared_boundary=0))
    assert a.info()["n_shared_trees"] == 1 and b.info()["n_shared_trees"] == 0
    ia, sa = a.render(s, rt.mkStdGen(seed), want_stats=True)
    ib, sb = b.render(s, rt.mkStdGen(seed), want_stats=True)
    assert np.isfinite(ia).all()
    assert abs(int(sa["rays"]) - int(sb["rays"])) <= 2e-4 * sb["rays"] + 2
    assert np.isclose(ia, ib, rtol=2e-4, atol=2e-5).all(axis=2).mean() >= 0.998
    assert abs(ia.mean() - ib.mean()) <= 2e-3 * ib.mean()
    ida, ta = a.probe_primary(s.replace(cs_imageWidth=320))
    idb, tb = b.probe_primary(s.replace(cs_imageWidth=320))
    assert np.array_equal(ida, idb) and np.array_equal(ta, tb)
