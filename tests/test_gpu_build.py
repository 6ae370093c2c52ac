"""Device-side tree build (raytrace_b200/csrc/rt_build.cuh, RtOptions.gpu_build) - `-m gpu`.

A closest hit does not depend on how the set of primitives is searched, and exact ties go to the primitive the reference
evaluates first (ref_order) whatever the shape of the tree (Geometry.hs:340-345,356-361): so the tree the GPU builds must
give, pixel for pixel and BIT FOR BIT, the primitive ids and distances that the host-built tree gives - that is the test
(plus the structural invariants of the node array, and the oracle's all-pixel gate on the device-built tree).  The
BASELINE-size parity gates (tests/test_gpu_parity.py) run pawn, bunny and demo2 with the default options, i.e. on
device-built trees as well."""
import math

import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import scenes
from raytrace_b200.ray import CompiledScene
from oracle.harness import OracleScene, primary_gate
from util import check_device_tree, luminance
from test_gpu_random_scenes import _random_world

pytestmark = pytest.mark.gpu


def _leaves_in_evaluation_order(cs, nodes, canon, n_tree):
    """inside every leaf the primitives are stored in the reference's evaluation order (the scan takes the first on a tie)"""
    order = cs.handle.get_eval_order()
    pos = {int(c): r for r, c in enumerate(order[:n_tree])}          # canonical id -> evaluation position (surface part)
    ref = nodes[:, 3].view(np.uint32)
    for k in np.nonzero(ref & 0x80000000)[0]:
        first, c = int(ref[k] & 0x7FFFFFFF) >> 5, int(ref[k] & 31) + 1
        if first + c > n_tree:
            continue                                                  # a medium's own tree
        p = [pos[int(x)] for x in canon[first:first + c]]
        assert p == sorted(p), (k, p)


CASES = [("pawn", {}, {}), ("bunny_volume", {}, {}), ("demo2", {}, {}), ("demo1", {}, dict(gpu_build=1)),
         ("demo2", {}, dict(gpu_build=1, aggregate_boxes=0)), ("demo2", {}, dict(gpu_build=1, collapse=8, split_moving=1))]


@pytest.mark.parametrize("name,kw,opts", CASES, ids=["pawn", "bunny", "demo2", "demo1", "demo2-faces", "demo2-movlist"])
def test_device_built_tree_answers_like_the_host_built_one(built, name, kw, opts):
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    s = s.replace(cs_imageWidth=320, cs_samplesPerPixel=8)
    host_opts = dict(opts, gpu_build=0)
    dev, host = CompiledScene(w, upload=False, options=opts), CompiledScene(w, options=host_opts)
    assert dev.info()["gpu_build"] == 1 and host.info()["gpu_build"] == 0
    dev.handle.upload(0)
    di, hi = dev.info(), host.info()
    assert di["gpu_build"] == 2 and di["n_prims"] == hi["n_prims"] and di["max_depth"] < 32 and di["n_bvh_nodes"] > 0
    assert np.array_equal(dev.handle.get_eval_order(), host.handle.get_eval_order())
    nodes, canon = dev.handle.get_bvh()
    check_device_tree(nodes, di)
    n_tree = di["n_prims"] // 2 if di["n_shared_trees"] else di["n_prims"]
    if len(set(canon[:n_tree].tolist())) == n_tree:
        _leaves_in_evaluation_order(dev, nodes, canon, n_tree)
    # same cost model, same bins: the two builders end up with trees of the same size (fp32 vs fp64 centroids may move a few items)
    assert abs(di["n_bvh_nodes"] - hi["n_bvh_nodes"]) <= 0.05 * hi["n_bvh_nodes"] + 4, (di["n_bvh_nodes"], hi["n_bvh_nodes"])
    for time in (0.0, 0.37, 1.0):
        ids_d, t_d = dev.probe_primary(s, time=time)
        ids_h, t_h = host.probe_primary(s, time=time)
        assert np.array_equal(ids_d, ids_h)
        assert np.array_equal(t_d.view(np.uint32), t_h.view(np.uint32))          # bit for bit
    assert np.array_equal(dev.probe_surface(s).view(np.uint32), host.probe_surface(s).view(np.uint32))
    a, sa = dev.render(s, rt.mkStdGen(seed), want_stats=True)
    b, sb = host.render(s, rt.mkStdGen(seed), want_stats=True)
    assert sa["rays"] == sb["rays"] and sa["samples"] == sb["samples"]            # the same paths, segment for segment
    assert np.allclose(a, b, rtol=2e-4, atol=2e-5)                                # (fp32 summation order)


@pytest.mark.parametrize("seed", [11, 37, 96])
def test_device_build_on_random_scenes_against_the_oracle(built, seed):
    """nested groups / bvhTrees / transforms / moving instances / aggregated cuboids (items of six primitives that a leaf
    may not split), every tree built on the device: the oracle's all-pixel id + distance gate."""
    world = _random_world(seed)
    s = rt.defaultCameraSettings.replace(cs_imageWidth=240, cs_aspectRatio=1.5, cs_vfov=math.radians(70), cs_samplesPerPixel=16,
                                         cs_maxRecursionDepth=10, cs_background=scenes.sky)
    ora = OracleScene(world)
    for opts in (dict(gpu_build=1), dict(gpu_build=1, leaf_max=4, collapse=2), dict(gpu_build=1, aggregate_boxes=0, collapse=16)):
        cs = CompiledScene(world, options=opts)
        if cs.info()["n_prims"] <= opts.get("leaf_max", 24):
            continue
        assert cs.info()["gpu_build"] == 2
        nodes, canon = cs.handle.get_bvh()
        check_device_tree(nodes, cs.info())
        assert np.array_equal(cs.handle.get_eval_order()[: len(ora.leaf_order())], ora.leaf_order())
        for time in (0.0, 1.0):
            ids5, ts5 = ora.probe_tilted(s, time=time)
            ids, ts = cs.probe_primary(s, time=time)
            g = primary_gate(ids, ts, ids5, ts5)
            assert g["pass_frac"] >= 0.9995 and g["id_unexplained"] <= 3, (seed, opts, {k: v for k, v in g.items() if k != "passed_mask"})
    img = cs.render(s, rt.mkStdGen(seed))
    ref, _ = ora.render(s, seed)
    lg, lo = luminance(img.astype(np.float64)).mean(), luminance(ref).mean()
    assert np.isfinite(img).all() and abs(lg - lo) <= 0.05 * lo + 1e-3


@pytest.mark.parametrize("case", ["coincident", "peeling", "collinear", "two", "grid"])
def test_device_builder_pathological_inputs(built, case):
    """identical centroids (no bin split exists: the segment is halved as it stands), sizes growing geometrically along a line
    (SAH peels one primitive per level until the depth rule takes over), thousands of primitives on a line, the smallest
    tree, and a regular grid with thousands of exact centroid ties: the tree stays within the 32-entry traversal stack,
    reaches every primitive once, and answers exactly like the host-built tree."""
    mat = rt.lambertian(rt.constantTexture(0.5))
    if case == "coincident":
        prims = [mat << rt.sphere((0, 0, -5), 0.5 + 1e-3 * k) for k in range(300)]
    elif case == "peeling":
        prims = [mat << rt.sphere((1.5 ** k, 0, -5), 0.4 * 1.5 ** k) for k in range(60)]
    elif case == "collinear":
        prims = [mat << rt.sphere((0.01 * k - 25, 0.1, -30), 0.3) for k in range(5000)]
    elif case == "two":
        prims = [mat << rt.sphere((-1, 0, -5), 0.5), mat << rt.sphere((1, 0, -5), 0.5)]
    else:
        prims = [mat << rt.sphere((0.25 * i - 4, 0.25 * j - 4, -12 - 0.25 * k), 0.12) for i in range(32) for j in range(32) for k in range(4)]
    world = rt.bvhTree(prims)
    opts = dict(gpu_build=1, leaf_max=1) if case == "two" else dict(gpu_build=1)
    dev, host = CompiledScene(world, options=opts), CompiledScene(world, options=dict(opts, gpu_build=0))
    info = dev.info()
    assert info["gpu_build"] == 2 and info["n_prims"] == len(prims) and info["max_depth"] < 32
    nodes, canon = dev.handle.get_bvh()
    seen = check_device_tree(nodes, info)
    assert seen.all() and sorted(canon.tolist()) == list(range(len(prims)))
    s = rt.defaultCameraSettings.replace(cs_imageWidth=200, cs_aspectRatio=2.0, cs_vfov=math.radians(60), cs_samplesPerPixel=4)
    ids_d, t_d = dev.probe_primary(s)
    ids_h, t_h = host.probe_primary(s)
    assert np.array_equal(ids_d, ids_h) and np.array_equal(t_d.view(np.uint32), t_h.view(np.uint32))
    assert (ids_d != 0xFFFFFFFF).any()


def test_builds_back_to_back_share_one_scratch(built):
    """The builder's scratch (bins, index arrays) is one grow-only allocation per device, and every build must leave its bins
    clean for the next: large, small, large again, then two scenes built from two threads at once."""
    import threading
    s_p, pawn = scenes.pawn(width=64, spp=1)
    s_b, bunny = scenes.bunny_volume(width=64, spp=1)
    ref = {}
    for name, w, s in (("bunny", bunny, s_b), ("pawn", pawn, s_p)):
        ref[name] = CompiledScene(w, options=dict(gpu_build=0)).probe_primary(s)
    for name, w, s in (("bunny", bunny, s_b), ("pawn", pawn, s_p), ("bunny", bunny, s_b), ("pawn", pawn, s_p)):
        cs = CompiledScene(w)
        assert cs.info()["gpu_build"] == 2
        ids, t = cs.probe_primary(s)
        assert np.array_equal(ids, ref[name][0]) and np.array_equal(t.view(np.uint32), ref[name][1].view(np.uint32))
    out = {}

    def work(name, w, s):
        cs = CompiledScene(w, upload=False)
        out[name] = cs.probe_primary(s)                     # the probe's upload builds the tree

    th = [threading.Thread(target=work, args=a) for a in (("bunny", bunny, s_b), ("pawn", pawn, s_p))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for name in ("bunny", "pawn"):
        assert np.array_equal(out[name][0], ref[name][0]) and np.array_equal(out[name][1].view(np.uint32), ref[name][1].view(np.uint32))


def test_one_build_serves_every_device_of_a_context(built):
    """rt_render_multi uploads a fresh scene from one thread per device: the first builds the tree, the others wait and copy
    it (one tree, identical on every GPU).  With one GPU the context lists it twice."""
    import ctypes
    from raytrace_b200 import _ffi
    n = ctypes.c_int32(0)
    _ffi.lib().rt_device_count(ctypes.byref(n))
    devices = list(range(n.value)) if n.value >= 2 else [0, 0]
    s, w = scenes.pawn(width=96, spp=16)
    ctx = rt.RenderContext(devices)
    try:
        cs = CompiledScene(w, upload=False)
        img = ctx.render(cs, s, rt.mkStdGen(5))
        assert cs.info()["gpu_build"] == 2
    finally:
        ctx.close()
    one = CompiledScene(w, options=dict(gpu_build=0)).render(s, rt.mkStdGen(5))
    assert np.allclose(img, one, rtol=2e-4, atol=2e-5)
