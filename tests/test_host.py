"""CPU-only tests: the C-ABI library loads and exports what include/rt_b200.h declares, the host
logic behind it (description tree -> flattened scene, BVH policy, .obj loader, error behaviour)
and the Python mirror of the reference API.  No compute call is made (no GPU here)."""
import ctypes
import math
import os
import re

import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import _ffi, scenes
from raytrace_b200.ray import CompiledScene, classify_background, BG_CONSTANT, BG_YGRADIENT
from oracle.harness import OracleScene
from oracle import obj_oracle
from util import check_device_tree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built):
    L = _ffi.lib()
    header = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    declared = set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_ffi.EXPORTS), declared ^ set(_ffi.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.rt_version() == 0x00020001


def test_no_gpu_is_a_loud_error_not_a_fallback(built):
    n = ctypes.c_int32(-1)
    assert _ffi.lib().rt_device_count(ctypes.byref(n)) == 0
    if n.value > 0:
        pytest.skip("a GPU is visible")
    s, w = scenes.readme_example(width=16, spp=1)
    cs = CompiledScene(w, upload=False)
    with pytest.raises(_ffi.RtError) as e:
        cs.render(s, rt.mkStdGen(1))
    assert e.value.status == 7 and "no CPU fallback" in e.value.message


def test_page_locked_buffers_need_the_driver(built):
    """rt_host_alloc without a CUDA device is RT_ERR_NO_DEVICE (never a silent malloc); rt_host_free refuses pointers
    it did not hand out."""
    L = _ffi.lib()
    n = ctypes.c_int32(-1)
    L.rt_device_count(ctypes.byref(n))
    p = ctypes.c_void_p()
    if n.value == 0:
        assert L.rt_host_alloc(4096, ctypes.byref(p)) == 7 and not p.value
        assert b"no CPU fallback" in L.rt_last_error()
    assert L.rt_host_free(ctypes.c_void_p(4096)) == 1
    assert L.rt_host_free(None) == 0
    assert L.rt_host_alloc(16, None) == 1


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "raytrace_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("no oracle", ""), f


@pytest.mark.parametrize("policy", ["sah", "reference"])
@pytest.mark.parametrize("name", ["readme", "cornell", "demo1", "demo2", "pawn", "bunny_volume"])
def test_flattened_tree_matches_reference_evaluation_order(built, name, policy):
    """bvhTree policy (Geometry.hs:369-377) + group order: the evaluation order the flattener records (it decides exact
    ties) must equal the oracle's evaluation-order leaf sequence (same split axis, same stable sort, same n div 2
    split) — whether the device tree is the default SAH tree (its own storage order) or the reference-shaped one
    (storage order = evaluation order)."""
    ctor, _ = scenes.CONFIGS[name]
    s, w = ctor()
    cs = CompiledScene(w, upload=False, options=dict(bvh_reference=int(policy == "reference"), gpu_build=0))     # host builders
    nodes, canon = cs.handle.get_bvh()
    ora = OracleScene(w)
    order = ora.leaf_order()
    assert np.array_equal(cs.handle.get_eval_order()[: len(order)], order)
    assert sorted(canon.tolist()) == sorted(cs.handle.get_eval_order().tolist())
    if policy == "reference":
        assert np.array_equal(canon[: len(order)], order)
    check_device_tree(nodes, cs.info())


def test_large_surface_tree_is_left_to_the_device(built):
    """RtOptions.gpu_build (default: trees of >= 1024 items): commit stops after the reference-policy pass - the evaluation
    order is there, the tree is not - and the first upload builds it on the GPU (tests/test_gpu_build.py).  Small trees,
    the reference-shaped tree and gpu_build = 0 stay on the host."""
    s, w = scenes.bunny_volume(width=8, spp=1)
    cs = CompiledScene(w, upload=False)
    info = cs.info()
    assert info["gpu_build"] == 1 and info["n_bvh_nodes"] == 0 and info["surface_root"] == 0 and info["n_shared_trees"] == 1
    with pytest.raises(_ffi.RtError) as e:
        cs.handle.get_bvh()
    assert e.value.status == 8 and "built on the device" in e.value.message
    order = OracleScene(w).leaf_order()
    assert np.array_equal(cs.handle.get_eval_order()[: len(order)], order)
    host = CompiledScene(w, upload=False, options=dict(gpu_build=0))
    assert host.info()["gpu_build"] == 0 and host.info()["n_bvh_nodes"] > 0
    assert np.array_equal(host.handle.get_eval_order(), cs.handle.get_eval_order())
    assert CompiledScene(w, upload=False, options=dict(bvh_reference=1)).info()["gpu_build"] == 0
    s, w = scenes.demo1(width=8, spp=1)                                    # 484 items: below the default threshold
    assert CompiledScene(w, upload=False).info()["gpu_build"] == 0
    assert CompiledScene(w, upload=False, options=dict(gpu_build=1)).info()["gpu_build"] == 1
    s, w = scenes.cornell_box()                                            # one scanned leaf: there is no tree to build
    assert CompiledScene(w, upload=False, options=dict(gpu_build=1)).info()["gpu_build"] == 0
    s, w = scenes.demo2(width=8, spp=1)                                    # the device writes its pairs behind whatever the host built
    info = CompiledScene(w, upload=False).info()
    assert info["gpu_build"] == 1 and info["surface_root"] == info["n_bvh_nodes"]


def test_boundary_that_is_the_surface_shares_its_tree(built):
    """`dielectric <$ mesh` + `constantMedium rho mesh` (pawn, bunny): the medium's boundary is exactly the surface geometry,
    so its root IS the surface root and no second tree is built; with the shortcut off (or when the surfaces hold more than
    the boundary, demo2) the boundary keeps a tree of its own.  Evaluation order and primitive bookkeeping are unchanged."""
    for name in ("pawn", "bunny_volume"):
        ctor, _ = scenes.CONFIGS[name]
        s, w = ctor()
        a = CompiledScene(w, upload=False, options=dict(gpu_build=0))
        b = CompiledScene(w, upload=False, options=dict(shared_boundary=0, gpu_build=0))
        ia, ib = a.info(), b.info()
        assert ia["n_shared_trees"] == 1 and ib["n_shared_trees"] == 0
        assert ia["n_prims"] == ib["n_prims"] and ia["n_bvh_nodes"] * 2 == ib["n_bvh_nodes"]      # one tree instead of two equal ones
        assert np.array_equal(a.handle.get_eval_order(), b.handle.get_eval_order())
        assert sorted(a.handle.get_bvh()[1].tolist()) == sorted(b.handle.get_bvh()[1].tolist())
        r = CompiledScene(w, upload=False, options=dict(bvh_reference=1)).info()
        assert r["n_shared_trees"] == 1
    s, w = scenes.demo2(width=16, spp=1)
    assert CompiledScene(w, upload=False).info()["n_shared_trees"] == 0


def test_scene_counts(built):
    s, w = scenes.demo2(width=16, spp=1)
    info = CompiledScene(w, upload=False).info()
    assert info["n_media"] == 2 and info["n_planars"] == 2401 and info["n_spheres"] == 1008
    assert info["n_moving_list"] == 0        # default: the motion-blurred sphere stays in the tree
    assert CompiledScene(w, upload=False, options=dict(split_moving=1)).info()["n_moving_list"] == 1     # TF_MOVLIST, opt-in
    assert CompiledScene(w, upload=False, options=dict(split_moving=1, bvh_reference=1)).info()["n_moving_list"] == 0
    s, w = scenes.cornell_box()
    info = CompiledScene(w, upload=False).info()
    assert info["n_prims"] == 18 and info["n_media"] == 0


def test_user_defined_closures_are_rejected(built):
    ball = rt.sphere((0, 0, -1), 0.5)
    cases = [
        rt.Material(lambda d, h: None) << ball,
        rt.lambertian(rt.solidTexture(lambda p: p)) << ball,
        rt.lambertian(rt.uvTexture(lambda uv: (1, 1, 1))) << ball,
        rt.lambertian(rt.Texture(lambda p, uv: (1, 1, 1))) << ball,
        rt.lambertian(rt.constantTexture(1)) << rt.planeShape((0, 0, 0), (1, 0, 0), (0, 1, 0), lambda a, b: a * a + b * b < 1,
                                                             lambda a, b: (a, b), rt.fromCorners((0, 0, 0), (1, 1, 0))),
    ]
    for w in cases:
        with pytest.raises(_ffi.UserDefinedError) as e:
            CompiledScene(w, upload=False)
        assert "no CPU fallback" in e.value.message
    with pytest.raises(_ffi.UserDefinedError):
        classify_background(lambda ray: (ray.direction[0] ** 2, 0, 0))


def test_error_codes(built):
    white = rt.lambertian(rt.constantTexture(1))
    with pytest.raises(ValueError, match="bvhTree: empty list"):
        rt.bvhTree([])
    with pytest.raises(_ffi.RtError) as e:      # scale must not be used with transform (Geometry.hs:434)
        CompiledScene(white << rt.transform(rt.scale(2.0), rt.sphere((0, 0, 0), 1)), upload=False)
    assert e.value.status == 4
    with pytest.raises(TypeError):               # Geometry m () passed to raytrace
        CompiledScene(rt.sphere((0, 0, 0), 1), upload=False)
    with pytest.raises(TypeError):               # group [Geometry Identity a, Geometry (State StdGen) a]
        rt.group([white << rt.sphere((0, 0, 0), 1), white << rt.constantMedium(1.0, rt.sphere((0, 0, 0), 1))])
    # raw C ABI: empty bvhTree -> RT_ERR_EMPTY_BVH ; unknown ids -> RT_ERR_INVALID_ARG
    L = _ffi.lib()
    h = ctypes.c_void_p()
    assert L.rt_scene_create(ctypes.byref(h)) == 0
    d = _ffi.RtNodeDesc()
    d.kind = 8
    out = ctypes.c_uint32()
    assert L.rt_scene_add_node(h, ctypes.byref(d), ctypes.byref(out)) == 0
    assert L.rt_scene_commit(h, out.value) == 2
    assert b"bvhTree: empty list" in L.rt_last_error()
    assert L.rt_scene_commit(h, 99) == 1
    info = _ffi.RtSceneInfo()
    assert L.rt_scene_info(h, ctypes.byref(info)) == 8
    L.rt_scene_destroy(h)


def test_batched_description_equals_the_object_walk(built):
    """rt_scene_add_batch: the records of a world's first serialisation, kept on the (immutable) world, give every later scene
    the same description in one call - same flattened scene, same device tree - and a bad record stops the batch with the
    error the single call reports."""
    for name in ("demo1", "demo2", "cornell"):
        ctor, _ = scenes.CONFIGS[name]
        s, w = ctor()
        w.__dict__.pop("_rt_flat", None)
        a = CompiledScene(w, upload=False, options=dict(gpu_build=0))        # walks the objects, leaves the arrays on the world
        assert "_rt_flat" in w.__dict__
        b = CompiledScene(w, upload=False, options=dict(gpu_build=0))        # one rt_scene_add_batch call
        ia, ib = a.info(), b.info()
        assert ia == ib and ia["n_prims"] > 0
        (na, ca), (nb, cb) = a.handle.get_bvh(), b.handle.get_bvh()
        assert np.array_equal(na, nb) and np.array_equal(ca, cb)
    L = _ffi.lib()
    h = ctypes.c_void_p()
    assert L.rt_scene_create(ctypes.byref(h)) == 0
    nodes = (_ffi.RtNodeDesc * 2)()
    nodes[0].kind, nodes[1].kind = 1, 77                                       # a sphere, then an unknown kind
    nodes[0].s[0] = 1.0
    first = (ctypes.c_uint32 * 3)()
    assert L.rt_scene_add_batch(h, None, 0, None, 0, nodes, 2, first) == 1 and b"unknown node kind" in L.rt_last_error()
    assert tuple(first) == (0, 0, 0)
    assert L.rt_scene_add_batch(h, None, 0, None, 0, nodes, 1, first) == 0 and first[2] == 1   # the sphere of the failed batch stayed
    assert L.rt_scene_add_batch(h, None, 3, None, 0, None, 0, first) == 1                    # a count without an array
    L.rt_scene_destroy(h)


def test_outermost_material_wins_and_shared_geometry(built):
    a, b = rt.lambertian(rt.constantTexture(0.1)), rt.mirror(rt.constantTexture(0.9))
    inner = a << rt.sphere((0, 0, 0), 1)
    cs = CompiledScene(b << rt.group([inner, a << rt.sphere((3, 0, 0), 1)]), upload=False)
    assert cs.info()["n_prims"] == 2
    s, w = scenes.pawn(spp=1, width=8)           # the same mesh object is surface and medium boundary
    nodes, canon = CompiledScene(w, upload=False, options=dict(gpu_build=0)).handle.get_bvh()
    # surface records in the tree's leaf order, then the boundary's copies of the SAME canonical primitives (they share the
    # surface tree, so they are stored in evaluation order behind it)
    assert np.array_equal(np.sort(canon[:1200]), np.arange(1200)) and np.array_equal(np.sort(canon[1200:]), np.arange(1200))


def test_image_height_round_half_even(built):
    L = _ffi.lib()
    for w, aspect, expect in [(600, 16 / 9, 338), (1200, 16 / 9, 675), (100, 1.0, 100), (5, 2.0, 2), (7, 2.0, 4), (1920, 16 / 9, 1080)]:
        s = rt.defaultCameraSettings.replace(cs_imageWidth=w, cs_aspectRatio=aspect)
        cam = _ffi.make_camera(s, 1, (1, 1, 1), (1, 1, 1))
        h = ctypes.c_int32()
        assert L.rt_image_height(ctypes.byref(cam), ctypes.byref(h)) == 0
        assert h.value == expect == rt.imageHeight(s)


def test_background_classification():
    assert classify_background(rt.defaultCameraSettings.cs_background)[0] == BG_CONSTANT
    k, c0, c1 = classify_background(scenes.sky)
    assert k == BG_YGRADIENT and np.allclose(c0, (1, 1, 1)) and np.allclose(c1, (0.5, 0.7, 1.0))
    k, c0, c1 = classify_background(scenes.grayFade)
    assert k == BG_YGRADIENT and np.allclose(c0, 0) and np.allclose(c1, 1)


# ---- .obj loader (parseObj, Geometry.hs:207-285) -------------------------------------------------------
OBJ_OK = """# comment
v 0 0 0
v 1 0 0   # trailing comment
v 1 1 0 9 9
v 0 1 0
v -0.5e1 2.5E-1 3
vt 0.5
vt 0.25 0.75 0
vn 0 0 1
f 1 2 3 4
f -1 -2 -3
f 1/1 2/2 3//7
f 1//1 2/-1 5
g ignored
"""
OBJ_BAD = [
    ("v 1 2\n", "line 1: invalid 'v' statement"),
    ("v 1 2 .5\n", "line 1: invalid 'v' statement"),
    ("v 0 0 0\nvt x\n", "line 2: invalid 'vt' statement"),
    ("v 0 0 0\nv 1 0 0\nf 1 2\n", "line 3: invalid 'f' statement (fewer than 3 vertices)"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 4\n", "line 4: index out of bounds: 4"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 0\n", "line 4: index out of bounds: 0"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 3x\n", "line 4: unexpected character 'x'"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 a\n", "line 4: expected number"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1/ 2 3\n", "line 4: expected number"),
    ("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1/1 2 3\n", "line 4: index out of bounds: 1"),
]


def test_obj_loader_matches_restatement(built):
    m = rt.parseObj(OBJ_OK)
    v, t, f = obj_oracle.parse_obj(OBJ_OK)
    assert np.array_equal(m.verts, v) and np.array_equal(m.uvs, t) and np.array_equal(m.faces, f)
    assert len(f) == 2 + 1 + 1 + 1 and tuple(f[0]) == (0, -1, 1, -1, 2, -1) and tuple(f[1]) == (0, -1, 2, -1, 3, -1)
    assert tuple(f[2]) == (4, -1, 3, -1, 2, -1) and tuple(f[3]) == (0, 0, 1, 1, 2, -1) and tuple(f[4]) == (0, -1, 1, 1, 4, -1)
    assert np.array_equal(t, [[0.5, 0.0], [0.25, 0.75]])
    for text, msg in OBJ_BAD:
        with pytest.raises(ValueError) as e1:
            obj_oracle.parse_obj(text)
        with pytest.raises(ValueError) as e2:
            rt.parseObj(text)
        assert str(e1.value) == msg == str(e2.value), (text, str(e1.value), str(e2.value))


def test_obj_fixture_meshes():
    """The product parses the REAL .obj text of the reference's images/pawn.obj and images/bunny.obj (package data) with
    rt_obj_parse; the result must equal, bit for bit, what the pure-Python restatement of parseObj (oracle/obj_oracle.py,
    Geometry.hs:207-285) made of the same files when the fixtures were generated (tests/golden/assets/*.npz), and what it
    makes of the package's copy now."""
    golden = os.path.join(ROOT, "tests", "golden", "assets")
    for name, nv, nf in (("pawn", 602, 1200), ("bunny", 2503, 4968)):
        mesh = scenes.load_mesh(name)                                        # readObj -> rt_obj_parse on the text
        assert mesh.verts.shape == (nv, 3) and mesh.faces.shape == (nf, 6)
        z = np.load(os.path.join(golden, f"{name}.npz"))
        assert np.array_equal(mesh.verts, z["verts"]) and np.array_equal(mesh.uvs.reshape(-1, 2), z["uvs"].reshape(-1, 2))
        assert np.array_equal(mesh.faces, z["faces"])
        with open(os.path.join(scenes.ASSET_DIR, f"{name}.obj")) as f:
            v, t, fc = obj_oracle.parse_obj(f.read())
        assert np.array_equal(mesh.verts, v) and np.array_equal(mesh.faces, fc)
    pawn = scenes.load_mesh("pawn")
    assert (pawn.faces[:, 1::2] == -1).all()
    # round trip through the library's loader
    lines = ["v %r %r %r" % tuple(float(x) for x in v) for v in pawn.verts] + ["f %d %d %d" % tuple(f[0::2] + 1) for f in pawn.faces]
    m = rt.parseObj("\n".join(lines) + "\n")
    assert np.array_equal(m.verts, pawn.verts) and np.array_equal(m.faces, pawn.faces)
    # the JPEG asset decodes to the fixture made when the goldens were generated (decoders may differ by a code value or two)
    z = np.load(os.path.join(golden, "earthmap.npz"))["rgb8"].astype(np.int32)
    from PIL import Image
    now = np.asarray(Image.open(os.path.join(scenes.ASSET_DIR, "earthmap.jpg")).convert("RGB"), dtype=np.int32)
    assert now.shape == z.shape == (512, 1024, 3) and np.abs(now - z).max() <= 2


def test_bounding_boxes_follow_the_reference():
    assert rt.boundingBox(rt.sphere((1, 2, 3), 2)) == ((-1, 3), (0, 4), (1, 5))
    b = rt.boundingBox(rt.parallelogram((0, 0, 0), (1, 0, 0), (0, 1, 0)))
    assert b == ((-0.0001, 1.0001), (-0.0001, 1.0001), (-0.0001, 0.0001))                      # padBox, Geometry.hs:144
    g = rt.moving((0, 0, 0), (30, 0, 0), rt.sphere((400, 400, 200), 50))
    assert rt.boundingBox(g) == ((350, 480), (350, 450), (150, 250))
    t = rt.transform(rt.translate((1, 0, 0)) @ rt.rotateY(math.pi / 2), rt.sphere((0, 0, 0), 1))
    assert np.allclose(np.array(rt.boundingBox(t)), [[0, 2], [-1, 1], [-1, 1]])
    assert rt.longestDim(((0, 1), (0, 1), (0, 1))) == 2 and rt.longestDim(((0, 2), (0, 2), (0, 1))) == 1  # argMax ties, Core.hs:43-46


@pytest.mark.parametrize("case", ["coincident", "peeling", "single", "collinear"])
def test_sah_builder_pathological_inputs(built, case):
    """The SAH device tree must stay within the 32-entry traversal stack and reach every primitive whatever the input:
    identical centroids (no SAH split exists), sizes growing geometrically along a line (SAH peels one primitive per
    level), a single primitive, thousands of primitives on a line."""
    mat = rt.lambertian(rt.constantTexture(0.5))
    if case == "coincident":
        prims = [mat << rt.sphere((0, 0, -5), 0.5 + 1e-3 * k) for k in range(300)]
    elif case == "peeling":
        prims = [mat << rt.sphere((1.5 ** k, 0, -5), 0.4 * 1.5 ** k) for k in range(60)]
    elif case == "single":
        prims = [mat << rt.sphere((0, 0, -5), 1)]
    else:
        prims = [mat << rt.sphere((0.01 * k, 0, -5), 0.004) for k in range(5000)]
    cs = CompiledScene(rt.bvhTree(prims), upload=False, options=dict(gpu_build=0))
    info = cs.info()
    nodes, canon = cs.handle.get_bvh()
    assert info["n_prims"] == len(prims) and info["max_depth"] < 32
    assert sorted(cs.handle.get_eval_order().tolist()) == sorted(canon.tolist())
    if len(prims) <= 24:
        assert info["n_bvh_nodes"] == 0          # short list: one leaf
        return
    ref = nodes[:, 3].view(np.uint32)
    seen = np.zeros(len(prims), bool)
    for k in range(len(nodes)):
        if ref[k] & 0x80000000:
            first, c = int(ref[k] & 0x7FFFFFFF) >> 5, int(ref[k] & 31) + 1
            assert not seen[first:first + c].any()
            seen[first:first + c] = True
    assert seen.all()


def _volume(mesh_geom, density=5.0):
    from raytrace_b200 import V3, constantTexture, dielectric, isotropic, group, pureGeometry, constantMedium
    return group([pureGeometry(dielectric(1.5) << mesh_geom),
                  isotropic(constantTexture(V3(1, 0, 0))) << constantMedium(density, mesh_geom)])


def test_closed_boundary_detection(built):
    """Interior queries are bounded by the free-flight distance only when the medium's boundary is part of the surfaces
    AND provably closed and outward oriented (DMedium::closed): every directed edge once, its reverse once, positive
    volume in world space - or one sphere.  Anything else keeps the reference's unbounded boundary queries."""
    from raytrace_b200 import Mesh, triangleMesh, sphere, transform, scale, rotateY
    closed = lambda g: CompiledScene(_volume(g), upload=False).info()["n_closed_media"]   # noqa: E731
    assert CompiledScene(scenes.pawn(width=8, spp=1)[1], upload=False).info()["n_closed_media"] == 1      # watertight
    assert CompiledScene(scenes.bunny_volume(width=8, spp=1)[1], upload=False).info()["n_closed_media"] == 0  # holes
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], dtype=np.float64)
    tet = [(0, 2, 1), (0, 1, 3), (0, 3, 2), (1, 2, 3)]                  # outward-oriented tetrahedron
    faces = lambda tris: np.array([[a, -1, b, -1, c, -1] for a, b, c in tris], dtype=np.int32)   # noqa: E731
    assert closed(triangleMesh(Mesh(v, np.zeros((0, 2)), faces(tet)))) == 1
    assert closed(triangleMesh(Mesh(v, np.zeros((0, 2)), faces(tet[:3])))) == 0                       # one face missing
    assert closed(triangleMesh(Mesh(v, np.zeros((0, 2)), faces([(a, c, b) for a, b, c in tet])))) == 0  # inside out
    assert closed(triangleMesh(Mesh(v, np.zeros((0, 2)), faces(tet[:3] + [(1, 3, 2)])))) == 0          # one face flipped
    dup = np.vstack([v, v[1:2]])                                        # an obj-style duplicated vertex still matches
    assert closed(triangleMesh(Mesh(dup, np.zeros((0, 2)), faces([(0, 2, 4), (0, 1, 3), (0, 3, 2), (1, 2, 3)])))) == 1
    assert closed(transform(rotateY(0.3), triangleMesh(Mesh(v, np.zeros((0, 2)), faces(tet))))) == 1  # rotated instance
    assert closed(sphere((0, 0, 0), 1.0)) == 1
    assert closed(sphere((0, 0, 0), -1.0)) == 0                         # the hollow-sphere trick: normals point inwards
    assert CompiledScene(scenes.demo2(width=8, spp=1)[1], upload=False).info()["n_media"] == 2        # two media: hint unused


def test_options_struct_round_trip(built):
    """RtOptions (include/rt_b200.h) is the library's only tuning interface - it reads no environment variables:
    every field defaults to -1, set_options updates by name, unknown names are refused, and commit-time fields reach
    the flattener (reference-shaped tree: storage order = evaluation order)."""
    s, w = scenes.demo1(width=16, spp=1)
    cs = CompiledScene(w, upload=False)
    o = cs.handle.get_options()
    assert set(o) == set(_ffi.RtOptions.NAMES) and all(v < 0 for v in o.values())
    cs.handle.set_options(node_burst=2, pool=4096, sah_node_cost=0.5)
    o = cs.handle.get_options()
    assert o["node_burst"] == 2 and o["pool"] == 4096 and o["sah_node_cost"] == 0.5 and o["refill"] == -1
    cs.handle.set_options(node_burst=None)
    assert cs.handle.get_options()["node_burst"] == -1
    cs.handle.set_pool_size(0)
    assert cs.handle.get_options()["pool"] == -1
    with pytest.raises(KeyError):
        cs.handle.set_options(no_such_knob=1)
    ref = CompiledScene(w, upload=False, options=dict(bvh_reference=1))
    assert np.array_equal(ref.handle.get_bvh()[1], ref.handle.get_eval_order())
    assert not np.array_equal(cs.handle.get_bvh()[1], cs.handle.get_eval_order())
    for src in ("rt_api.cpp", "rt_flatten.cpp", "rt_kernels.cu", "rt_multi.cu"):
        path = os.path.join(ROOT, "raytrace_b200", "csrc", src)
        if os.path.exists(path):
            assert "getenv" not in open(path).read(), src


def test_primary_gate_counts_every_pixel():
    """oracle.harness.primary_gate: raw matches, tie-explained pixels (id equals a TILTED oracle answer and t agrees with
    the oracle's t for that id) and unexplained ones, over all pixels."""
    from oracle.harness import primary_gate
    MISS = 0xFFFFFFFF
    ids5 = np.zeros((5, 1, 6), np.uint32)
    ts5 = np.full((5, 1, 6), 2.0)
    ids5[:, 0, 1] = [3, 3, 4, 3, 3]          # pixel 1: an edge - the tilts see primitive 4 as well
    ts5[:, 0, 1] = [2.0, 2.0, 2.5, 2.0, 2.0]
    ids5[:, 0, 5] = MISS
    ts5[:, 0, 5] = np.inf
    gpu_ids = np.array([[0, 4, 7, 0, 4, MISS]], np.uint32)
    gpu_ts = np.array([[2.0, 2.5, 2.0, 2.0001, 2.0, np.inf]], np.float32)
    g = primary_gate(gpu_ids, gpu_ts, ids5, ts5)
    assert g["pixels"] == 6 and g["raw_match"] == 3           # pixels 0, 3, 5
    assert g["tie_explained"] == 1                            # pixel 1
    assert g["id_unexplained"] == 2                           # pixels 2 and 4: ids the oracle never returns there
    assert g["t_failed"] == 1                                 # pixel 3: right id, t off by 5e-5
    assert g["unexplained"] == 3 and abs(g["pass_frac"] - 0.5) < 1e-12
