"""ctypes binding of include/rt_b200.h (librt_b200.so, built in-tree by __graft_entry__.build()).

This is the Python spelling of the reference-side binding a maintainer would add
(INTEGRATION.md shows the Haskell `foreign import ccall` form).  The library is the product:
if it is missing, importing the render path fails loudly — there is no fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB: an alternative build of the SAME library (A/B runs of compile-time knobs, tools/gpu_sweep.py)
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")

RT_OK = 0
STATUS_NAMES = {
    0: "RT_OK", 1: "RT_ERR_INVALID_ARG", 2: "RT_ERR_EMPTY_BVH", 3: "RT_ERR_USER_DEFINED",
    4: "RT_ERR_NON_EUCLIDEAN", 5: "RT_ERR_TYPE", 6: "RT_ERR_CUDA", 7: "RT_ERR_NO_DEVICE",
    8: "RT_ERR_NOT_COMMITTED", 9: "RT_ERR_PARSE", 10: "RT_ERR_UNSUPPORTED",
}
PRIM_MISS = 0xFFFFFFFF
PRIM_MEDIUM = 0xFFFFFFFE


class RtError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {message}")
        self.status = status
        self.message = message


class UserDefinedError(RtError):
    """Scene uses a user-defined closure (material / texture / shape / background)."""


D3 = C.c_double * 3


class RtTextureDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("i0", C.c_int32), ("i1", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32),
                ("freq", C.c_double), ("c0", D3), ("c1", D3), ("shift", D3), ("dir", D3),
                ("image", C.POINTER(C.c_float))]


class RtMaterialDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("texture", C.c_uint32), ("param", C.c_double)]


class RtNodeDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("material", C.c_uint32), ("n_children", C.c_uint32),
                ("n_verts", C.c_uint32), ("n_uvs", C.c_uint32), ("n_faces", C.c_uint32),
                ("children", C.POINTER(C.c_uint32)), ("verts", C.POINTER(C.c_double)),
                ("uvs", C.POINTER(C.c_double)), ("faces", C.POINTER(C.c_int32)),
                ("v", (C.c_double * 3) * 3), ("s", C.c_double * 2),
                ("uv", (C.c_double * 2) * 3), ("m", C.c_double * 16)]


class RtRedirectTarget(C.Structure):
    _fields_ = [("p", C.c_double), ("q", D3), ("s0", D3), ("s1", D3)]


class RtCamera(C.Structure):
    _fields_ = [("center", D3), ("look_at", D3), ("up", D3),
                ("vfov", C.c_double), ("aspect_ratio", C.c_double),
                ("defocus_angle", C.c_double), ("focus_dist", C.c_double),
                ("bg_c0", D3), ("bg_c1", D3),
                ("bg_kind", C.c_int32), ("image_width", C.c_int32),
                ("samples_per_pixel", C.c_int32), ("max_recursion_depth", C.c_int32),
                ("n_redirect_targets", C.c_int32), ("_pad", C.c_int32),
                ("redirect_targets", C.POINTER(RtRedirectTarget))]


class RtStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("launches", C.c_uint64),
                ("iterations", C.c_uint32), ("pool", C.c_uint32),
                ("ms_total", C.c_double), ("ms_traverse", C.c_double), ("ms_shade", C.c_double),
                ("ms_raygen", C.c_double), ("ms_copy", C.c_double), ("ms_finish", C.c_double),
                ("rays_finish", C.c_uint64), ("paths_finish", C.c_uint64),
                ("node_pairs", C.c_uint64), ("sphere_tests", C.c_uint64), ("planar_tests", C.c_uint64),
                ("box_tests", C.c_uint64), ("rays_counted", C.c_uint64)]

    def as_dict(self) -> Dict[str, float]:
        return {n: getattr(self, n) for n, _ in self._fields_}


class RtSceneInfo(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("n_spheres", C.c_uint32), ("n_planars", C.c_uint32),
                ("n_bvh_nodes", C.c_uint32), ("n_media", C.c_uint32), ("n_materials", C.c_uint32),
                ("n_textures", C.c_uint32), ("n_xforms", C.c_uint32), ("surface_root", C.c_uint32),
                ("max_depth", C.c_uint32), ("device_bytes", C.c_uint64),
                ("n_closed_media", C.c_uint32), ("n_moving_list", C.c_uint32),
                ("n_shared_trees", C.c_uint32), ("gpu_build", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class RtContextInfo(C.Structure):
    _fields_ = [("n_devices", C.c_int32), ("reduce_mode", C.c_int32), ("devices", C.c_int32 * 16), ("reduce_note", C.c_char * 160)]

    MODES = {0: "single", 1: "multicast", 2: "peer", 3: "gather"}

    def as_dict(self):
        return {"n_devices": self.n_devices, "reduce_mode": self.MODES.get(self.reduce_mode, str(self.reduce_mode)),
                "devices": list(self.devices[: self.n_devices]), "reduce_note": self.reduce_note.decode("utf-8", "replace")}


class RtOptions(C.Structure):
    """include/rt_b200.h RtOptions: negative = library default."""
    _fields_ = [("bvh_reference", C.c_int32), ("leaf_max", C.c_int32), ("collapse", C.c_int32),
                ("aggregate_boxes", C.c_int32), ("shared_boundary", C.c_int32), ("interior_bound", C.c_int32),
                ("world_box", C.c_int32), ("split_moving", C.c_int32), ("trace_commit", C.c_int32), ("gpu_build", C.c_int32),
                ("sah_node_cost", C.c_double),
                ("pool", C.c_int64), ("finish_max", C.c_int64),
                ("finish_refill", C.c_int32), ("tail_batch", C.c_int32), ("node_burst", C.c_int32),
                ("refill", C.c_int32), ("trav_blocks", C.c_int32), ("tile", C.c_int32),
                ("primary_cull", C.c_int32), ("shade_sort", C.c_int32), ("timing", C.c_int32),
                ("count_work", C.c_int32), ("debug_sync", C.c_int32), ("raygen_block", C.c_int32), ("material_sampling", C.c_int32), ("smem_carveout", C.c_int32), ("reserved", C.c_int32 * 5)]

    NAMES = ("bvh_reference", "leaf_max", "collapse", "aggregate_boxes", "shared_boundary", "interior_bound", "world_box",
             "split_moving", "trace_commit", "gpu_build", "sah_node_cost", "pool", "finish_max", "finish_refill", "tail_batch",
             "node_burst", "refill", "trav_blocks", "tile", "primary_cull", "shade_sort", "timing", "count_work", "debug_sync", "raygen_block", "material_sampling", "smem_carveout")


# every symbol include/rt_b200.h declares (tests check that the .so exports all of them)
EXPORTS = [
    "rt_version", "rt_last_error", "rt_device_count",
    "rt_scene_create", "rt_scene_destroy", "rt_scene_add_texture", "rt_scene_add_material",
    "rt_scene_add_node", "rt_scene_add_batch", "rt_scene_set_noise_tables", "rt_scene_commit", "rt_scene_info", "rt_scene_get_eval_order",
    "rt_scene_get_bvh", "rt_scene_upload", "rt_image_height", "rt_render", "rt_render_device",
    "rt_probe_primary", "rt_probe_surface", "rt_set_pool_size", "rt_tonemap_u8",
    "rt_obj_parse", "rt_mesh_sizes", "rt_mesh_data", "rt_mesh_destroy", "rt_host_alloc", "rt_host_free", "rt_reduce_image",
    "rt_options_init", "rt_scene_set_options", "rt_scene_get_options",
    "rt_context_create", "rt_context_destroy", "rt_context_info", "rt_render_multi",
    "rt_tonemap_u8_static",
]

_lib: Optional[C.CDLL] = None


def lib() -> C.CDLL:
    """Load librt_b200.so; raise loudly if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA backend has not been built "
            "(run `python -c 'import __graft_entry__ as g; g.build()'`). "
            "raytrace_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, u32p = C.c_void_p, C.POINTER(C.c_uint32)
    L.rt_version.restype = C.c_uint32
    L.rt_last_error.restype = C.c_char_p
    L.rt_device_count.argtypes = [C.POINTER(C.c_int32)]
    L.rt_scene_create.argtypes = [C.POINTER(vp)]
    L.rt_scene_destroy.argtypes = [vp]
    L.rt_scene_add_texture.argtypes = [vp, C.POINTER(RtTextureDesc), u32p]
    L.rt_scene_add_material.argtypes = [vp, C.POINTER(RtMaterialDesc), u32p]
    L.rt_scene_add_node.argtypes = [vp, C.POINTER(RtNodeDesc), u32p]
    L.rt_scene_add_batch.argtypes = [vp, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, u32p]
    L.rt_scene_set_noise_tables.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                            C.POINTER(C.c_int32), C.POINTER(C.c_double)]
    L.rt_scene_commit.argtypes = [vp, C.c_uint32]
    L.rt_scene_get_eval_order.argtypes = [vp, u32p]
    L.rt_scene_info.argtypes = [vp, C.POINTER(RtSceneInfo)]
    L.rt_scene_get_bvh.argtypes = [vp, C.POINTER(C.c_float), u32p]
    L.rt_scene_upload.argtypes = [vp, C.c_int32]
    L.rt_image_height.argtypes = [C.POINTER(RtCamera), C.POINTER(C.c_int32)]
    L.rt_render.argtypes = [vp, C.c_int32, C.POINTER(RtCamera), C.c_uint64, C.c_uint64, C.c_uint64,
                            C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(RtStats)]
    L.rt_render_device.argtypes = [vp, C.c_int32, C.POINTER(RtCamera), C.c_uint64, C.c_uint64,
                                   C.c_uint64, vp, vp, vp, C.POINTER(RtStats)]
    L.rt_probe_primary.argtypes = [vp, C.c_int32, C.POINTER(RtCamera), C.c_double, u32p,
                                   C.POINTER(C.c_float)]
    L.rt_probe_surface.argtypes = [vp, C.c_int32, C.POINTER(RtCamera), C.c_double, C.POINTER(C.c_float)]
    L.rt_set_pool_size.argtypes = [vp, C.c_uint32]
    L.rt_context_create.argtypes = [C.POINTER(C.c_int32), C.c_int32, C.POINTER(vp)]
    L.rt_context_destroy.argtypes = [vp]
    L.rt_context_info.argtypes = [vp, C.POINTER(RtContextInfo)]
    L.rt_render_multi.argtypes = [vp, vp, C.POINTER(RtCamera), C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(RtStats)]
    L.rt_options_init.argtypes = [C.POINTER(RtOptions)]
    L.rt_scene_set_options.argtypes = [vp, C.POINTER(RtOptions)]
    L.rt_scene_get_options.argtypes = [vp, C.POINTER(RtOptions)]
    L.rt_tonemap_u8.argtypes = [C.c_int32, C.POINTER(C.c_float), C.c_int32, C.c_int32,
                                C.POINTER(C.c_uint8)]
    L.rt_tonemap_u8_static.argtypes = [C.c_int32, vp, C.c_int32, C.c_int32, vp, vp]
    L.rt_obj_parse.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(vp)]
    L.rt_mesh_sizes.argtypes = [vp, u32p, u32p, u32p]
    L.rt_mesh_data.argtypes = [vp, C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.POINTER(C.c_double)),
                               C.POINTER(C.POINTER(C.c_int32))]
    L.rt_mesh_destroy.argtypes = [vp]
    L.rt_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    L.rt_host_free.argtypes = [vp]
    L.rt_reduce_image.argtypes = [C.c_int32, vp, C.POINTER(vp), C.c_int32, vp, C.c_uint64, C.c_float, vp]
    for name in EXPORTS:
        fn = getattr(L, name)
        if name not in ("rt_version", "rt_last_error"):
            fn.restype = C.c_int32
    _lib = L
    return L


def check(status: int):
    if status != RT_OK:
        msg = lib().rt_last_error()
        msg = msg.decode("utf-8", "replace") if msg else ""
        if status == 3:
            raise UserDefinedError(status, msg)
        raise RtError(status, msg)


def _d3(v) -> D3:
    return D3(float(v[0]), float(v[1]), float(v[2]))


def obj_parse(text: str):
    """rt_obj_parse → raytrace_b200.geometry.Mesh."""
    from .geometry import Mesh
    L = lib()
    data = text.encode("utf-8")
    h = C.c_void_p()
    st = L.rt_obj_parse(data, len(data), C.byref(h))
    if st != RT_OK:
        msg = L.rt_last_error().decode("utf-8", "replace")
        raise ValueError(msg)
    try:
        nv, nt, nf = C.c_uint32(), C.c_uint32(), C.c_uint32()
        check(L.rt_mesh_sizes(h, C.byref(nv), C.byref(nt), C.byref(nf)))
        pv, pt, pf = C.POINTER(C.c_double)(), C.POINTER(C.c_double)(), C.POINTER(C.c_int32)()
        check(L.rt_mesh_data(h, C.byref(pv), C.byref(pt), C.byref(pf)))
        verts = np.ctypeslib.as_array(pv, shape=(nv.value, 3)).copy() if nv.value else np.zeros((0, 3))
        uvs = np.ctypeslib.as_array(pt, shape=(nt.value, 2)).copy() if nt.value else np.zeros((0, 2))
        faces = np.ctypeslib.as_array(pf, shape=(nf.value, 6)).copy() if nf.value else np.zeros((0, 6), np.int32)
    finally:
        L.rt_mesh_destroy(h)
    return Mesh(verts, uvs, faces)


class SceneHandle:
    """Owns an RtScene*; serialises a description tree (raytrace_b200.geometry.Geometry)."""

    def __init__(self):
        self._L = lib()
        self._h = C.c_void_p()
        check(self._L.rt_scene_create(C.byref(self._h)))
        self._tex_ids: Dict[int, int] = {}
        self._mat_ids: Dict[int, int] = {}
        self._node_ids: Dict[int, int] = {}
        self._keep: List[object] = []   # keep numpy buffers alive until commit
        self._rec = None                # first serialisation of a world into a fresh scene: the records in add order
        self.uploaded = set()
        self.committed = False

    def close(self):
        if self._h:
            self._L.rt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- serialisation (memoised on object identity: a shared subtree is sent once) --------
    # Description objects are immutable (frozen dataclasses), so the handle-independent part of each struct - kind, vectors,
    # scalars, matrix, colours - is filled ONCE and kept on the object as bytes (`_rt_desc`); serialising the same world again
    # (another RtScene: every `raytrace` call, every bench step) copies those bytes and sets only what belongs to this scene:
    # child / material / texture ids and array pointers.  demo1 (969 nodes, 462 textures, 484 materials): 15 -> 5 ms.
    @staticmethod
    def _cached(obj, cls, fill):
        raw = obj.__dict__.get("_rt_desc")
        if raw is None:
            d = cls()
            fill(d)
            raw = bytes(d)
            object.__setattr__(obj, "_rt_desc", raw)      # frozen dataclass: a private cache, not a field
        return cls.from_buffer_copy(raw)

    def add_texture(self, tex) -> int:
        key = id(tex)
        if key in self._tex_ids:
            return self._tex_ids[key]

        def fill(d):
            d.kind = int(tex.kind)
            d.i0, d.i1, d.freq = tex.i0, tex.i1, tex.freq
            d.c0, d.c1, d.shift, d.dir = _d3(tex.c0), _d3(tex.c1), _d3(tex.shift), _d3(tex.dir)
        d = self._cached(tex, RtTextureDesc, fill)
        if tex.image is not None:
            img = np.ascontiguousarray(tex.image, dtype=np.float32)
            self._keep.append(img)
            d.height, d.width = img.shape[0], img.shape[1]
            d.image = img.ctypes.data_as(C.POINTER(C.c_float))
        out = C.c_uint32()
        check(self._L.rt_scene_add_texture(self._h, C.byref(d), C.byref(out)))
        if self._rec is not None:
            self._rec[0].append(d)
        self._keep.append(tex)
        self._tex_ids[key] = out.value
        return out.value

    def add_material(self, mat) -> int:
        key = id(mat)
        if key in self._mat_ids:
            return self._mat_ids[key]

        def fill(d):
            d.kind = int(mat.kind)
            d.param = float(mat.param)
        d = self._cached(mat, RtMaterialDesc, fill)
        d.texture = self.add_texture(mat.texture) if mat.texture is not None else 0
        out = C.c_uint32()
        check(self._L.rt_scene_add_material(self._h, C.byref(d), C.byref(out)))
        if self._rec is not None:
            self._rec[1].append(d)
        self._keep.append(mat)
        self._mat_ids[key] = out.value
        return out.value

    def add_node(self, g) -> int:
        key = id(g)
        if key in self._node_ids:
            return self._node_ids[key]
        child_ids = [self.add_node(c) for c in g.children]

        def fill(d):
            d.kind = int(g.kind)
            for k, vec in enumerate(g.v[:3]):
                d.v[k] = _d3(vec)
            for k, sc in enumerate(g.s[:2]):
                d.s[k] = float(sc)
            for k, uv in enumerate(g.uv[:3]):
                d.uv[k][0], d.uv[k][1] = float(uv[0]), float(uv[1])
            if g.m is not None:
                flat = g.m.a.reshape(-1)
                for k in range(16):
                    d.m[k] = float(flat[k])
        d = self._cached(g, RtNodeDesc, fill)
        d.n_children = len(child_ids)
        if child_ids:
            arr = (C.c_uint32 * len(child_ids))(*child_ids)
            self._keep.append(arr)
            d.children = C.cast(arr, C.POINTER(C.c_uint32))
        if g.mesh is not None:
            m = g.mesh
            self._keep.append(m)
            d.n_verts, d.n_uvs, d.n_faces = len(m.verts), len(m.uvs), len(m.faces)
            d.verts = m.verts.ctypes.data_as(C.POINTER(C.c_double))
            if len(m.uvs):
                d.uvs = m.uvs.ctypes.data_as(C.POINTER(C.c_double))
            d.faces = m.faces.ctypes.data_as(C.POINTER(C.c_int32))
        if g.material is not None:
            d.material = self.add_material(g.material)
        out = C.c_uint32()
        check(self._L.rt_scene_add_node(self._h, C.byref(d), C.byref(out)))
        if self._rec is not None:
            self._rec[2].append(d)
        self._keep.append(g)
        self._node_ids[key] = out.value
        return out.value

    def set_noise_tables(self, perm_x, perm_y, perm_z, gradients):
        def ip(a):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=np.int32)
            self._keep.append(a)
            return a.ctypes.data_as(C.POINTER(C.c_int32))
        g = None
        if gradients is not None:
            ga = np.ascontiguousarray(gradients, dtype=np.float64)
            self._keep.append(ga)
            g = ga.ctypes.data_as(C.POINTER(C.c_double))
        check(self._L.rt_scene_set_noise_tables(self._h, ip(perm_x), ip(perm_y), ip(perm_z), g))

    def commit(self, world) -> None:
        # A world is immutable, and in a FRESH scene ids are handed out in add order from 0: the records of its first
        # serialisation (textures, materials, nodes - child / material / texture ids and array pointers included) are
        # kept on the world as three C arrays and go to every later scene in ONE rt_scene_add_batch call instead of a
        # Python walk with one ctypes call per object (demo2: 3.9 -> 0.3 ms per `raytrace` call).
        fresh = not (self._tex_ids or self._mat_ids or self._node_ids)
        flat = world.__dict__.get("_rt_flat") if fresh else None
        if flat is not None:
            tex, mat, nodes, root, keep = flat
            first = (C.c_uint32 * 3)()
            check(self._L.rt_scene_add_batch(self._h, tex, len(tex), mat, len(mat), nodes, len(nodes), first))
            if tuple(first) != (0, 0, 0):
                raise RtError("rt_scene_add_batch: the scene was not empty")
            self._keep.append(flat)
            self._node_ids[id(world)] = root
        else:
            self._rec = ([], [], []) if fresh else None
            try:
                root = self.add_node(world)
                if self._rec is not None:
                    t, m, n = self._rec
                    flat = ((RtTextureDesc * len(t))(*t), (RtMaterialDesc * len(m))(*m), (RtNodeDesc * len(n))(*n), root, list(self._keep))
                    try:
                        object.__setattr__(world, "_rt_flat", flat)
                    except (AttributeError, TypeError):
                        pass
            finally:
                self._rec = None
        check(self._L.rt_scene_commit(self._h, root))
        self.committed = True

    def info(self) -> dict:
        i = RtSceneInfo()
        check(self._L.rt_scene_info(self._h, C.byref(i)))
        return i.as_dict()

    def get_bvh(self):
        i = self.info()
        nodes = np.zeros((i["n_bvh_nodes"], 8), dtype=np.float32)
        canon = np.zeros(i["n_prims"], dtype=np.uint32)
        check(self._L.rt_scene_get_bvh(self._h, nodes.ctypes.data_as(C.POINTER(C.c_float)),
                                       canon.ctypes.data_as(C.POINTER(C.c_uint32))))
        return nodes, canon

    def get_eval_order(self):
        """canonical primitive ids in the reference's evaluation order (the order that decides exact ties)"""
        out = np.zeros(self.info()["n_prims"], dtype=np.uint32)
        check(self._L.rt_scene_get_eval_order(self._h, out.ctypes.data_as(C.POINTER(C.c_uint32))))
        return out

    def upload(self, device: int = 0):
        if device not in self.uploaded:
            check(self._L.rt_scene_upload(self._h, device))
            self.uploaded.add(device)

    def set_pool_size(self, n: int):
        check(self._L.rt_set_pool_size(self._h, int(n)))

    def set_options(self, **kw):
        """Update RtOptions fields by name (None / negative = library default).  Commit-time fields (bvh_reference,
        leaf_max, collapse, aggregate_boxes, shared_boundary, interior_bound, world_box, split_moving, sah_node_cost)
        must be set before ``commit``."""
        o = RtOptions()
        check(self._L.rt_scene_get_options(self._h, C.byref(o)))
        for k, v in kw.items():
            if k not in RtOptions.NAMES:
                raise KeyError(f"unknown option {k!r}")
            setattr(o, k, -1 if v is None else (float(v) if k == "sah_node_cost" else int(v)))
        check(self._L.rt_scene_set_options(self._h, C.byref(o)))

    def get_options(self) -> dict:
        o = RtOptions()
        check(self._L.rt_scene_get_options(self._h, C.byref(o)))
        return {k: getattr(o, k) for k in RtOptions.NAMES}

    @property
    def handle(self):
        return self._h


def make_camera(settings, bg_kind: int, bg_c0, bg_c1):
    """CameraSettings → RtCamera (+ the ctypes array that must outlive it)."""
    cam = RtCamera()
    cam.center, cam.look_at, cam.up = _d3(settings.cs_center), _d3(settings.cs_lookAt), _d3(settings.cs_up)
    cam.vfov, cam.aspect_ratio = float(settings.cs_vfov), float(settings.cs_aspectRatio)
    cam.defocus_angle, cam.focus_dist = float(settings.cs_defocusAngle), float(settings.cs_focusDist)
    cam.bg_kind, cam.bg_c0, cam.bg_c1 = int(bg_kind), _d3(bg_c0), _d3(bg_c1)
    cam.image_width = int(settings.cs_imageWidth)
    cam.samples_per_pixel = int(settings.cs_samplesPerPixel)
    cam.max_recursion_depth = int(settings.cs_maxRecursionDepth)
    tg = list(settings.cs_redirectTargets)
    cam.n_redirect_targets = len(tg)
    arr = (RtRedirectTarget * max(1, len(tg)))()
    for k, (p, q, s0, s1) in enumerate(tg):
        arr[k].p, arr[k].q, arr[k].s0, arr[k].s1 = float(p), _d3(q), _d3(s0), _d3(s1)
    cam.redirect_targets = C.cast(arr, C.POINTER(RtRedirectTarget))
    cam._keepalive = arr  # type: ignore[attr-defined]
    return cam
