"""ctypes harness for the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Feeds the SAME description tree objects (raytrace_b200.geometry.Geometry, …) that the product
serialises to librt_b200.so into the double-precision restatement of the reference, so the
two sides render identical scenes.  May be imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / `--impl reference` legs — never by raytrace_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Optional

import numpy as np

from raytrace_b200 import _ffi as F  # struct layouts of include/rt_b200.h only
from raytrace_b200.ray import CameraSettings, classify_background, imageHeight

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")


class OrcCounters(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("box_tests", C.c_uint64),
                ("sphere_tests", C.c_uint64), ("planar_tests", C.c_uint64),
                ("medium_queries", C.c_uint64), ("seconds", C.c_double),
                ("threads", C.c_int32), ("_pad", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "_pad"}


_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"])
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(LIB_PATH)
    vp, u32p, dp = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_double)
    L.orc_scene_create.restype = vp
    L.orc_scene_destroy.argtypes = [vp]
    L.orc_last_error.argtypes = [vp]
    L.orc_last_error.restype = C.c_char_p
    L.orc_scene_set_noise_tables.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), dp]
    L.orc_noise_tables.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), dp]
    L.orc_scene_add_texture.argtypes = [vp, C.POINTER(F.RtTextureDesc), u32p]
    L.orc_scene_add_material.argtypes = [vp, C.POINTER(F.RtMaterialDesc), u32p]
    L.orc_scene_add_node.argtypes = [vp, C.POINTER(F.RtNodeDesc), u32p]
    L.orc_scene_commit.argtypes = [vp, C.c_uint32]
    L.orc_scene_num_prims.argtypes = [vp]
    L.orc_scene_num_prims.restype = C.c_uint32
    L.orc_scene_leaf_order.argtypes = [vp, u32p, C.c_uint32, u32p]
    L.orc_image_height.argtypes = [C.POINTER(F.RtCamera)]
    L.orc_render.argtypes = [vp, C.POINTER(F.RtCamera), C.c_uint64, C.c_int32, C.c_int32, dp, dp,
                             C.POINTER(OrcCounters)]
    L.orc_probe_primary.argtypes = [vp, C.POINTER(F.RtCamera), C.c_double, C.c_double, C.c_double, u32p, dp]
    L.orc_probe_surface.argtypes = [vp, C.POINTER(F.RtCamera), C.c_double, dp]
    L.orc_perlin.argtypes = [vp, C.c_double, C.c_double, C.c_double]
    L.orc_perlin.restype = C.c_double
    L.orc_fractal.argtypes = [vp, C.c_int32, C.c_double, C.c_double, C.c_double]
    L.orc_fractal.restype = C.c_double
    L.orc_texture_eval.argtypes = [vp, C.c_uint32, dp, C.c_double, C.c_double, dp]
    L.orc_hit.argtypes = [vp, dp, dp, C.c_double, C.c_double, C.c_double, dp, dp, dp,
                          C.POINTER(C.c_int32), dp, u32p]
    L.orc_splitmix_next.argtypes = [C.c_uint64, C.c_int32]
    L.orc_splitmix_next.restype = C.c_uint64
    _lib = L
    return L


class OracleScene:
    """Same serialisation protocol as raytrace_b200._ffi.SceneHandle, against orc_* entry points."""

    def __init__(self, world=None):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_scene_create())
        self._tex: Dict[int, int] = {}
        self._mat: Dict[int, int] = {}
        self._node: Dict[int, int] = {}
        self._keep: List[object] = []
        if world is not None:
            self.commit(world)

    def __del__(self):
        try:
            if self.h:
                self.L.orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def _check(self, st: int):
        if st != 0:
            msg = self.L.orc_last_error(self.h)
            raise F.RtError(st, (msg or b"").decode())

    def add_texture(self, tex) -> int:
        if id(tex) in self._tex:
            return self._tex[id(tex)]
        d = F.RtTextureDesc()
        d.kind, d.i0, d.i1, d.freq = int(tex.kind), tex.i0, tex.i1, tex.freq
        d.c0, d.c1, d.shift, d.dir = F._d3(tex.c0), F._d3(tex.c1), F._d3(tex.shift), F._d3(tex.dir)
        if tex.image is not None:
            img = np.ascontiguousarray(tex.image, dtype=np.float32)
            self._keep.append(img)
            d.height, d.width = img.shape[0], img.shape[1]
            d.image = img.ctypes.data_as(C.POINTER(C.c_float))
        out = C.c_uint32()
        self._check(self.L.orc_scene_add_texture(self.h, C.byref(d), C.byref(out)))
        self._keep.append(tex)
        self._tex[id(tex)] = out.value
        return out.value

    def add_material(self, mat) -> int:
        if id(mat) in self._mat:
            return self._mat[id(mat)]
        d = F.RtMaterialDesc()
        d.kind, d.param = int(mat.kind), float(mat.param)
        d.texture = self.add_texture(mat.texture) if mat.texture is not None else 0
        out = C.c_uint32()
        self._check(self.L.orc_scene_add_material(self.h, C.byref(d), C.byref(out)))
        self._keep.append(mat)
        self._mat[id(mat)] = out.value
        return out.value

    def add_node(self, g) -> int:
        if id(g) in self._node:
            return self._node[id(g)]
        child_ids = [self.add_node(c) for c in g.children]
        d = F.RtNodeDesc()
        d.kind, d.n_children = int(g.kind), len(child_ids)
        if child_ids:
            arr = (C.c_uint32 * len(child_ids))(*child_ids)
            self._keep.append(arr)
            d.children = C.cast(arr, C.POINTER(C.c_uint32))
        for k, vec in enumerate(g.v[:3]):
            d.v[k] = F._d3(vec)
        for k, sc in enumerate(g.s[:2]):
            d.s[k] = float(sc)
        for k, uv in enumerate(g.uv[:3]):
            d.uv[k][0], d.uv[k][1] = float(uv[0]), float(uv[1])
        if g.m is not None:
            flat = g.m.a.reshape(-1)
            for k in range(16):
                d.m[k] = float(flat[k])
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
        self._check(self.L.orc_scene_add_node(self.h, C.byref(d), C.byref(out)))
        self._keep.append(g)
        self._node[id(g)] = out.value
        return out.value

    def commit(self, world):
        root = self.add_node(world)
        self._check(self.L.orc_scene_commit(self.h, root))

    # -- queries ---------------------------------------------------------------------------------
    def camera(self, settings: CameraSettings):
        kind, c0, c1 = classify_background(settings.cs_background)
        return F.make_camera(settings, kind, c0, c1)

    def num_prims(self) -> int:
        return int(self.L.orc_scene_num_prims(self.h))

    def leaf_order(self) -> np.ndarray:
        n = C.c_uint32()
        cap = max(1, self.num_prims() * 4)
        out = np.zeros(cap, dtype=np.uint32)
        self._check(self.L.orc_scene_leaf_order(self.h, out.ctypes.data_as(C.POINTER(C.c_uint32)), cap, C.byref(n)))
        return out[: n.value].copy()

    def noise_tables(self):
        px, py, pz = (np.zeros(256, np.int32) for _ in range(3))
        g = np.zeros((256, 3), np.float64)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        self.L.orc_noise_tables(self.h, ip(px), ip(py), ip(pz), g.ctypes.data_as(C.POINTER(C.c_double)))
        return px, py, pz, g

    def render(self, settings: CameraSettings, seed: int, spp: Optional[int] = None, threads: int = 0,
               want_sumsq: bool = False):
        """→ (mean image H×W×3 float64, [sumsq], counters dict)."""
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        mean = np.zeros((h, w, 3), np.float64)
        sq = np.zeros((h, w, 3), np.float64) if want_sumsq else None
        cnt = OrcCounters()
        dp = C.POINTER(C.c_double)
        self._check(self.L.orc_render(self.h, C.byref(cam), C.c_uint64(int(seed) & (2**64 - 1)), int(spp or 0), int(threads),
                                      mean.ctypes.data_as(dp), sq.ctypes.data_as(dp) if sq is not None else None,
                                      C.byref(cnt)))
        if want_sumsq:
            return mean, sq, cnt.as_dict()
        return mean, cnt.as_dict()

    def probe_primary(self, settings: CameraSettings, time: float = 0.5, dx: float = 0.5, dy: float = 0.5):
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        ids = np.zeros((h, w), np.uint32)
        ts = np.zeros((h, w), np.float64)
        self._check(self.L.orc_probe_primary(self.h, C.byref(cam), float(time), float(dx), float(dy),
                                             ids.ctypes.data_as(C.POINTER(C.c_uint32)),
                                             ts.ctypes.data_as(C.POINTER(C.c_double))))
        return ids, ts

    def probe_surface(self, settings: CameraSettings, time: float = 0.5):
        """(H, W, 10): normal(3), uv(2), front, texture colour(3), material kind of the pixel-centre primary hit."""
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        out = np.zeros((h, w, 10), np.float64)
        self._check(self.L.orc_probe_surface(self.h, C.byref(cam), float(time), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def probe_tilted(self, settings: CameraSettings, time: float = 0.5, delta: float = None, eps_rad: float = 2e-6):
        """The pixel-centre probe and the same probe tilted by ±eps_rad in x and y (2e-6 rad ~ 16 fp32 ulps of a
        direction component; `delta` overrides it in pixel units): ids (5, H, W) uint32 and ts (5, H, W) float64,
        index 0 = the centre ray."""
        import math
        if delta is None:
            h = imageHeight(settings)
            pixel_angle = 2.0 * math.tan(settings.cs_vfov / 2.0) / h
            delta = eps_rad / pixel_angle
        out_i, out_t = [], []
        for ox, oy in ((0.0, 0.0), (-delta, 0.0), (delta, 0.0), (0.0, -delta), (0.0, delta)):
            i2, t2 = self.probe_primary(settings, time, 0.5 + ox, 0.5 + oy)
            out_i.append(i2)
            out_t.append(t2)
        return np.stack(out_i), np.stack(out_t)

    def probe_stable(self, settings: CameraSettings, time: float = 0.5, delta: float = None, t_tol: float = 1e-4,
                     eps_rad: float = 2e-6, tilted=None):
        """Centre probe plus the mask of WELL-CONDITIONED pixels: the oracle's own answer (primitive id, and t
        within t_tol relative) does not change when the probe ray is tilted by ±eps_rad.  A pixel that passes has a
        condition number (dt/t)/dtheta < t_tol/eps_rad = 50, so fp32 input rounding alone (~1.2e-7 rad) cannot move t by
        1e-5.  Pixels whose centre ray runs through a primitive edge or grazes a silhouette fail this; their reference
        answer is fp64 rounding noise (README scene column 46: a = 2.2e-16 on the mirror's edge)."""
        ids5, ts5 = tilted if tilted is not None else self.probe_tilted(settings, time, delta, eps_rad)
        ids, ts = ids5[0], ts5[0]
        stable = np.ones(ids.shape, bool)
        for i2, t2 in zip(ids5[1:], ts5[1:]):
            stable &= i2 == ids
            hit = (ids != 0xFFFFFFFF) & (i2 == ids)
            with np.errstate(invalid="ignore", divide="ignore"):
                stable &= ~hit | (np.abs(t2 - ts) <= t_tol * np.abs(ts))
        return ids, ts, stable

    def perlin(self, x, y, z) -> float:
        return float(self.L.orc_perlin(self.h, x, y, z))

    def fractal(self, depth, x, y, z) -> float:
        return float(self.L.orc_fractal(self.h, depth, x, y, z))

    def texture_eval(self, tex, p, u=0.0, v=0.0):
        tid = self.add_texture(tex)
        pp = (C.c_double * 3)(*map(float, p))
        out = (C.c_double * 3)()
        self._check(self.L.orc_texture_eval(self.h, tid, pp, float(u), float(v), out))
        return tuple(out)

    def hit(self, origin, direction, time=0.0, tmin=1e-4, tmax=float("inf")):
        o = (C.c_double * 3)(*map(float, origin))
        d = (C.c_double * 3)(*map(float, direction))
        t = C.c_double()
        p, n = (C.c_double * 3)(), (C.c_double * 3)()
        uv = (C.c_double * 2)()
        front, prim = C.c_int32(), C.c_uint32()
        r = self.L.orc_hit(self.h, o, d, time, tmin, tmax, C.byref(t), p, n, C.byref(front), uv, C.byref(prim))
        if not r:
            return None
        return dict(t=t.value, p=tuple(p), n=tuple(n), front=bool(front.value), uv=tuple(uv), prim=prim.value)


def primary_gate(gpu_ids: np.ndarray, gpu_ts: np.ndarray, ids5: np.ndarray, ts5: np.ndarray, t_tol: float = 1e-5) -> dict:
    """north_star gates 1 + 2 over ALL pixels of a primary-ray probe (no pixel is masked out).

    A pixel PASSES when the GPU's primitive id is one the oracle itself returns for that pixel - at the centre ray or at
    one of the four ±2e-6 rad tilts (`OracleScene.probe_tilted`) - AND the GPU's hit distance is within t_tol relative
    of the oracle's t for THAT id (a miss must be a miss: t = inf).  Pixels whose id equals the centre answer are `raw`
    matches; pixels that only match a tilted answer are `tie_explained` (the centre ray runs through a primitive edge:
    which side wins is fp64 rounding noise in the reference itself, Geometry.hs:340-345,356-361 + Core.hs:83-84); every
    other pixel is `unexplained`.  The gate is pass_frac >= 0.9999."""
    MISS = 0xFFFFFFFF
    gi = np.asarray(gpu_ids)
    gt = np.asarray(gpu_ts, dtype=np.float64)
    same = ids5 == gi[None]                                   # (5, H, W): the oracle returns the GPU's id here
    with np.errstate(invalid="ignore", divide="ignore"):
        rel = np.abs(gt[None] - ts5) / np.abs(ts5)
    t_ok = np.where(gi[None] == MISS, np.isinf(gt)[None], rel <= t_tol)
    ok5 = same & t_ok
    raw = same[0]
    passed = ok5.any(axis=0)
    id_ok = same.any(axis=0)
    rel_centre = np.where(raw & (gi != MISS), rel[0], 0.0)
    n = gi.size
    return {
        "pixels": int(n), "raw_match": int(raw.sum()), "raw_frac": float(raw.mean()),
        "tie_explained": int((passed & ~raw).sum()),          # id differs at the centre, equals a tilted answer, t agrees
        "id_unexplained": int((~id_ok).sum()),                # id the oracle never returns around this pixel
        "t_failed": int((id_ok & ~passed).sum()),             # id fine, distance off by more than t_tol
        "unexplained": int((~passed).sum()), "pass_frac": float(passed.mean()),
        "max_rel_dt_raw": float(np.nanmax(rel_centre)) if n else 0.0,
        "passed_mask": passed,
    }
