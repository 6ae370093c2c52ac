"""Host-side mirror of Graphics.Ray.Geometry (reference: src/Graphics/Ray/Geometry.hs).

The reference's ``Geometry m a`` is a bounding box plus a hit closure (Geometry.hs:42).  Here
each combinator builds a node of a DESCRIPTION TREE (a deep embedding of the same
vocabulary); ``raytrace`` serialises that tree through the C ABI (include/rt_b200.h), where
it is validated, flattened and rendered on the GPU.  Nothing in this module intersects rays.

The two type parameters of ``Geometry m a`` are tracked dynamically so that ill-typed scene
programs fail where GHC would reject them:
  * ``monad``   — "Identity" or "State" (``State StdGen``; only ``constantMedium`` introduces it)
  * ``payload`` — "unit" (``()``) or "material" (after ``material <$ geometry``)
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from .core import (Box, Vec3, allCorners, as_vec3, boxHull, boxJoin, fromCorners, padBox,
                   shiftBox, vadd, vneg, vsub)

# RtNodeKind (include/rt_b200.h)
(NODE_SPHERE, NODE_PARALLELOGRAM, NODE_TRIANGLE, NODE_CUBOID, NODE_MESH, NODE_GROUP,
 NODE_BVH_NODE, NODE_BVH_TREE, NODE_TRANSFORM, NODE_MOVING, NODE_MEDIUM, NODE_WITH_MATERIAL,
 NODE_PURE, NODE_USER) = range(1, 15)


# ---- 4x4 matrices (linear's M44 Double; Geometry.hs:394-440) -----------------------------

class M44:
    """Row-major 4x4 matrix; ``a @ b`` is linear's ``a !*! b``."""
    __slots__ = ("a",)

    def __init__(self, rows):
        self.a = np.array(rows, dtype=np.float64).reshape(4, 4)

    def __matmul__(self, other: "M44") -> "M44":
        return M44(self.a @ other.a)

    def apply_point(self, p: Vec3) -> Vec3:
        v = self.a @ np.array([p[0], p[1], p[2], 1.0])
        return (float(v[0]), float(v[1]), float(v[2]))

    def __repr__(self):
        return f"M44({self.a.tolist()})"


def translate(v) -> M44:
    x, y, z = as_vec3(v)
    return M44([[1, 0, 0, x], [0, 1, 0, y], [0, 0, 1, z], [0, 0, 0, 1]])


def rotateX(angle: float) -> M44:
    c, s = math.cos(angle), math.sin(angle)
    return M44([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1]])


def rotateY(angle: float) -> M44:
    c, s = math.cos(angle), math.sin(angle)
    return M44([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1]])


def rotateZ(angle: float) -> M44:
    c, s = math.cos(angle), math.sin(angle)
    return M44([[c, -s, 0, 0], [s, c, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]])


def scale(a: float) -> M44:
    """Geometry.hs:434-440 — must not be used with ``transform`` (commit rejects it)."""
    return M44([[a, 0, 0, 0], [0, a, 0, 0], [0, 0, a, 0], [0, 0, 0, 1]])


# ---- meshes (Geometry.hs:179-190) -----------------------------------------------------------

@dataclass(frozen=True, eq=False)
class Mesh:
    """``Mesh verts uvs faces``: faces is an (n, 6) int array (v0,t0,v1,t1,v2,t2), 0-based,
    t = -1 standing for ``Nothing``."""
    verts: np.ndarray
    uvs: np.ndarray
    faces: np.ndarray

    def __post_init__(self):
        object.__setattr__(self, "verts", np.ascontiguousarray(self.verts, dtype=np.float64).reshape(-1, 3))
        object.__setattr__(self, "uvs", np.ascontiguousarray(self.uvs, dtype=np.float64).reshape(-1, 2))
        object.__setattr__(self, "faces", np.ascontiguousarray(self.faces, dtype=np.int32).reshape(-1, 6))


def transformVertices(m: M44, mesh: Mesh) -> Mesh:
    """Geometry.hs:187-190."""
    v = mesh.verts @ m.a[:3, :3].T + m.a[:3, 3]
    return Mesh(v, mesh.uvs, mesh.faces)


def parseObj(text: str) -> Mesh:
    """Geometry.hs:207-285, implemented in the library's C++ loader (rt_obj_parse).
    Raises ValueError carrying the reference-style message on a parse error
    (the reference returns ``Left msg``)."""
    from . import _ffi
    return _ffi.obj_parse(text)


def readObj(path: str) -> Mesh:
    """Geometry.hs:194-195 (error message is prefixed with the path)."""
    with open(path, "r") as f:
        text = f.read()
    try:
        return parseObj(text)
    except ValueError as e:
        raise ValueError(f"{path}, {e}") from None


# ---- the description tree --------------------------------------------------------------------

@dataclass(frozen=True, eq=False)
class Geometry:
    kind: int
    monad: str = "Identity"
    payload: str = "unit"
    bbox: Optional[Box] = None
    children: Tuple["Geometry", ...] = ()
    v: Tuple[Vec3, ...] = ()
    s: Tuple[float, ...] = ()
    uv: Tuple[Tuple[float, float], ...] = ()
    m: Optional[M44] = None
    mesh: Optional[Mesh] = field(default=None, repr=False)
    material: Optional[object] = None
    user_fn: Optional[Callable] = field(default=None, repr=False)

    # `fmap`/`<$` spelled from the geometry side:  geometry.withMaterial(mat)
    def withMaterial(self, mat):
        return withMaterial(mat, self)


def boundingBox(g: Geometry) -> Box:
    """Geometry.hs:54-55."""
    return g.bbox  # type: ignore[return-value]


def pureGeometry(g: Geometry) -> Geometry:
    """Geometry.hs:50-51."""
    if g.monad != "Identity":
        raise TypeError("pureGeometry: expected Geometry Identity a")
    return Geometry(NODE_PURE, monad="State", payload=g.payload, bbox=g.bbox, children=(g,))


def withMaterial(mat, g: Geometry) -> Geometry:
    """``mat <$ g`` (Functor instance, Geometry.hs:44-47)."""
    return Geometry(NODE_WITH_MATERIAL, monad=g.monad, payload="material", bbox=g.bbox,
                    children=(g,), material=mat)


def sphere(center, radius: float) -> Geometry:
    """Geometry.hs:58-94."""
    c = as_vec3(center)
    r = float(radius)
    d = (r, r, r)
    return Geometry(NODE_SPHERE, bbox=fromCorners(vsub(c, d), vadd(c, d)), v=(c,), s=(r,))


def planeShape(q, u, v, test: Callable, getUV: Callable, bbox: Box) -> Geometry:
    """Geometry.hs:109-144 with arbitrary predicate / uv functions: user-defined, rejected at
    commit.  ``parallelogram`` and ``triangle`` are the library's own instances."""
    return Geometry(NODE_USER, bbox=padBox(0.0001, bbox), v=(as_vec3(q), as_vec3(u), as_vec3(v)),
                    user_fn=test)


def parallelogram(q, u, v) -> Geometry:
    """Geometry.hs:147-151."""
    q, u, v = as_vec3(q), as_vec3(u), as_vec3(v)
    bbox = boxHull([q, vadd(q, u), vadd(q, v), vadd(vadd(q, u), v)])
    return Geometry(NODE_PARALLELOGRAM, bbox=padBox(0.0001, bbox), v=(q, u, v))


def cuboid(box: Box) -> Geometry:
    """Geometry.hs:154-166 (a `group` of six parallelograms; face order fixed there)."""
    (x0, x1), (y0, y1), (z0, z1) = box
    # bbox of the group = join of the six padded face boxes
    bbox = padBox(0.0001, ((x0, x1), (y0, y1), (z0, z1)))
    return Geometry(NODE_CUBOID, bbox=bbox, v=((float(x0), float(y0), float(z0)),
                                              (float(x1), float(y1), float(z1))))


def triangle(a, b, c) -> Geometry:
    """Geometry.hs:169-176; each argument is (point, (u, v))."""
    (p0, uv0), (p1, uv1), (p2, uv2) = a, b, c
    p0, p1, p2 = as_vec3(p0), as_vec3(p1), as_vec3(p2)
    bbox = padBox(0.0001, boxHull([p0, p1, p2]))
    return Geometry(NODE_TRIANGLE, bbox=bbox, v=(p0, p1, p2),
                    uv=(tuple(map(float, uv0)), tuple(map(float, uv1)), tuple(map(float, uv2))))


def triangleMesh(mesh: Mesh) -> Geometry:
    """Geometry.hs:288-294 (a `bvhTree` of triangles in face order)."""
    if len(mesh.faces) == 0:
        raise ValueError("bvhTree: empty list")
    vi = mesh.faces[:, 0::2].reshape(-1)
    used = mesh.verts[vi]
    lo, hi = used.min(axis=0) - 0.0001, used.max(axis=0) + 0.0001
    bbox = tuple((float(a), float(b)) for a, b in zip(lo, hi))
    return Geometry(NODE_MESH, bbox=bbox, mesh=mesh)


def constantMedium(density: float, boundary: Geometry) -> Geometry:
    """Geometry.hs:298-330."""
    if boundary.monad != "Identity" or boundary.payload != "unit":
        raise TypeError("constantMedium: boundary must be a Geometry Identity ()")
    return Geometry(NODE_MEDIUM, monad="State", payload="unit", bbox=boundary.bbox,
                    children=(boundary,), s=(float(density),))


def _check_same(name: str, obs: Sequence[Geometry]):
    monads = {o.monad for o in obs}
    payloads = {o.payload for o in obs}
    if len(monads) > 1:
        raise TypeError(f"{name}: children mix Geometry Identity and Geometry (State StdGen); use pureGeometry")
    if len(payloads) > 1:
        raise TypeError(f"{name}: children mix Geometry m () and Geometry m Material")


def group(obs: Sequence[Geometry]) -> Geometry:
    """Geometry.hs:336-347."""
    obs = tuple(obs)
    if not obs:
        raise ValueError("group: empty list (boxJoin of an empty list is undefined in the reference)")
    _check_same("group", obs)
    return Geometry(NODE_GROUP, monad=obs[0].monad, payload=obs[0].payload,
                    bbox=boxJoin(o.bbox for o in obs), children=obs)


def bvhNode(left: Geometry, right: Geometry) -> Geometry:
    """Geometry.hs:352-363."""
    _check_same("bvhNode", (left, right))
    return Geometry(NODE_BVH_NODE, monad=left.monad, payload=left.payload,
                    bbox=boxJoin([left.bbox, right.bbox]), children=(left, right))


def bvhTree(obs: Sequence[Geometry]) -> Geometry:
    """Geometry.hs:369-377.  The tree itself is built by the library at commit time with the
    reference's split policy; an empty list is the reference's runtime error."""
    obs = tuple(obs)
    if not obs:
        raise ValueError("bvhTree: empty list")
    _check_same("bvhTree", obs)
    return Geometry(NODE_BVH_TREE, monad=obs[0].monad, payload=obs[0].payload,
                    bbox=boxJoin(o.bbox for o in obs), children=obs)


def transform(m: M44, g: Geometry) -> Geometry:
    """Geometry.hs:382-391."""
    corners = [m.apply_point(c) for c in allCorners(g.bbox)]
    return Geometry(NODE_TRANSFORM, monad=g.monad, payload=g.payload, bbox=boxHull(corners),
                    children=(g,), m=m)


def moving(v0, v1, g: Geometry) -> Geometry:
    """Geometry.hs:449-456."""
    v0, v1 = as_vec3(v0), as_vec3(v1)
    bbox = boxJoin([shiftBox(v0, g.bbox), shiftBox(v1, g.bbox)])
    return Geometry(NODE_MOVING, monad=g.monad, payload=g.payload, bbox=bbox, children=(g,),
                    v=(v0, v1))
