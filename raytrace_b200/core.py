"""Host-side mirror of Graphics.Ray.Core (reference: src/Graphics/Ray/Core.hs).

Only the names a scene program needs are mirrored: numbers, vectors, intervals, boxes.
Vectors are plain 3-tuples of Python floats (the reference's ``V3 Double``); all arithmetic
here is host-side scene description, never the render path.
"""
from __future__ import annotations

import math
from typing import Iterable, Sequence, Tuple

Vec3 = Tuple[float, float, float]
Interval = Tuple[float, float]
Box = Tuple[Interval, Interval, Interval]

infinity = math.inf  # Core.hs:22-23


def degrees(x: float) -> float:
    """``degrees x`` = x degrees in radians (Core.hs:26-27)."""
    return x * math.pi / 180.0


def V3(x, y=None, z=None) -> Vec3:
    """``V3 x y z``; a single number broadcasts like linear's ``Num (V3 a)`` literals."""
    if y is None and z is None:
        if isinstance(x, (int, float)):
            return (float(x), float(x), float(x))
        x, y, z = x
    return (float(x), float(y), float(z))


def as_vec3(v) -> Vec3:
    if isinstance(v, (int, float)):
        return (float(v), float(v), float(v))
    x, y, z = v
    return (float(x), float(y), float(z))


def vadd(a: Vec3, b: Vec3) -> Vec3:
    return (a[0] + b[0], a[1] + b[1], a[2] + b[2])


def vsub(a: Vec3, b: Vec3) -> Vec3:
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def vscale(s: float, a: Vec3) -> Vec3:
    return (s * a[0], s * a[1], s * a[2])


def vneg(a: Vec3) -> Vec3:
    return (-a[0], -a[1], -a[2])


def dot(a: Vec3, b: Vec3) -> float:
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def cross(a: Vec3, b: Vec3) -> Vec3:
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def norm(a: Vec3) -> float:
    return math.sqrt(dot(a, a))


# ---- intervals and boxes (Core.hs:76-144) ------------------------------------------------

def midpoint(i: Interval) -> float:
    return (i[0] + i[1]) / 2.0


def padInterval(padding: float, i: Interval) -> Interval:
    return (i[0] - padding, i[1] + padding)


def fromCorners(p: Vec3, q: Vec3) -> Box:
    """Core.hs:112-113."""
    p, q = as_vec3(p), as_vec3(q)
    return tuple((a, b) if a < b else (b, a) for a, b in zip(p, q))  # type: ignore[return-value]


def boxJoin(boxes: Iterable[Box]) -> Box:
    """Core.hs:116-117 (foldl1': errors on an empty list, like the reference)."""
    it = iter(boxes)
    try:
        acc = next(it)
    except StopIteration:
        raise ValueError("boxJoin: empty list")
    for b in it:
        acc = tuple((min(a[0], c[0]), max(a[1], c[1])) for a, c in zip(acc, b))  # type: ignore[assignment]
    return acc


def boxHull(pts: Sequence[Vec3]) -> Box:
    """Core.hs:120-125."""
    return tuple((min(p[k] for p in pts), max(p[k] for p in pts)) for k in range(3))  # type: ignore[return-value]


def allCorners(b: Box):
    """Core.hs:128-132 (x outermost, z innermost; fst before snd)."""
    return [(fx, fy, fz) for fx in b[0] for fy in b[1] for fz in b[2]]


def padBox(padding: float, b: Box) -> Box:
    return tuple(padInterval(padding, i) for i in b)  # type: ignore[return-value]


def shiftBox(v: Vec3, b: Box) -> Box:
    return tuple((i[0] + x, i[1] + x) for x, i in zip(v, b))  # type: ignore[return-value]


def longestDim(b: Box) -> int:
    """Core.hs:143-144 with argMax's tie rule (Core.hs:43-46). Returns 0/1/2 for X/Y/Z."""
    x, y, z = (i[1] - i[0] for i in b)
    if x > y:
        return 0 if x > z else 2
    return 1 if y > z else 2
