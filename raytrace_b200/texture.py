"""Host-side mirror of Graphics.Ray.Texture (reference: src/Graphics/Ray/Texture.hs).

In the reference a ``Texture`` is a closure ``Point3 -> V2 Double -> Color`` (Texture.hs:15).
Here every library constructor is reified as data so that it can be flattened to a device
record; textures built from user functions (``solidTexture`` / ``uvTexture`` / ``Texture``)
are representable but are rejected at commit time with RT_ERR_USER_DEFINED — no CPU fallback.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Optional

import numpy as np

from .core import Vec3, as_vec3

# RtTextureKind (include/rt_b200.h)
TEX_CONSTANT, TEX_CHECKER, TEX_IMAGE, TEX_NOISE, TEX_MARBLE, TEX_USER = 1, 2, 3, 4, 5, 6


@dataclass(frozen=True, eq=False)
class Texture:
    """Reified ``Texture``. Calling ``Texture(fn)`` with a Python callable mirrors the raw
    constructor (Texture.hs:15) and yields a user-defined texture."""
    kind: Any = TEX_USER
    i0: int = 0
    i1: int = 0
    freq: float = 0.0
    c0: Vec3 = (0.0, 0.0, 0.0)
    c1: Vec3 = (0.0, 0.0, 0.0)
    shift: Vec3 = (0.0, 0.0, 0.0)
    dir: Vec3 = (0.0, 0.0, 0.0)
    image: Optional[np.ndarray] = field(default=None, repr=False)
    user_fn: Optional[Callable] = field(default=None, repr=False)

    def __post_init__(self):
        if callable(self.kind):  # Texture(\p uv -> ...)
            object.__setattr__(self, "user_fn", self.kind)
            object.__setattr__(self, "kind", TEX_USER)


def constantTexture(color) -> Texture:
    """Texture.hs:18-19."""
    return Texture(kind=TEX_CONSTANT, c0=as_vec3(color))


def solidTexture(f: Callable) -> Texture:
    """Texture.hs:22-23 — wraps an arbitrary function of the point: user-defined."""
    return Texture(kind=TEX_USER, user_fn=f)


def uvTexture(f: Callable) -> Texture:
    """Texture.hs:26-27 — wraps an arbitrary function of (u, v): user-defined."""
    return Texture(kind=TEX_USER, user_fn=f)


def imageTexture(image) -> Texture:
    """Texture.hs:31-41. ``image`` is an H×W×3 array of LINEAR RGB (what ``readImage`` returns,
    Ray.hs:241-245), row 0 = top row."""
    arr = np.ascontiguousarray(np.asarray(image, dtype=np.float32))
    if arr.ndim != 3 or arr.shape[2] != 3:
        raise ValueError("imageTexture: expected an H x W x 3 array of linear RGB")
    return Texture(kind=TEX_IMAGE, image=arr)


def checkerTexture(n_u: int, n_v: int, c0, c1) -> Texture:
    """Texture.hs:45-53."""
    return Texture(kind=TEX_CHECKER, i0=int(n_u), i1=int(n_v), c0=as_vec3(c0), c1=as_vec3(c1))


def noiseTexture(k: int, freq: float, shift, color0, color1) -> Texture:
    """Texture.hs:56-67."""
    return Texture(kind=TEX_NOISE, i0=int(k), freq=float(freq), shift=as_vec3(shift),
                   c0=as_vec3(color0), c1=as_vec3(color1))


def marbleTexture(dir, freq: float, shift) -> Texture:
    """Texture.hs:70-79."""
    return Texture(kind=TEX_MARBLE, freq=float(freq), shift=as_vec3(shift), dir=as_vec3(dir))
