"""Host-side mirror of Graphics.Ray.Material (reference: src/Graphics/Ray/Material.hs).

A reference ``Material`` is a closure ``Vec3 -> HitRecord -> (Color, State StdGen MaterialResult)``
(Material.hs:17).  The ten library constructors are reified as data; ``Material(fn)`` with a
user function is representable but rejected at commit (RT_ERR_USER_DEFINED), as north_star
requires — there is no CPU fallback.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Optional

from .texture import Texture, constantTexture

# RtMaterialKind (include/rt_b200.h)
(MAT_LIGHT, MAT_PITCH_BLACK, MAT_LAMBERTIAN, MAT_LOMMEL_SEELIGER, MAT_MIRROR, MAT_METAL,
 MAT_DIELECTRIC, MAT_TRANSPARENT, MAT_ISOTROPIC, MAT_ANISOTROPIC, MAT_USER) = range(1, 12)


@dataclass(frozen=True, eq=False)
class Material:
    kind: Any = MAT_USER
    texture: Optional[Texture] = None
    param: float = 0.0
    user_fn: Optional[Callable] = field(default=None, repr=False)

    def __post_init__(self):
        if callable(self.kind):  # Material(\dir hit -> ...)
            object.__setattr__(self, "user_fn", self.kind)
            object.__setattr__(self, "kind", MAT_USER)

    # `material <$ geometry` (Geometry.hs:44-47); Python spelling: material << geometry
    def __lshift__(self, geometry):
        from .geometry import withMaterial
        return withMaterial(self, geometry)


_BLACK = constantTexture(0.0)


def lightSource(tex: Texture) -> Material:
    """Material.hs:41-42."""
    return Material(MAT_LIGHT, tex)


pitchBlack = Material(MAT_PITCH_BLACK, _BLACK)  # Material.hs:46-47


def lambertian(tex: Texture) -> Material:
    """Material.hs:51-53."""
    return Material(MAT_LAMBERTIAN, tex)


def lommelSeeliger(tex: Texture) -> Material:
    """Material.hs:56-61."""
    return Material(MAT_LOMMEL_SEELIGER, tex)


def mirror(tex: Texture) -> Material:
    """Material.hs:64-67."""
    return Material(MAT_MIRROR, tex)


def metal(fuzz: float, tex: Texture) -> Material:
    """Material.hs:72-78."""
    return Material(MAT_METAL, tex, float(fuzz))


def dielectric(ior: float) -> Material:
    """Material.hs:89-106."""
    return Material(MAT_DIELECTRIC, _BLACK, float(ior))


def transparent(tex: Texture) -> Material:
    """Material.hs:109-112."""
    return Material(MAT_TRANSPARENT, tex)


def isotropic(tex: Texture) -> Material:
    """Material.hs:116-118."""
    return Material(MAT_ISOTROPIC, tex)


def anisotropic(g: float, tex: Texture) -> Material:
    """Material.hs:124-129."""
    return Material(MAT_ANISOTROPIC, tex, float(g))
