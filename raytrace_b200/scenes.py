"""The reference's scene programs restated against the mirrored API
(test/Main.hs and README.md:41-55; parameters per SURVEY.md Appendix C).

Each function returns ``(settings, world)``; the render seed the reference fixes for that
scene (where it fixes one) is in ``SEEDS``.  Scene-generation randomness (demo1 / demo2)
comes from numpy's PCG64 with a fixed seed: distributionally identical to the reference's
``randomR`` calls, not stream-identical (the `random` package is not in the reference tree).

Assets (earth map, pawn / bunny meshes) are read from tests/golden/assets/*.npz, produced
from the reference's images/ by tests/golden/make_fixtures.py.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple

import numpy as np

from .core import V3, degrees, fromCorners, midpoint
from .geometry import (Geometry, Mesh, boundingBox, bvhTree, constantMedium, cuboid, group, moving,
                       parallelogram, pureGeometry, rotateX, rotateY, scale, sphere, transform,
                       transformVertices, translate, triangleMesh, withMaterial)
from .material import (dielectric, isotropic, lambertian, lightSource, lommelSeeliger, metal,
                       mirror, pitchBlack)
from .ray import CameraSettings, Ray, defaultCameraSettings
from .texture import (checkerTexture, constantTexture, imageTexture, marbleTexture, noiseTexture)

# package data: the reference's own input files (images/pawn.obj, images/bunny.obj, images/earthmap.jpg), read the way the
# reference reads them - readObj / parseObj through the library's C++ loader (rt_obj_parse) and readImage
ASSET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")

SEEDS = {"readme": 100, "cornell": 234, "demo1": 1, "demo2": 1234, "pawn": 55, "bunny": 55,
         "lommel": 55, "sphere_uv": 12}


# ---- backgrounds (test/Main.hs:19-28) -----------------------------------------------------------

def sky(ray: Ray):
    a = 0.5 * (ray.direction[1] + 1)
    return tuple((1 - a) * 1.0 + a * c for c in (0.5, 0.7, 1.0))


def grayFade(ray: Ray):
    t = (ray.direction[1] + 1) * 0.5
    return (t, t, t)


def black(_ray: Ray):
    return (0.0, 0.0, 0.0)


# ---- assets ---------------------------------------------------------------------------------------
_EARTH = None
_MESHES: dict = {}

def _srgb8_to_linear(img_u8: np.ndarray) -> np.ndarray:
    c = img_u8.astype(np.float64) / 255.0
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4).astype(np.float32)


def load_earthmap() -> np.ndarray:
    """images/earthmap.jpg through `readImage` (decode, then sRGB -> linear; Ray.hs:241-245)."""
    global _EARTH
    if _EARTH is None:
        from .ray import readImage
        _EARTH = readImage(os.path.join(ASSET_DIR, "earthmap.jpg"))
    return _EARTH


def load_mesh(name: str) -> Mesh:
    """images/<name>.obj through readObj: the .obj TEXT parsed by the library's loader (rt_obj_parse, Geometry.hs:207-285)."""
    if name not in _MESHES:
        from .geometry import readObj
        _MESHES[name] = readObj(os.path.join(ASSET_DIR, f"{name}.obj"))
    return _MESHES[name]


# ---- config 1: README example (README.md:41-55) --------------------------------------------------

def readme_example(width: int = 600, spp: int = 50) -> Tuple[CameraSettings, Geometry]:
    world = group([
        lambertian(checkerTexture(20, 10, 0.2, 0.8)) << sphere(V3(0, 0, 0), 1),
        lambertian(constantTexture(V3(0, 0.2, 0.5))) << sphere(V3(0, -1000, 0), 999),
        mirror(constantTexture(0.8)) << parallelogram(V3(-3.25, -1, -0.75), V3(1.25, 0, -1.25), V3(0, 2, 0)),
    ])
    settings = defaultCameraSettings.replace(
        cs_center=V3(-0.75, 0, 2), cs_lookAt=V3(0, 0, -1), cs_aspectRatio=16 / 9,
        cs_imageWidth=width, cs_samplesPerPixel=spp)
    return settings, world


# ---- config 2: Cornell box (test/Main.hs:188-218) -------------------------------------------------

def cornell_box(spp: int = 200, depth: int = 50, redirect: bool = True, width: int = 600):
    red = lambertian(constantTexture(V3(0.65, 0.05, 0.05)))
    white = lambertian(constantTexture(V3(0.73, 0.73, 0.73)))
    green = lambertian(constantTexture(V3(0.12, 0.45, 0.15)))
    light = lightSource(constantTexture(V3(15, 15, 15)))
    world = group([
        green << parallelogram(V3(555, 0, 0), V3(0, 555, 0), V3(0, 0, 555)),
        red << parallelogram(V3(0, 0, 0), V3(0, 555, 0), V3(0, 0, 555)),
        light << parallelogram(V3(343, 554, 332), V3(-130, 0, 0), V3(0, 0, -105)),
        white << parallelogram(V3(0, 0, 0), V3(555, 0, 0), V3(0, 0, 555)),
        white << parallelogram(V3(555, 555, 555), V3(-555, 0, 0), V3(0, 0, -555)),
        white << parallelogram(V3(0, 0, 555), V3(555, 0, 0), V3(0, 555, 0)),
        transform(translate(V3(265, 0, 295)) @ rotateY(degrees(15)),
                  white << cuboid(fromCorners(V3(0, 0, 0), V3(165, 330, 165)))),
        transform(translate(V3(130, 0, 65)) @ rotateY(degrees(-18)),
                  white << cuboid(fromCorners(V3(0, 0, 0), V3(165, 165, 165)))),
    ])
    targets = [(0.25, V3(343, 554, 332), V3(-130, 0, 0), V3(0, 0, -105))] if redirect else []
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=1.0, cs_imageWidth=width, cs_samplesPerPixel=spp, cs_maxRecursionDepth=depth,
        cs_background=black, cs_vfov=degrees(40), cs_center=V3(278, 278, -800),
        cs_lookAt=V3(278, 278, 0), cs_redirectTargets=targets)
    return settings, world


# ---- config 3: demo1 (test/Main.hs:136-186) -------------------------------------------------------

def demo1(width: int = 1200, spp: int = 500, depth: int = 50, scene_seed: int = 1):
    rng = np.random.default_rng(scene_seed)
    materialGround = lambertian(constantTexture(V3(0.5, 0.5, 0.5)))
    materialGlass = dielectric(1.5)
    materialDiffuse = lambertian(constantTexture(V3(0.4, 0.2, 0.1)))
    materialMirror = mirror(constantTexture(V3(0.7, 0.6, 0.5)))
    objs = [
        materialGround << sphere(V3(0, -1000, 0), 1000),
        materialGlass << sphere(V3(0, 1, 0), 1),
        materialDiffuse << sphere(V3(-4, 1, 0), 1),
        materialMirror << sphere(V3(4, 1, 0), 1),
    ]
    for a in range(-11, 11):
        for b in range(-11, 11):
            ox, oz = rng.uniform(0, 0.9), rng.uniform(0, 0.9)
            center = (a + ox, 0.2, b + oz)
            if math.dist(center, (4, 0.2, 0)) <= 0.9:
                continue
            choose = rng.random()
            if choose < 0.8:
                color = tuple(rng.random(3) * rng.random(3))
                mat = lambertian(constantTexture(color))
            elif choose < 0.95:
                fuzz = rng.uniform(0, 0.5)
                color = tuple(rng.uniform(0.5, 1.0, 3))
                mat = metal(fuzz, constantTexture(color))
            else:
                mat = materialGlass
            objs.append(mat << sphere(center, 0.2))
    world = bvhTree(objs)
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=16 / 9, cs_imageWidth=width, cs_samplesPerPixel=spp,
        cs_maxRecursionDepth=depth, cs_vfov=degrees(20), cs_center=V3(13, 2, 3),
        cs_lookAt=V3(0, 0, 0), cs_defocusAngle=degrees(0.6), cs_focusDist=10, cs_background=sky)
    return settings, world


# ---- config 4: demo2 (test/Main.hs:259-321) -------------------------------------------------------

def demo2(width: int = 800, spp: int = 250, depth: int = 50, scene_seed: int = 1234,
          earth: Optional[np.ndarray] = None):
    rng = np.random.default_rng(scene_seed)
    ground = lambertian(constantTexture(V3(0.48, 0.83, 0.53)))
    white = lambertian(constantTexture(V3(0.73, 0.73, 0.73)))
    if earth is None:
        earth = load_earthmap()
    box_list = []
    for i in range(20):
        for j in range(20):
            x0, z0 = -1000 + i * 100, -1000 + j * 100
            y1 = rng.uniform(1, 101)
            box_list.append(cuboid(fromCorners(V3(x0, 0, z0), V3(x0 + 100, y1, z0 + 100))))
    boxes = ground << bvhTree(box_list)
    ball_list = [sphere(tuple(rng.uniform(0, 165, 3)), 10) for _ in range(1000)]
    balls = white << transform(translate(V3(-100, 270, 395)) @ rotateY(degrees(15)), bvhTree(ball_list))
    boundary = sphere(V3(360, 150, 145), 70)
    lq, ls0, ls1 = V3(123, 554, 147), V3(300, 0, 0), V3(0, 0, 265)
    large = [
        lightSource(constantTexture(V3(7, 7, 7))) << parallelogram(lq, ls0, ls1),
        lambertian(constantTexture(V3(0.7, 0.3, 0.1))) << moving(V3(0, 0, 0), V3(30, 0, 0), sphere(V3(400, 400, 200), 50)),
        dielectric(1.5) << sphere(V3(260, 150, 45), 50),
        dielectric(1.5) << boundary,
        metal(1.0, constantTexture(V3(0.8, 0.8, 0.9))) << sphere(V3(0, 150, 145), 50),
        lambertian(imageTexture(earth)) << transform(translate(V3(400, 0, 400)) @ rotateY(math.pi / 2),
                                                     sphere(V3(0, 200, 0), 100)),
        lambertian(marbleTexture(V3(0, 0, 0.05), 4, 0)) << sphere(V3(220, 280, 300), 80),
    ]
    world = group([
        pureGeometry(group([boxes, balls] + large)),
        isotropic(constantTexture(1)) << constantMedium(0.0001, sphere(V3(0, 0, 0), 5000)),
        isotropic(constantTexture(V3(0.2, 0.4, 0.9))) << constantMedium(0.2, boundary),
    ])
    settings = defaultCameraSettings.replace(
        cs_center=V3(478, 278, -600), cs_lookAt=V3(278, 278, 0), cs_vfov=degrees(40),
        cs_aspectRatio=1.0, cs_imageWidth=width, cs_samplesPerPixel=spp,
        cs_maxRecursionDepth=depth, cs_background=black,
        cs_redirectTargets=[(0.25, lq, ls0, ls1)])
    return settings, world


# ---- config 5: meshes with a subsurface volume (test/Main.hs:323-344, 362-378) --------------------

def _volume_world(mesh_geom: Geometry) -> Geometry:
    return group([
        pureGeometry(dielectric(1.5) << mesh_geom),
        isotropic(constantTexture(V3(1, 0, 0))) << constantMedium(5, mesh_geom),
    ])


def pawn(width: int = 500, spp: int = 400, depth: int = 20, aspect: float = 1.0,
         mesh: Optional[Mesh] = None):
    """pawnTest; the benchmark form is width=1920, aspect=16/9, spp=1024."""
    mesh = load_mesh("pawn") if mesh is None else mesh
    mesh = transformVertices(scale(100), mesh)
    world = _volume_world(triangleMesh(mesh))
    settings = defaultCameraSettings.replace(
        cs_center=V3(0, 3.75, 5), cs_lookAt=V3(0, 2.75, 0), cs_imageWidth=width,
        cs_aspectRatio=aspect, cs_vfov=degrees(80), cs_samplesPerPixel=spp,
        cs_maxRecursionDepth=depth, cs_background=grayFade)
    return settings, world


def _bunny_mesh(mesh: Optional[Mesh]) -> Mesh:
    mesh = load_mesh("bunny") if mesh is None else mesh
    center = tuple(midpoint(i) for i in boundingBox(triangleMesh(mesh)))
    m = rotateY(degrees(30)) @ scale(12) @ translate(tuple(-c for c in center))
    return transformVertices(m, mesh)


def bunny_test(width: int = 600, spp: int = 100, mesh: Optional[Mesh] = None):
    """bunnyTest (test/Main.hs:362-378): lambertian bunny, default depth 10."""
    world = lambertian(constantTexture(V3(0.3, 0.3, 1))) << triangleMesh(_bunny_mesh(mesh))
    settings = defaultCameraSettings.replace(
        cs_center=V3(0, 0.5, 2), cs_lookAt=V3(0, 0, 0), cs_imageWidth=width,
        cs_samplesPerPixel=spp, cs_background=grayFade)
    return settings, world


def bunny_volume(width: int = 1920, spp: int = 1024, depth: int = 20, aspect: float = 16 / 9,
                 mesh: Optional[Mesh] = None):
    """Config 5's bunny variant: bunnyTest's mesh and camera with pawnTest's world recipe."""
    world = _volume_world(triangleMesh(_bunny_mesh(mesh)))
    settings = defaultCameraSettings.replace(
        cs_center=V3(0, 0.5, 2), cs_lookAt=V3(0, 0, 0), cs_imageWidth=width, cs_aspectRatio=aspect,
        cs_samplesPerPixel=spp, cs_maxRecursionDepth=depth, cs_background=grayFade)
    return settings, world


# ---- feature fixtures (the reference's other scene programs) ---------------------------------------

def metal_test(width: int = 400, spp: int = 100):
    """metalTest (test/Main.hs:30-61)."""
    world = group([
        lambertian(constantTexture(V3(0.8, 0.8, 0.0))) << sphere(V3(0, -100.5, -1), 100),
        lambertian(constantTexture(V3(0.1, 0.2, 0.5))) << sphere(V3(0, 0, -1.2), 0.5),
        dielectric(1.5) << sphere(V3(-1, 0, -1), 0.5),
        dielectric(1 / 1.5) << sphere(V3(-1, 0, -1), 0.4),
        metal(1.0, constantTexture(V3(0.8, 0.6, 0.2))) << sphere(V3(1, 0, -1), 0.5),
    ])
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=16 / 9, cs_imageWidth=width, cs_samplesPerPixel=spp, cs_maxRecursionDepth=50,
        cs_background=sky, cs_center=V3(-2, 2, 1), cs_lookAt=V3(0, 0, -1), cs_vfov=degrees(20),
        cs_defocusAngle=degrees(10), cs_focusDist=3.4)
    return settings, world


def noise_test(width: int = 400, spp: int = 100):
    """noiseTest (test/Main.hs:63-86)."""
    world = group([
        lambertian(noiseTexture(2, 2.0, V3(10, 0, 0), 0, 1)) << sphere(V3(0, -1000, 0), 1000),
        lambertian(marbleTexture(V3(0, 0, 1), 4, 0)) << sphere(V3(0, 2, 0), 2),
    ])
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=16 / 9, cs_imageWidth=width, cs_samplesPerPixel=spp, cs_maxRecursionDepth=50,
        cs_background=sky, cs_vfov=degrees(20), cs_center=V3(13, 2, 3), cs_lookAt=V3(0, 0, 0))
    return settings, world


def quad_test(width: int = 400, spp: int = 100):
    """quadTest (test/Main.hs:88-117)."""
    def lam(c):
        return lambertian(constantTexture(c))
    world = group([
        lam(V3(1.0, 0.2, 0.2)) << parallelogram(V3(-3, -2, 5), V3(0, 0, -4), V3(0, 4, 0)),
        lam(V3(0.2, 1.0, 0.2)) << parallelogram(V3(-2, -2, 0), V3(4, 0, 0), V3(0, 4, 0)),
        lam(V3(0.2, 0.2, 1.0)) << parallelogram(V3(3, -2, 1), V3(0, 0, 4), V3(0, 4, 0)),
        lam(V3(1.0, 0.5, 0.0)) << parallelogram(V3(-2, 3, 1), V3(4, 0, 0), V3(0, 0, 4)),
        lam(V3(0.2, 0.8, 0.8)) << parallelogram(V3(-2, -3, 5), V3(4, 0, 0), V3(0, 0, -4)),
    ])
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=1, cs_imageWidth=width, cs_samplesPerPixel=spp, cs_maxRecursionDepth=50,
        cs_background=sky, cs_vfov=degrees(80), cs_center=V3(0, 0, 9), cs_lookAt=V3(0, 0, 0))
    return settings, world


def cuboid_test(width: int = 300, spp: int = 10, earth: Optional[np.ndarray] = None):
    """cuboidTest (test/Main.hs:119-126)."""
    earth = load_earthmap() if earth is None else earth
    obj = lambertian(imageTexture(earth)) << cuboid(fromCorners(V3(-1, -2, -0.5), V3(1, 2, 0.5)))
    world = transform(translate(V3(0, 0, -3)) @ rotateX(degrees(60)), obj)
    return defaultCameraSettings.replace(cs_imageWidth=width, cs_samplesPerPixel=spp), world


def sphere_uv_test(earth: Optional[np.ndarray] = None):
    """sphereUVTest (test/Main.hs:128-134): the one fully deterministic reference program."""
    earth = load_earthmap() if earth is None else earth
    world = lambertian(imageTexture(earth)) << group([sphere(V3(0, 0, -2), 0.4), sphere(V3(0, 0, -1), 0.4)])
    settings = defaultCameraSettings.replace(cs_imageWidth=1, cs_samplesPerPixel=1, cs_vfov=0.0001)
    return settings, world


def cornell_smoke(width: int = 600, spp: int = 200, depth: int = 50):
    """cornellSmoke (test/Main.hs:220-257)."""
    red = lambertian(constantTexture(V3(0.65, 0.05, 0.05)))
    white = lambertian(constantTexture(V3(0.73, 0.73, 0.73)))
    green = lambertian(constantTexture(V3(0.12, 0.45, 0.15)))
    light = lightSource(constantTexture(V3(7, 7, 7)))
    surfaces = group([
        green << parallelogram(V3(555, 0, 0), V3(0, 555, 0), V3(0, 0, 555)),
        red << parallelogram(V3(0, 0, 0), V3(0, 555, 0), V3(0, 0, 555)),
        light << parallelogram(V3(113, 554, 127), V3(330, 0, 0), V3(0, 0, 305)),
        white << parallelogram(V3(0, 0, 0), V3(555, 0, 0), V3(0, 0, 555)),
        white << parallelogram(V3(555, 555, 555), V3(-555, 0, 0), V3(0, 0, -555)),
        white << parallelogram(V3(0, 0, 555), V3(555, 0, 0), V3(0, 555, 0)),
    ])
    cube1 = transform(translate(V3(265, 0, 295)) @ rotateY(degrees(15)),
                      cuboid(fromCorners(V3(0, 0, 0), V3(165, 330, 165))))
    cube2 = transform(translate(V3(130, 0, 65)) @ rotateY(degrees(-18)),
                      cuboid(fromCorners(V3(0, 0, 0), V3(165, 165, 165))))
    world = group([
        pureGeometry(surfaces),
        pitchBlack << constantMedium(0.01, cube1),
        isotropic(constantTexture(1)) << constantMedium(0.01, cube2),
    ])
    settings = defaultCameraSettings.replace(
        cs_aspectRatio=1.0, cs_imageWidth=width, cs_samplesPerPixel=spp, cs_maxRecursionDepth=depth,
        cs_background=black, cs_vfov=degrees(40), cs_center=V3(278, 278, -800),
        cs_lookAt=V3(278, 278, 0),
        cs_redirectTargets=[(0.25, V3(113, 554, 127), V3(330, 0, 0), V3(0, 0, 305))])
    return settings, world


def lommel_seeliger_test(width: int = 500, spp: int = 500):
    """lommelSeeligerTest (test/Main.hs:346-360)."""
    world = group([
        lommelSeeliger(constantTexture(1)) << sphere(V3(0, 0, -2), 1),
        lightSource(constantTexture(160)) << sphere(V3(0, 0, 22), 1),
    ])
    settings = defaultCameraSettings.replace(
        cs_imageWidth=width, cs_samplesPerPixel=spp, cs_background=black,
        cs_redirectTargets=[(0.5, V3(-1, -1, 21), V3(2, 0, 0), V3(0, 2, 0))])
    return settings, world


# name → (constructor, render seed); the five BASELINE configs first
CONFIGS = {
    "readme": (readme_example, SEEDS["readme"]),
    "cornell": (lambda **kw: cornell_box(redirect=True, **kw), SEEDS["cornell"]),
    "cornell_noredirect": (lambda **kw: cornell_box(redirect=False, **kw), SEEDS["cornell"]),
    "demo1": (demo1, SEEDS["demo1"]),
    "demo2": (demo2, SEEDS["demo2"]),
    "pawn": (pawn, SEEDS["pawn"]),
    "bunny_volume": (bunny_volume, SEEDS["bunny"]),
}
