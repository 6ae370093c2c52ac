"""raytrace_b200 — B200-native backend for the `raytrace` path-tracing loop.

The package mirrors the export lists of the reference's six modules
(Graphics.Ray / .Core / .Geometry / .Material / .Texture / .Noise) with the same names and
argument meaning, so a scene program reads like its Haskell original.  Every combinator
builds data (a description tree); `raytrace` hands that tree to librt_b200.so (C ABI,
include/rt_b200.h), which flattens it and renders with hand-written sm_100a CUDA kernels.
There is no CPU fallback: without the built library `raytrace` raises ImportError.
"""
from .core import (V3, Box, Interval, Vec3, allCorners, boxHull, boxJoin, cross, degrees, dot,
                   fromCorners, infinity, longestDim, midpoint, norm, padBox, padInterval,
                   shiftBox)
from .geometry import (M44, Geometry, Mesh, boundingBox, bvhNode, bvhTree, constantMedium,
                       cuboid, group, moving, parallelogram, parseObj, planeShape, pureGeometry,
                       readObj, rotateX, rotateY, rotateZ, scale, sphere, transform,
                       transformVertices, translate, triangle, triangleMesh, withMaterial)
from .material import (Material, anisotropic, dielectric, isotropic, lambertian, lightSource,
                       lommelSeeliger, metal, mirror, pitchBlack, transparent)
from .ray import (CameraSettings, CompiledScene, Ray, RenderContext, StdGen, defaultCameraSettings,
                  imageHeight, mkStdGen, raytrace, readImage, writeImage, writeImageSqrt)
from .texture import (Texture, checkerTexture, constantTexture, imageTexture, marbleTexture,
                      noiseTexture, solidTexture, uvTexture)
from ._ffi import RtError, UserDefinedError

__version__ = "0.1.0"
