"""Host-side mirror of Graphics.Ray (reference: src/Graphics/Ray.hs): CameraSettings,
defaultCameraSettings, raytrace, readImage, writeImage, writeImageSqrt.

``raytrace settings world seed`` keeps the reference's meaning (Ray.hs:121) but runs on the
GPU through the C ABI; the result is an H×W×3 float32 array (row 0 = top, linear RGB,
unclamped, divided by cs_samplesPerPixel) instead of a delayed massiv matrix of Doubles.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import math
import random as _pyrandom
from dataclasses import dataclass, field
from typing import Callable, List, NamedTuple, Optional, Sequence, Tuple

import numpy as np

from . import _ffi
from .core import Vec3, as_vec3
from .geometry import Geometry


class Ray(NamedTuple):
    """``Ray origin direction`` (Core.hs:72)."""
    origin: Vec3
    direction: Vec3


@dataclass(frozen=True)
class StdGen:
    """Stand-in for System.Random.StdGen: the render seed.  The GPU uses a Philox counter RNG
    keyed by this value; stream identity with `random`'s SplitMix is neither possible (its
    source is not in the reference) nor needed (parity is statistical)."""
    seed: int


def mkStdGen(n: int) -> StdGen:
    return StdGen(int(n) & 0xFFFFFFFFFFFFFFFF)


def _white(_ray: Ray) -> Vec3:
    return (1.0, 1.0, 1.0)


@dataclass(frozen=True)
class CameraSettings:
    """Ray.hs:40-68, field for field."""
    cs_center: Vec3 = (0.0, 0.0, 0.0)
    cs_lookAt: Vec3 = (0.0, 0.0, -1.0)
    cs_up: Vec3 = (0.0, 1.0, 0.0)
    cs_vfov: float = math.pi / 2
    cs_aspectRatio: float = 1.0
    cs_imageWidth: int = 100
    cs_samplesPerPixel: int = 10
    cs_maxRecursionDepth: int = 10
    cs_background: Callable[[Ray], Vec3] = _white
    cs_defocusAngle: float = 0.0
    cs_focusDist: float = 10.0
    cs_redirectTargets: Sequence[Tuple[float, Vec3, Vec3, Vec3]] = ()

    def replace(self, **kw) -> "CameraSettings":
        """Record update: ``defaultCameraSettings { cs_imageWidth = 600 }``."""
        for k in ("cs_center", "cs_lookAt", "cs_up"):
            if k in kw:
                kw[k] = as_vec3(kw[k])
        if "cs_redirectTargets" in kw:
            kw["cs_redirectTargets"] = tuple(
                (float(p), as_vec3(q), as_vec3(s0), as_vec3(s1)) for p, q, s0, s1 in kw["cs_redirectTargets"])
        return dataclasses.replace(self, **kw)


defaultCameraSettings = CameraSettings()  # Ray.hs:84-98


def imageHeight(settings: CameraSettings) -> int:
    """Ray.hs:123 — Haskell ``round`` is round-half-even (Python's ``round`` is too)."""
    return int(round(settings.cs_imageWidth / settings.cs_aspectRatio))


# ---- background classification ---------------------------------------------------------------

BG_CONSTANT, BG_YGRADIENT = 1, 2


def classify_background(fn: Callable[[Ray], Vec3]):
    """``cs_background :: Ray -> Color`` keeps its type; the host classifies it by probing
    (SURVEY §7.3 item 1): constant, or affine in dir.y (``sky`` and ``grayFade``,
    test/Main.hs:19-28).  Anything else is a user-defined closure → rejected."""
    rng = _pyrandom.Random(12345)

    def unit():
        while True:
            v = (rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1))
            q = v[0] * v[0] + v[1] * v[1] + v[2] * v[2]
            if 1e-4 <= q <= 1:
                s = 1 / math.sqrt(q)
                return (v[0] * s, v[1] * s, v[2] * s)

    def ev(o, d):
        return as_vec3(fn(Ray(o, d)))

    o0 = (0.0, 0.0, 0.0)
    c_dn, c_up = ev(o0, (0.0, -1.0, 0.0)), ev(o0, (0.0, 1.0, 0.0))
    probes = [((rng.uniform(-100, 100), rng.uniform(-100, 100), rng.uniform(-100, 100)), unit())
              for _ in range(64)]
    vals = [ev(o, d) for o, d in probes]
    tol = 1e-9
    if all(max(abs(v[k] - c_up[k]) for k in range(3)) <= tol for v in vals) and \
            max(abs(c_dn[k] - c_up[k]) for k in range(3)) <= tol:
        return BG_CONSTANT, c_up, c_up
    ok = True
    for (o, d), v in zip(probes, vals):
        a = 0.5 * (d[1] + 1.0)
        for k in range(3):
            if abs((1 - a) * c_dn[k] + a * c_up[k] - v[k]) > tol:
                ok = False
    if ok:
        return BG_YGRADIENT, c_dn, c_up
    raise _ffi.UserDefinedError(
        3, "cs_background is neither constant nor affine in the ray direction's y component; "
           "user-defined backgrounds are rejected (no CPU fallback)")


_BG_CACHE: dict = {}


def _classified(fn):
    """classify_background probes the closure ~70 times (0.5 ms of Python): remember the answer per function object."""
    try:
        hit = _BG_CACHE.get(fn)
    except TypeError:          # unhashable callable
        return classify_background(fn)
    if hit is None:
        hit = classify_background(fn)
        if len(_BG_CACHE) < 256:
            _BG_CACHE[fn] = hit
    return hit


# ---- raytrace -----------------------------------------------------------------------------------

class _PinnedBuffer:
    """Owner of one page-locked float32 buffer from rt_host_alloc; numpy arrays built over it keep it alive
    (``__array_interface__``) and it goes back to the library's free list when the last of them is dropped."""

    def __init__(self, n_floats: int):
        self._lib = _ffi.lib()
        p = C.c_void_p()
        _ffi.check(self._lib.rt_host_alloc(max(1, n_floats) * 4, C.byref(p)))
        self._ptr = p.value
        self.__array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (self._ptr, False), "version": 3}

    def __del__(self):
        try:
            if self._ptr:
                self._lib.rt_host_free(C.c_void_p(self._ptr))
                self._ptr = None
        except Exception:      # interpreter shutdown
            pass


def _image_buffer(h: int, w: int, pinned: bool = True) -> np.ndarray:
    """H×W×3 float32 result array over page-locked memory: rt_render's copy engine writes it directly (no staging
    memcpy).  ``pinned=False``: an ordinary pageable numpy array (the library then stages the copy)."""
    if not pinned:
        return np.empty((h, w, 3), dtype=np.float32)
    return np.asarray(_PinnedBuffer(h * w * 3)).reshape(h, w, 3)



# RtOptions applied to every scene compiled without explicit ``options`` (sweep tools set this; the library itself reads
# no environment variables)
DEFAULT_OPTIONS: dict = {}


class CompiledScene:
    """A committed + uploaded scene; reusable across renders (the upload happens once).
    ``options``: RtOptions fields by name (include/rt_b200.h), e.g. ``dict(bvh_reference=1)``."""

    def __init__(self, world: Geometry, device: int = 0, noise_tables=None, upload: bool = True, options: Optional[dict] = None):
        if world.payload != "material":
            raise TypeError("raytrace: world must be a Geometry m Material (attach materials with <$)")
        self.handle = _ffi.SceneHandle()
        if noise_tables is not None:
            self.handle.set_noise_tables(*noise_tables)
        opts = dict(DEFAULT_OPTIONS)
        opts.update(options or {})
        if opts:
            self.handle.set_options(**opts)
        self.handle.commit(world)
        self.device = device
        if upload:
            self.handle.upload(device)

    def info(self):
        return self.handle.info()

    def camera(self, settings: CameraSettings):
        kind, c0, c1 = _classified(settings.cs_background)
        return _ffi.make_camera(settings, kind, c0, c1)

    def render(self, settings: CameraSettings, seed: StdGen, first_sample: int = 0,
               n_samples: Optional[int] = None, want_sumsq: bool = False, want_stats: bool = False, pinned_out: bool = True):
        L = _ffi.lib()
        self.handle.upload(self.device)
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        n = settings.cs_samplesPerPixel if n_samples is None else int(n_samples)
        out = _image_buffer(h, w, pinned_out)
        sq = _image_buffer(h, w, pinned_out) if want_sumsq else None
        stats = _ffi.RtStats()
        _ffi.check(L.rt_render(self.handle.handle, self.device, C.byref(cam), C.c_uint64(seed.seed),
                               C.c_uint64(first_sample), C.c_uint64(n),
                               out.ctypes.data_as(C.POINTER(C.c_float)),
                               sq.ctypes.data_as(C.POINTER(C.c_float)) if sq is not None else None,
                               C.byref(stats)))
        res = [out]
        if want_sumsq:
            res.append(sq)
        if want_stats:
            res.append(stats.as_dict())
        return res[0] if len(res) == 1 else tuple(res)

    def render_device(self, settings: CameraSettings, seed: StdGen, first_sample: int, n_samples: int,
                      d_sum_ptr: int, d_sumsq_ptr: int = 0, stream_ptr: int = 0, want_stats: bool = True):
        """Accumulate un-normalised sums into caller-owned DEVICE buffers (torch tensors).  Completion is ordered on
        the stream.  With ``want_stats=False`` nothing is read back, and a render whose launch sequence is known up
        front (flat scene, sample range within the pool: one GPU's share of a small render) returns without a host
        synchronisation, so the caller's next stream operation (the reduce) queues behind it with no idle gap."""
        L = _ffi.lib()
        cam = self.camera(settings)
        stats = _ffi.RtStats() if want_stats else None
        _ffi.check(L.rt_render_device(self.handle.handle, self.device, C.byref(cam), C.c_uint64(seed.seed),
                                      C.c_uint64(first_sample), C.c_uint64(n_samples),
                                      C.c_void_p(d_sum_ptr), C.c_void_p(d_sumsq_ptr or None),
                                      C.c_void_p(stream_ptr or None), C.byref(stats) if want_stats else None))
        return stats.as_dict() if want_stats else None

    def probe_primary(self, settings: CameraSettings, time: float = 0.5):
        L = _ffi.lib()
        self.handle.upload(self.device)
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        ids = np.empty((h, w), dtype=np.uint32)
        ts = np.empty((h, w), dtype=np.float32)
        _ffi.check(L.rt_probe_primary(self.handle.handle, self.device, C.byref(cam), float(time),
                                      ids.ctypes.data_as(C.POINTER(C.c_uint32)),
                                      ts.ctypes.data_as(C.POINTER(C.c_float))))
        return ids, ts

    def probe_surface(self, settings: CameraSettings, time: float = 0.5):
        """(H, W, 10) float32: normal(3), uv(2), front, texture colour(3), material kind at the pixel-centre primary hit."""
        L = _ffi.lib()
        self.handle.upload(self.device)
        cam = self.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        out = np.empty((h, w, 10), dtype=np.float32)
        _ffi.check(L.rt_probe_surface(self.handle.handle, self.device, C.byref(cam), float(time),
                                      out.ctypes.data_as(C.POINTER(C.c_float))))
        return out


class RenderContext:
    """A set of GPUs driven by ONE process (rt_context_create): ``render`` is rt_render_multi - samples per pixel split
    across the devices, reduce-scatter of the image over NVLink inside the library, every device copying its slice of
    the result to the host.  ``devices=None``: every visible sm_100 device."""

    def __init__(self, devices: Optional[Sequence[int]] = None):
        self._L = _ffi.lib()
        self._h = C.c_void_p()
        if devices:
            arr = (C.c_int32 * len(devices))(*[int(d) for d in devices])
            _ffi.check(self._L.rt_context_create(arr, len(devices), C.byref(self._h)))
        else:
            _ffi.check(self._L.rt_context_create(None, 0, C.byref(self._h)))

    def close(self):
        if self._h:
            self._L.rt_context_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        i = _ffi.RtContextInfo()
        _ffi.check(self._L.rt_context_info(self._h, C.byref(i)))
        return i.as_dict()

    def render(self, scene: "CompiledScene", settings: CameraSettings, seed: StdGen, want_sumsq: bool = False,
               want_stats: bool = False, pinned_out: bool = True):
        cam = scene.camera(settings)
        h, w = imageHeight(settings), settings.cs_imageWidth
        out = _image_buffer(h, w, pinned_out)
        sq = _image_buffer(h, w, pinned_out) if want_sumsq else None
        stats = _ffi.RtStats() if want_stats else None
        _ffi.check(self._L.rt_render_multi(self._h, scene.handle.handle, C.byref(cam), C.c_uint64(seed.seed),
                                           out.ctypes.data_as(C.POINTER(C.c_float)),
                                           sq.ctypes.data_as(C.POINTER(C.c_float)) if sq is not None else None,
                                           C.byref(stats) if want_stats else None))
        res = [out]
        if want_sumsq:
            res.append(sq)
        if want_stats:
            res.append(stats.as_dict())
        return res[0] if len(res) == 1 else tuple(res)


_ALL_GPUS: Optional[RenderContext] = None


def raytrace(settings: CameraSettings, world: Geometry, seed: StdGen, device: Optional[int] = None) -> np.ndarray:
    """``raytrace :: CameraSettings -> Geometry m Material -> StdGen -> Matrix D Color`` (Ray.hs:121).
    Like the reference - one call, the whole machine - the render uses EVERY visible GPU (rt_render_multi; one GPU when
    only one is visible); ``device=k`` pins it to GPU k (rt_render)."""
    if device is not None:
        return CompiledScene(world, device=device).render(settings, seed)
    global _ALL_GPUS
    if _ALL_GPUS is None:
        _ALL_GPUS = RenderContext()
    return _ALL_GPUS.render(CompiledScene(world, upload=False), settings, seed)


# ---- image IO (Ray.hs:241-260) ---------------------------------------------------------------------

def _srgb_to_linear(c: np.ndarray) -> np.ndarray:
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def readImage(path: str) -> np.ndarray:
    """Ray.hs:241-245: decode, then sRGB → linear RGB; H×W×3 float32, row 0 = top."""
    from PIL import Image
    img = np.asarray(Image.open(path).convert("RGB"), dtype=np.float64) / 255.0
    return _srgb_to_linear(img).astype(np.float32)


def _tonemap(m: np.ndarray, mode: int, device: int = 0) -> np.ndarray:
    L = _ffi.lib()
    a = np.ascontiguousarray(m, dtype=np.float32)
    out = np.empty(a.shape, dtype=np.uint8)
    _ffi.check(L.rt_tonemap_u8(device, a.ctypes.data_as(C.POINTER(C.c_float)), a.size // 3, mode,
                               out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def writeImage(path: str, m: np.ndarray, device: int = 0) -> None:
    """Ray.hs:248-252: linear RGB → sRGB OETF → 8-bit (device kernel), then PNG/JPEG by extension."""
    from PIL import Image
    Image.fromarray(_tonemap(m, 0, device), "RGB").save(path)


def writeImageSqrt(path: str, m: np.ndarray, device: int = 0) -> None:
    """Ray.hs:256-260: sqrt 'gamma' instead of the sRGB OETF."""
    from PIL import Image
    Image.fromarray(_tonemap(m, 1, device), "RGB").save(path)
