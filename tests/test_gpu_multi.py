"""-m gpu tests of the in-library multi-GPU path (rt_context_create / rt_render_multi, include/rt_b200.h): one process,
one host thread + stream per share, reduce-scatter of the image inside the library, every share copying its slice to the
host.  A context may list a device more than once, so the G > 1 logic (sample split, slice layout with padding, peer-load
reduce, per-slice read-back) runs on a one-GPU box too; the NVSwitch multicast variant needs distinct GPUs
(tools/gpu_multi_check.py, profiles/r02*_multi_*.txt)."""
import ctypes as C

import numpy as np
import pytest

import raytrace_b200 as rt
from raytrace_b200 import _ffi, scenes
from raytrace_b200.ray import CompiledScene, RenderContext

pytestmark = pytest.mark.gpu


def _n_gpus():
    n = C.c_int32(0)
    _ffi.lib().rt_device_count(C.byref(n))
    return n.value


CASES = [
    ("cornell", dict(width=50, spp=30)),       # flat scene, single-shot shares; 7500 floats: not a multiple of 4 G for G = 3
    ("demo2", dict(width=41, spp=10, depth=12)),   # BVH + two media + moving sphere + image / Perlin textures; odd size
    ("pawn", dict(width=35, spp=7)),           # spp not divisible by G; primary-ray culling; subsurface medium
]


@pytest.mark.parametrize("name,kw", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("shares", [1, 2, 3, 8])
def test_render_multi_equals_single_device_render(built, name, kw, shares):
    ctor, seed = scenes.CONFIGS[name]
    s, w = ctor(**kw)
    cs = CompiledScene(w)
    want, want_sq, st1 = cs.render(s, rt.mkStdGen(seed), want_sumsq=True, want_stats=True)
    ctx = RenderContext([0] * shares)
    img, sq, stn = ctx.render(cs, s, rt.mkStdGen(seed), want_sumsq=True, want_stats=True)
    info = ctx.info()
    assert info["n_devices"] == shares and info["reduce_mode"] == ("single" if shares == 1 else "peer"), info
    assert stn["samples"] == st1["samples"] and stn["rays"] == st1["rays"]       # the same path segments, however they are split
    assert np.isfinite(img).all()
    assert np.allclose(img, want, rtol=2e-4, atol=2e-5)
    assert np.allclose(sq, want_sq, rtol=1e-3, atol=1e-4)
    # pageable destination (staged copy) and a second render on the same context (buffers reused)
    again = ctx.render(cs, s, rt.mkStdGen(seed), pinned_out=False)
    assert np.allclose(again, want, rtol=2e-4, atol=2e-5)
    # a larger image on the same context re-allocates its buffers
    s2 = s.replace(cs_imageWidth=s.cs_imageWidth + 13, cs_samplesPerPixel=4)
    assert np.allclose(ctx.render(cs, s2, rt.mkStdGen(3)), cs.render(s2, rt.mkStdGen(3)), rtol=2e-4, atol=2e-5)
    ctx.close()


def test_raytrace_uses_every_visible_gpu(built):
    """`raytrace` is one call that uses the whole machine (Ray.hs:121,238): with no device argument it goes through
    rt_render_multi on every visible sm_100 GPU (NVSwitch multicast reduce-scatter when there are several)."""
    s, w = scenes.cornell_box(spp=24, width=64)
    img = rt.raytrace(s, w, rt.mkStdGen(234))
    one = rt.raytrace(s, w, rt.mkStdGen(234), device=0)
    assert np.allclose(img, one, rtol=2e-4, atol=2e-5)
    ctx = RenderContext()
    info = ctx.info()
    assert info["n_devices"] == _n_gpus()
    ctx.render(CompiledScene(w, upload=False), s, rt.mkStdGen(1))
    info = ctx.info()
    print("all-GPU context:", info)
    if info["n_devices"] > 1:
        assert info["reduce_mode"] in ("multicast", "peer"), info      # NVLink systems never need the gather fallback
    ctx.close()


def test_context_argument_errors(built):
    L = _ffi.lib()
    h = C.c_void_p()
    bad = (C.c_int32 * 1)(99)
    assert L.rt_context_create(bad, 1, C.byref(h)) == 1 and not h.value          # device index out of range
    assert L.rt_context_create(None, 1, C.byref(h)) == 1
    assert L.rt_context_create(None, 0, None) == 1
    assert L.rt_context_destroy(None) == 0
    assert L.rt_render_multi(None, None, None, 0, None, None, None) == 1
