"""The drop-in boundary is a plain C ABI: a C program (tests/cabi/readme_scene.c: no Python, no torch) builds the README
scene through include/rt_b200.h, links librt_b200.so and renders.  Without a GPU it must fail loudly with
RT_ERR_NO_DEVICE (there is no CPU fallback); on the GPU box its result must match the Python front end's."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def client(built, tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("cabi") / "readme_scene")
    libdir = os.path.join(ROOT, "raytrace_b200")
    subprocess.check_call(["gcc", "-std=c11", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cabi", "readme_scene.c"), "-o", exe,
                           "-L", libdir, "-lrt_b200", "-lm", f"-Wl,-rpath,{libdir}"])
    return exe


def _gpu_visible():
    import ctypes
    from raytrace_b200 import _ffi
    n = ctypes.c_int32(0)
    _ffi.lib().rt_device_count(ctypes.byref(n))
    return n.value > 0


def test_c_client_links_and_fails_loudly_without_gpu(client):
    if _gpu_visible():
        pytest.skip("a GPU is visible")
    r = subprocess.run([client, "32", "4"], capture_output=True, text=True)
    assert "committed: 3 prims" in r.stdout              # host half (validate, flatten) works without a GPU
    assert r.returncode == 7 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_c_client_matches_python_front_end(client):
    import raytrace_b200 as rt
    from raytrace_b200 import scenes
    r = subprocess.run([client, "160", "64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    mean = float(re.search(r"mean ([0-9.]+)", r.stdout).group(1))
    s, w = scenes.readme_example(width=160, spp=64)
    img = rt.raytrace(s, w, rt.mkStdGen(100))
    # same scene, same seed, same sample set: equal up to fp32 summation order
    assert abs(mean - float(np.mean(img, dtype=np.float64))) < 2e-5 * mean, (mean, img.mean())
