"""The drop-in boundary is a plain C ABI: a C program (tests/cabi/readme_scene.c: no Python, no torch) builds the README
scene through include/rt_b200.h, links librt_b200.so and renders.  Without a GPU it must fail loudly with
RT_ERR_NO_DEVICE (there is no CPU fallback); on the GPU box its result must match the Python front end's."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _compile(tmp_path_factory, name):
    exe = str(tmp_path_factory.mktemp("cabi") / name)
    libdir = os.path.join(ROOT, "raytrace_b200")
    subprocess.check_call(["gcc", "-std=c11", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cabi", name + ".c"), "-o", exe,
                           "-L", libdir, "-lrt_b200", "-lm", f"-Wl,-rpath,{libdir}"])
    return exe


@pytest.fixture(scope="module")
def client(built, tmp_path_factory):
    return _compile(tmp_path_factory, "readme_scene")


@pytest.fixture(scope="module")
def multi_client(built, tmp_path_factory):
    return _compile(tmp_path_factory, "cornell_multi")


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


def test_multi_gpu_c_client_fails_loudly_without_gpu(multi_client):
    if _gpu_visible():
        pytest.skip("a GPU is visible")
    r = subprocess.run([multi_client, "32", "4"], capture_output=True, text=True)
    assert "committed: 18 prims" in r.stdout
    assert r.returncode == 7 and "rt_context_create" in r.stderr and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_multi_gpu_c_client_matches_single_gpu_render(multi_client):
    """rt_context_create(NULL, 0) + rt_render_multi from plain C, one process, ALL visible GPUs (one on the driver's test
    box; tools/gpu_multi_check.sh runs the same binary on 2 / 8 GPUs): same image as rt_render on device 0 up to fp32
    summation order, same ray count, sum of squares included."""
    r = subprocess.run([multi_client, "120", "40"], capture_output=True, text=True)
    print(r.stdout.strip())
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "pixels beyond tolerance 0" in r.stdout
