"""rt_tonemap_u8_static (include/rt_b200.h): the allocation-free variant of rt_tonemap_u8, whose staging memory is a file-scope
`static __device__` array inside the library (for callers that cannot call cudaMalloc; it has no counterpart in the
reference).  Called through the C ABI and checked against the existing, allocating entry point - exactly: both run the same
arithmetic, and the 8-bit tone-map output is compared bit for bit everywhere else in this suite (test_image_output_step).

This file is named so that pytest collects it LAST.  It passes alone in a fresh process
(the first call resolves the array's device address) and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from raytrace_b200 import _ffi

MAX_PIXELS = 2097152       # RT_TONEMAP_STATIC_MAX_PIXELS


def _n_gpus():
    n = C.c_int32(0)
    _ffi.lib().rt_device_count(C.byref(n))
    return n.value


def _image(h, w, seed):
    rng = np.random.default_rng(seed)
    m = rng.random((h, w, 3), dtype=np.float32) ** 3 * 1.3 - 0.05      # below 0 and above 1 occur: the clamp is exercised
    m[0, 0] = (np.nan, np.inf, -np.inf)
    if w > 1:
        m[0, 1] = (0.74, 15.0, 0.0031308)
    return np.ascontiguousarray(m)


def _existing(m, mode):
    out = np.empty(m.shape, np.uint8)
    _ffi.check(_ffi.lib().rt_tonemap_u8(0, m.ctypes.data_as(C.POINTER(C.c_float)), m.size // 3, mode,
                                        out.ctypes.data_as(C.POINTER(C.c_uint8))))
    return out


def _static(m, mode, stream=None):
    out = np.full(m.shape, 0xAB, np.uint8)
    _ffi.check(_ffi.lib().rt_tonemap_u8_static(0, C.c_void_p(m.ctypes.data), m.size // 3, mode, C.c_void_p(out.ctypes.data),
                                               C.c_void_p(stream)))
    return out


@pytest.mark.gpu
def test_static_scratch_tonemap_matches_the_allocating_entry_point(built):
    """Two calls in a row share the one static array, which keeps the first call's contents: the second (smaller, other
    mode, other data) must not see any of it.  Then a full 1080p frame, sizes that are not multiples of anything, and
    the limit."""
    big, small = _image(270, 480, 1), _image(37, 53, 2)
    first = _static(big, 0)
    second = _static(small, 1)                       # right behind the first call, no other work in between
    assert np.array_equal(first, _existing(big, 0))
    assert np.array_equal(second, _existing(small, 1))
    assert tuple(first[0, 1]) == (224, 255, 10) and tuple(first[0, 0]) == (0, 255, 0)      # OETF pin, clamp, NaN -> 0
    again = _static(big, 0)                          # same input after another call used the array: same answer
    assert np.array_equal(again, first)
    frame = _image(1080, 1920, 3)                    # BASELINE's largest image fits the array
    assert np.array_equal(_static(frame, 0), _existing(frame, 0))
    one = _image(1, 1, 4)
    assert np.array_equal(_static(one, 1), _existing(one, 1))
    L = _ffi.lib()
    # device buffers are accepted as well (cudaMemcpyDefault), on a caller's stream
    import torch
    dev = torch.device("cuda", 0)
    side = torch.cuda.Stream(dev)
    d_in = torch.from_numpy(small).to(dev)
    d_out = torch.zeros(small.shape, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(dev)
    _ffi.check(L.rt_tonemap_u8_static(0, C.c_void_p(d_in.data_ptr()), small.size // 3, 1, C.c_void_p(d_out.data_ptr()),
                                      C.c_void_p(side.cuda_stream)))
    assert np.array_equal(d_out.cpu().numpy(), second)
    assert torch.cuda.current_device() == 0          # the caller's current device is left alone
    # sizes beyond the array are refused with a clear message, and nothing is written
    too_big = np.zeros(4, np.float32)
    guard = np.full(4, 7, np.uint8)
    st = L.rt_tonemap_u8_static(0, C.c_void_p(too_big.ctypes.data), MAX_PIXELS + 1, 0, C.c_void_p(guard.ctypes.data), None)
    assert st == 1 and b"exceed the static scratch array" in L.rt_last_error() and str(MAX_PIXELS).encode() in L.rt_last_error()
    assert (guard == 7).all()
    assert L.rt_tonemap_u8_static(0, None, 4, 0, C.c_void_p(guard.ctypes.data), None) == 1
    assert L.rt_tonemap_u8_static(99, C.c_void_p(too_big.ctypes.data), 1, 0, C.c_void_p(guard.ctypes.data), None) == 1
    empty = np.zeros((0, 3), np.float32)
    assert L.rt_tonemap_u8_static(0, C.c_void_p(too_big.ctypes.data), 0, 0, C.c_void_p(guard.ctypes.data), None) == 0 and empty.size == 0


def test_static_scratch_tonemap_needs_a_gpu(built):
    """Without a CUDA device the entry point fails loudly (RT_ERR_NO_DEVICE), like every other compute call; the size
    limit is checked before anything touches the driver."""
    L = _ffi.lib()
    a, o = np.zeros(3, np.float32), np.zeros(3, np.uint8)
    assert L.rt_tonemap_u8_static(0, C.c_void_p(a.ctypes.data), MAX_PIXELS + 1, 0, C.c_void_p(o.ctypes.data), None) == 1
    if _n_gpus() > 0:
        pytest.skip("a GPU is visible")
    assert L.rt_tonemap_u8_static(0, C.c_void_p(a.ctypes.data), 1, 0, C.c_void_p(o.ctypes.data), None) == 7
    assert b"no CPU fallback" in L.rt_last_error()
