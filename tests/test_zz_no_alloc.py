"""s35_vec_dot_device_noalloc: the dot product for callers that cannot allocate device memory (programs
running against a fixed, pre-reserved memory budget, callers tied to their own memory pool, solver loops
that must not suffer the device-wide stall of an allocation).  Its reduction scratch is a fixed array inside
the library, so it takes device pointers and never calls cudaMalloc.

The file name makes pytest collect it last: a run that stops at the first failure has then already run every
other test.  The device buffers come from torch here only because the test needs some caller-owned device
memory to hand in; the library itself does not use torch.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import seeded_vector

pytestmark = pytest.mark.gpu

MAX_N = 2097152  # stated limit of the built-in scratch array (include/step35a_b200.h)


def device_dot(pkg, torch, x_host, y_host, stream=None):
    x = torch.from_numpy(np.ascontiguousarray(x_host)).cuda()
    y = torch.from_numpy(np.ascontiguousarray(y_host)).cuda()
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    status = pkg.lib().s35_vec_dot_device_noalloc(x.numel(), C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                                  C.c_void_p(out.data_ptr()),
                                                  C.c_void_p(stream.cuda_stream) if stream is not None else None)
    pkg.check(status)
    torch.cuda.synchronize()
    return float(out.cpu()[0])


def test_matches_the_allocating_entry_point_on_two_calls_in_a_row(pkg):
    torch = pytest.importorskip("torch")
    # two calls back to back, different lengths and data: the second one finds the first one's partial sums
    # and ticket in the shared scratch array
    for n, seed in ((1412040, 7), (183624, 8)):
        x, y = seeded_vector(n, seed), seeded_vector(n, seed + 100)
        expected = pkg.vec_dot(x, y)  # the existing entry point (host vectors, allocates)
        got = device_dot(pkg, torch, x, y)
        # the tolerance test_gpu_kernels.py::test_vector_kernels uses for dot products ...
        assert abs(got - float(np.dot(x, y))) <= 1e-13 * np.sqrt(n) * max(1.0, abs(float(np.dot(x, y))))
        # ... and, the kernel and the launch shape being the same, the two entry points agree to the last bit
        assert got == expected


def test_sizes_beyond_the_scratch_array_are_rejected(pkg):
    torch = pytest.importorskip("torch")
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    status = pkg.lib().s35_vec_dot_device_noalloc(MAX_N + 1, C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()),
                                                  C.c_void_p(out.data_ptr()), None)
    assert status == 1
    message = pkg.lib().s35_last_error().decode()
    assert str(MAX_N) in message and "limit" in message
    # the largest admissible length still works
    x, y = seeded_vector(MAX_N, 1), seeded_vector(MAX_N, 2)
    assert device_dot(pkg, torch, x, y) == pkg.vec_dot(x, y)


def test_empty_vectors_and_a_stream_of_the_callers(pkg):
    torch = pytest.importorskip("torch")
    # n == 0: the result is 0.0 and the vectors are not touched (null pointers are fine)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    pkg.check(pkg.lib().s35_vec_dot_device_noalloc(0, None, None, C.c_void_p(out.data_ptr()), None))
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 0.0
    # a non-default stream owned by the caller, twice in a row on that stream
    stream = torch.cuda.Stream()
    for n, seed in ((50001, 3), (777, 4)):
        x, y = seeded_vector(n, seed), seeded_vector(n, seed + 50)
        assert device_dot(pkg, torch, x, y, stream) == pkg.vec_dot(x, y)
