"""``igg_reduce_max_sum_static``: the max/sum reduction for callers that may not allocate.

Some users of the library cannot hand it scratch memory and cannot let it allocate any: a monitoring
hook that runs once the application has filled the GPU with its fields, a host whose policy keeps all
device allocation inside its own pool, a time loop that must not take the device-wide stall of
``cudaMalloc``.  For them the scratch is a fixed array inside the library.  This test drives that entry
point the way such a caller would -- through the C ABI with raw pointers -- and checks it against the
ordinary entry point (``igg_reduce_max_sum`` with caller-provided scratch), on two calls in a row because
the built-in array is shared and keeps the previous call's contents.

The file name sorts last on purpose: with ``pytest -x`` every other test has run before this one can stop
the session.  It must pass both on its own in a fresh process (first call = address look-up) and after the
rest of the suite.
"""
import ctypes as C

import pytest
import torch

from rocm_mpi_b200.ops import native

pytestmark = pytest.mark.gpu

MAX_N = 1 << 29   # IGG_REDUCE_STATIC_MAX_N in include/igg_b200.h


def _entry():
    lib = native.require()
    fn = lib.igg_reduce_max_sum_static
    fn.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    return lib, fn


def _close(a: torch.Tensor, b: torch.Tensor) -> bool:
    # same tolerance as test_reduce_max_sum_repeated_calls: max exact, sum to 1e-9 relative
    return float(a[0]) == float(b[0]) and abs(float(a[1]) - float(b[1])) <= 1e-9 * max(1.0, abs(float(b[1])))


def test_static_scratch_matches_caller_scratch_on_consecutive_calls():
    assert torch.cuda.is_available()
    lib, fn = _entry()
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device="cuda").manual_seed(2024)
    # big then small then big: a shorter call must not pick up partials left by a longer one
    sizes = [(1 << 22) + 3, 1000, 1, (1 << 20), (1 << 22) + 3]
    out = torch.empty(2, dtype=torch.float64, device="cuda")
    for k, n in enumerate(sizes):
        x = torch.rand(n, generator=g, dtype=torch.float64, device="cuda") - 0.25 * (k + 1)
        for rep in range(2):                                  # two calls in a row on the same input ...
            out.fill_(float("nan"))
            rc = fn(x.data_ptr(), n, out.data_ptr(), stream)
            assert rc == 0, lib.igg_last_error().decode()
            want = native.reduce_max_sum(x)                    # ... against the existing entry point
            torch.cuda.synchronize()
            assert _close(out.cpu(), want.cpu()), (n, rep, out.cpu(), want.cpu())
        assert float(out[0]) == float(x.max())                 # and against plain torch
        assert abs(float(out[1]) - float(x.sum())) <= 1e-9 * max(1.0, abs(float(x.sum())))


def test_static_scratch_rejects_oversized_input_without_launching():
    lib, fn = _entry()
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    out = torch.full((2,), 5.0, dtype=torch.float64, device="cuda")
    before = native.launch_count()
    rc = fn(x.data_ptr(), MAX_N + 1, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == native.ERR_UNSUPPORTED
    msg = lib.igg_last_error().decode()
    assert "IGG_REDUCE_STATIC_MAX_N" in msg and str(MAX_N) in msg
    assert native.launch_count() == before
    torch.cuda.synchronize()
    assert out.tolist() == [5.0, 5.0]                          # untouched
    # the limit itself is accepted by the size check (exercised with a small array: only n is checked)
    rc = fn(x.data_ptr(), 8, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert out.tolist() == [0.0, 0.0]
