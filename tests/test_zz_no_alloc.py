"""The allocation-free dot product, called straight through the C ABI.

Some users of this library cannot allocate device memory at the point where they need a dot
product: their step runs inside a latency-critical loop that must not hit the device-wide
synchronisation an allocation brings, or a host framework owns every byte of device memory and
hands them fixed buffers only.  ``cgb_op_dot_static`` serves them with scratch that is built into
the library.  This test checks it against the ordinary entry point ``cgb_op_dot`` (which takes a
scratch buffer from the caller) on two consecutive calls — the second call sees the scratch the
first one left behind — and checks that inputs beyond the built-in capacity are refused.

The file sorts last on purpose: if it fails, a run that stops at the first failure has already
run everything else.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from cuda_mpi_parallel_b200 import _native
    return _native, _native.lib()


@pytest.mark.parametrize("n", [1, 1000, 32768, 32769, 5_000_017])
def test_static_dot_matches_scratch_dot_twice(n):
    import torch
    N, L = _lib()
    torch.cuda.set_device(0)
    g = torch.Generator(device="cuda").manual_seed(n)
    out_ref = torch.zeros(1, dtype=torch.float64, device="cuda")
    out_new = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    scratch = torch.zeros(L.cgb_op_scratch_bytes(), dtype=torch.uint8, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call in range(2):           # second call reuses whatever the first left in the built-in scratch
        a = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
        b = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
        N.check(L.cgb_op_dot(a.data_ptr(), b.data_ptr(), n, out_ref.data_ptr(), scratch.data_ptr(), stream))
        N.check(L.cgb_op_dot_static(a.data_ptr(), b.data_ptr(), n, out_new.data_ptr(), stream))
        torch.cuda.synchronize()
        want, got = out_ref.item(), out_new.item()
        bound = 1e-12 * (a.abs() * b.abs()).sum().item()       # tolerance used for reductions in these tests
        assert abs(got - want) <= bound, (call, n, got, want)
        exact = (a.cpu().numpy() * b.cpu().numpy()).sum()
        assert abs(got - exact) <= bound
        out_new.fill_(float("nan"))


def test_static_dot_is_repeatable_and_handles_empty_input():
    import torch
    N, L = _lib()
    torch.cuda.set_device(0)
    a = torch.randn(777_777, dtype=torch.float64, device="cuda")
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(L.cgb_op_dot_static(a.data_ptr(), a.data_ptr(), a.numel(), out.data_ptr(), stream))
    N.check(L.cgb_op_dot_static(a.data_ptr(), a.data_ptr(), a.numel(), out.data_ptr() + 8, stream))
    torch.cuda.synchronize()
    assert out[0].item() == out[1].item()
    N.check(L.cgb_op_dot_static(a.data_ptr(), a.data_ptr(), 0, out.data_ptr(), stream))
    torch.cuda.synchronize()
    assert out[0].item() == 0.0


def test_static_dot_rejects_sizes_beyond_its_scratch():
    import torch
    N, L = _lib()
    torch.cuda.set_device(0)
    limit = L.cgb_op_dot_static_max_n()
    assert limit == 2 ** 28
    a = torch.ones(8, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    rc = L.cgb_op_dot_static(a.data_ptr(), a.data_ptr(), limit + 1, out.data_ptr(), None)
    assert rc == -1
    msg = L.cgb_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert out.item() == 0.0        # nothing was launched
