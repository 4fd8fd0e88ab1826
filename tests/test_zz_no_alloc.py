"""mgb_noalloc_dot (csrc/noalloc.cu): the allocation-free dot product through the C ABI against the existing entry point
mgb_vec_dot, on two calls in a row (the static scratch array keeps the first call's contents), in both precisions, on a
second stream, and at the size limit's boundary.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("no GPU")
    return t


@pytest.mark.parametrize("precision", [1, 0])
def test_matches_vec_dot_on_two_calls_in_a_row(mgb, torch, precision):
    s = mgb.MultigridSolver(4, (1, 2, 1), 3)            # 16 x 32 x 16 cells of degree 4: 545 025 DoFs, 134 chunks
    lib = mgb.lib()
    dtype = torch.float64 if precision else torch.float32
    tol = 1e-12 if precision else 1e-5                  # the tolerances of tests/test_gpu_parity.py
    g = torch.Generator(device="cuda").manual_seed(5)
    out = torch.full((2,), -7.0, dtype=torch.float64, device="cuda")
    results = []
    # second call: fewer entries than the first, so stale partial sums of the first call sit behind the ones it reads
    for level in (s.n_levels - 1, s.n_levels - 2):
        n = s.n_dofs(level)
        x = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.3
        y = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.3
        torch.cuda.synchronize()
        ref = C.c_double()
        assert lib.mgb_vec_dot(s.h, level, precision, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), C.byref(ref)) == 0
        stream = torch.cuda.current_stream()
        code = lib.mgb_noalloc_dot(precision, n, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()),
                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream.cuda_stream))
        assert code == 0, lib.mgb_last_error()
        torch.cuda.synchronize()
        got = float(out[0])
        scale = float(x.double().abs() @ y.double().abs())
        assert abs(got - ref.value) <= tol * scale, (level, got, ref.value)
        assert abs(got - float(x.double() @ y.double())) <= tol * scale
        assert float(out[1]) == -7.0                    # writes one entry only
        results.append(got)
    # the same two calls again, back to back without a synchronisation in between: bitwise the same results
    again = torch.zeros(2, dtype=torch.float64, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(5)
    keep = []
    for k, level in enumerate((s.n_levels - 1, s.n_levels - 2)):
        n = s.n_dofs(level)
        x = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.3
        y = torch.rand(n, dtype=dtype, device="cuda", generator=g) - 0.3
        keep += [x, y]
        mgb.noalloc_dot(x, y, again[k:])
    torch.cuda.synchronize()
    assert [float(again[0]), float(again[1])] == results


def test_other_stream_ragged_length_and_limits(mgb, torch):
    lib = mgb.lib()
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    side = torch.cuda.Stream()
    for n in (1, 255, 4096, 4097, 3 * 4096 + 17):
        x = torch.arange(n, dtype=torch.float64, device="cuda") / n
        y = torch.ones(n, dtype=torch.float64, device="cuda")
        side.wait_stream(torch.cuda.current_stream())
        mgb.noalloc_dot(x, y, out, stream=side)
        side.synchronize()
        assert abs(float(out[0]) - (n - 1) / 2.) <= 1e-12 * n
    # n = 0: set-up only, nothing written, no pointers needed
    out.fill_(3.0)
    assert lib.mgb_noalloc_dot(1, 0, None, None, None, None) == 0
    torch.cuda.synchronize()
    assert float(out[0]) == 3.0
    # beyond the maximum: rejected before anything is queued, with a message that names the limit
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    assert lib.mgb_noalloc_dot(1, mgb.NOALLOC_DOT_MAX_N + 1, C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()),
                               C.c_void_p(out.data_ptr()), None) != 0
    assert str(mgb.NOALLOC_DOT_MAX_N) in lib.mgb_last_error().decode()
    assert lib.mgb_noalloc_dot(7, 8, C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), None) != 0
    assert lib.mgb_noalloc_dot(1, 8, None, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), None) != 0
    torch.cuda.synchronize()
    assert float(out[0]) == 3.0


def test_no_allocation_during_the_call(mgb, torch):
    """device memory in use (cudaMemGetInfo) does not move across calls once the first call has done its set-up"""
    x = torch.rand(1 << 20, dtype=torch.float32, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    mgb.noalloc_dot(x, x, out)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(50):
        mgb.noalloc_dot(x, x, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
    assert abs(float(out[0]) - float(x.double() @ x.double())) <= 1e-5 * float(x.double() @ x.double())
