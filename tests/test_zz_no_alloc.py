"""mgcg_dot_static — the dot product for callers that are not allowed to allocate GPU memory — against mgcg_dot.

Why the entry point exists: mgcg_dot needs a context, and a context reserves its reduction workspace with cudaMalloc.  A host
program that has already handed every byte of the GPU to its own arena, a time-step loop that cannot afford cudaMalloc's
device-wide synchronisation, or a plug-in whose embedding application forbids runtime allocation cannot make that call.
mgcg_dot_static takes no context and keeps its 33 KB workspace in a __device__ array inside libMgcgGpu.so (csrc/noalloc.cu), shared
by every call on the GPU and never cleared between them except for the one counter each call resets itself.

What is checked, through the C ABI only: two calls in a row (the second on different data and a different length, so anything the
first one left in the shared workspace would show) agree with mgcg_dot on the same inputs within the tolerance of
tests/test_gpu_kernels.py::test_dot_maxabs_deterministic (1e-13 * sum|a_i b_i|) and, since both use one summation order, bit for
bit; lengths beyond MGCG_DOT_STATIC_MAX_N are refused with a message; free device memory does not shrink across calls.

This file is named to be collected LAST (pytest -x stops at the first failure; a failure here must not hide other tests).  It keeps
no state of its own and passes alone in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 1 << 28   # MGCG_DOT_STATIC_MAX_N, include/mgcg.h


@pytest.fixture(scope="module")
def env():
    import torch
    from conjugategradient_b200 import _lib
    L = _lib.lib()
    ctx = L.mgcg_create(0)          # only the comparison (mgcg_dot) needs a context; mgcg_dot_static does not
    assert ctx, _lib.last_error()
    yield torch, L, ctx, _lib
    L.mgcg_destroy(ctx)


def _p(t, byte_offset=0):
    return C.c_void_p(t.data_ptr() + byte_offset)


def _both(torch, L, ctx, ta, tb, n, tout, stream=None, byte_offset=0):
    """(mgcg_dot_static, mgcg_dot) on the first n entries of the device vectors, each read back after a synchronise."""
    tout.fill_(float("nan"))
    torch.cuda.synchronize()
    assert L.mgcg_dot_static(0, n, _p(ta, byte_offset), _p(tb, byte_offset), _p(tout), C.c_void_p(stream)) == 0
    torch.cuda.synchronize()
    got = float(tout.item())
    ref = C.c_double(0)
    assert L.mgcg_dot(ctx, n, _p(ta, byte_offset), _p(tb, byte_offset), None, C.byref(ref)) == 0
    return got, ref.value


@pytest.mark.parametrize("n1,n2", [(1_000_003, 4097), (31, 1_000_003), (1, 2), (4_194_304, 257)])
def test_two_calls_in_a_row_match_mgcg_dot(env, n1, n2):
    torch, L, ctx, _ = env
    rng = np.random.default_rng(n1 + 7 * n2)
    tout = torch.empty(1, dtype=torch.float64, device="cuda:0")
    for n in (n1, n2):              # call 2 reuses the workspace call 1 left behind
        a, b = rng.standard_normal(n), rng.standard_normal(n)
        ta, tb = torch.as_tensor(a).to("cuda:0"), torch.as_tensor(b).to("cuda:0")
        got, ref = _both(torch, L, ctx, ta, tb, n, tout)
        tol = 1e-13 * np.abs(a * b).sum()
        assert abs(got - ref) <= tol, (n, got, ref)
        assert abs(got - float(np.dot(a, b))) <= 10 * tol   # and both are the dot product (numpy pairwise sum; looser bound)
        assert got == ref, (n, got, ref)                    # one summation order -> the same bits


def test_unaligned_views_and_caller_stream(env):
    torch, L, ctx, _ = env
    n = 65_537
    rng = np.random.default_rng(3)
    a, b = rng.standard_normal(n), rng.standard_normal(n)
    ta, tb = torch.as_tensor(a).to("cuda:0"), torch.as_tensor(b).to("cuda:0")
    tout = torch.empty(1, dtype=torch.float64, device="cuda:0")
    side = torch.cuda.Stream()
    for off, stream in ((8, None), (0, side.cuda_stream), (8, side.cuda_stream)):   # 8-byte-aligned views take the scalar kernel
        got, ref = _both(torch, L, ctx, ta, tb, n - off // 8, tout, stream=stream, byte_offset=off)
        assert abs(got - ref) <= 1e-13 * np.abs(a * b).sum()
        assert got == ref


def test_empty_and_too_long(env):
    torch, L, ctx, _lib = env
    tout = torch.full((1,), 5.0, dtype=torch.float64, device="cuda:0")
    ta = torch.ones(4, dtype=torch.float64, device="cuda:0")
    assert L.mgcg_dot_static(0, 0, _p(ta), _p(ta), _p(tout), None) == 0      # empty sum == 0.0, as mgcg_dot
    torch.cuda.synchronize()
    assert float(tout.item()) == 0.0
    # beyond the static workspace's limit: refused before anything is launched (the pointers are never dereferenced)
    assert L.mgcg_dot_static(0, MAX_N + 1, _p(ta), _p(ta), _p(tout), None) == _lib.ERR_INVALID
    assert "MGCG_DOT_STATIC_MAX_N" in _lib.last_error()
    assert L.mgcg_dot_static(0, -1, _p(ta), _p(ta), _p(tout), None) == _lib.ERR_INVALID
    torch.cuda.synchronize()
    assert float(tout.item()) == 0.0                                          # untouched by the refused calls


def test_free_device_memory_does_not_shrink(env):
    torch, L, ctx, _ = env
    n = 1_000_003
    ta = torch.rand(n, dtype=torch.float64, device="cuda:0")
    tb = torch.rand(n, dtype=torch.float64, device="cuda:0")
    tout = torch.empty(1, dtype=torch.float64, device="cuda:0")
    for off in (0, 8):               # warm both kernel variants: the driver may load their code on first use
        assert L.mgcg_dot_static(0, n - 1, _p(ta, off), _p(tb, off), _p(tout), None) == 0
    torch.cuda.synchronize()
    free0, _total = torch.cuda.mem_get_info(0)
    for _ in range(4):
        for off in (0, 8):
            assert L.mgcg_dot_static(0, n - 1, _p(ta, off), _p(tb, off), _p(tout), None) == 0
    torch.cuda.synchronize()
    free1, _total = torch.cuda.mem_get_info(0)
    assert free1 >= free0, (free0, free1)
