"""GPU kernels through the C ABI against the CPU oracle on the same seeded inputs."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    from conjugategradient_b200 import _lib, solver
    L = _lib.lib()
    ctx = L.mgcg_create(0)
    assert ctx, _lib.last_error()
    yield torch, L, ctx, solver
    L.mgcg_destroy(ctx)


def _dev(torch, a):
    return torch.as_tensor(a).to("cuda:0")


def _p(t):
    return C.c_void_p(t.data_ptr())


MATS = ["tridiag", "banded", "p2d", "p3d", "p27", "mps", "ragged"]


def _matrix(gen, name):
    if name == "tridiag":
        v, rp, c, _ = gen.tridiag121(5000)
    elif name == "banded":
        v, rp, c, _, _ = gen.banded_sin(3000, 160)
    elif name == "p2d":
        v, rp, c = gen.poisson2d5(128, 128)
    elif name == "p3d":
        v, rp, c = gen.poisson3d7(40, 33, 17)
    elif name == "p27":
        v, rp, c = gen.poisson3d27_varcoef(20)
    elif name == "mps":
        v, rp, c = gen.mps_particles(14)
    else:  # ragged: empty rows, a 3000-long row, random lengths
        rng = np.random.default_rng(7)
        n = 2000
        cnt = rng.integers(0, 40, n)
        cnt[::17] = 0
        cnt[5] = 3000
        rp = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(cnt, out=rp[1:])
        c = rng.integers(0, n, rp[-1]).astype(np.int32)
        v = rng.standard_normal(rp[-1])
        rp = rp.astype(np.int32)
    return v, rp, c


@pytest.mark.parametrize("name", MATS)
@pytest.mark.parametrize("kernel", [1, 2, 3, 0])       # STREAM (storage-order sums), VECTOR, ROWS (TMA-staged, lanes per row), AUTO
def test_spmv_parity(env, oracle, gen, name, kernel):
    torch, L, ctx, _ = env
    v, rp, c = _matrix(gen, name)
    n = len(rp) - 1
    x = np.random.default_rng(1).standard_normal(n)
    y_ref = oracle.spmv(v, rp, c, x)
    tv, tr, tc, tx = _dev(torch, v), _dev(torch, rp), _dev(torch, c), _dev(torch, x)
    ty = torch.full((n,), np.nan, dtype=torch.float64, device="cuda:0")
    assert L.mgcg_spmv(ctx, n, n, len(v), _p(tv), _p(tr), _p(tc), _p(tx), _p(ty), 1.0, 0.0, kernel) == 0
    torch.cuda.synchronize()
    y = ty.cpu().numpy()
    if kernel == 1:
        # STREAM sums each row in storage order with rounded products: bit-identical to SparseMatrix.Multiply
        assert np.array_equal(y, y_ref)
    else:
        scale = oracle.spmv(np.abs(v), rp, c, np.abs(x)) + 1e-300
        assert np.max(np.abs(y - y_ref) / scale) < 1e-14
    # y = alpha*A*x + beta*y  (CsrMV semantics, Mgcg.cu:10-19)
    y0 = np.random.default_rng(2).standard_normal(n)
    ty = _dev(torch, y0)
    assert L.mgcg_spmv(ctx, n, n, len(v), _p(tv), _p(tr), _p(tc), _p(tx), _p(ty), -2.0, 0.5, kernel) == 0
    torch.cuda.synchronize()
    np.testing.assert_allclose(ty.cpu().numpy(), -2.0 * y_ref + 0.5 * y0, rtol=1e-12, atol=1e-12 * np.abs(y_ref).max())


def test_spmv_rectangular_block(env, oracle, gen):
    # the multi-GPU use: countForDevice x count block with global column indices (Mgcg.cu:138)
    torch, L, ctx, _ = env
    v, rp, c, _, _ = gen.banded_sin(1000, 160)
    off, cnt = 250, 250
    lro, eo, ec, mn, mx = oracle.initialize_block(rp, c, off, cnt)
    x = np.random.default_rng(3).standard_normal(1000)
    y_ref = oracle.spmv(v, rp, c, x)[off:off + cnt]
    tv, tr, tc, tx = _dev(torch, v[eo:eo + ec]), _dev(torch, lro), _dev(torch, c[eo:eo + ec]), _dev(torch, x)
    ty = torch.empty(cnt, dtype=torch.float64, device="cuda:0")
    assert L.mgcg_spmv(ctx, cnt, 1000, ec, _p(tv), _p(tr), _p(tc), _p(tx), _p(ty), 1.0, 0.0, 1) == 0
    torch.cuda.synchronize()
    assert np.array_equal(ty.cpu().numpy(), y_ref)


@pytest.mark.parametrize("n", [1, 2, 31, 257, 4097, 1_000_003])
def test_dot_maxabs_deterministic(env, oracle, n):
    torch, L, ctx, _ = env
    rng = np.random.default_rng(n)
    a, b = rng.standard_normal(n), rng.standard_normal(n)
    ta, tb = _dev(torch, a), _dev(torch, b)
    out = C.c_double(0)
    vals = []
    for _ in range(3):
        assert L.mgcg_dot(ctx, n, _p(ta), _p(tb), None, C.byref(out)) == 0
        vals.append(out.value)
    assert vals[0] == vals[1] == vals[2]                       # fixed-order reduction: run-to-run bit-stable
    ref = oracle.dot(a, b)
    assert abs(vals[0] - ref) <= 1e-13 * np.abs(a * b).sum()
    assert L.mgcg_max_abs(ctx, n, _p(ta), None, C.byref(out)) == 0
    assert out.value == np.abs(a).max()
    # unaligned views (p + offsetForDevice in the split-phase exports)
    if n > 3:
        assert L.mgcg_dot(ctx, n - 1, C.c_void_p(ta.data_ptr() + 8), C.c_void_p(tb.data_ptr() + 8), None, C.byref(out)) == 0
        assert abs(out.value - oracle.dot(a[1:].copy(), b[1:].copy())) <= 1e-13 * np.abs(a * b).sum()


@pytest.mark.parametrize("n", [5, 1000, 65537])
def test_fused_updates(env, oracle, n):
    torch, L, ctx, _ = env
    rng = np.random.default_rng(n)
    x, r, p, Ap = (rng.standard_normal(n) for _ in range(4))
    alpha, beta = 0.37, 1.7
    tx, tr, tp, tA = (_dev(torch, a) for a in (x, r, p, Ap))
    rr = C.c_double(0)
    assert L.mgcg_update_xr(ctx, n, alpha, _p(tp), _p(tA), _p(tx), _p(tr), None, C.byref(rr)) == 0
    x1, r1 = x + alpha * p, r + (-alpha) * Ap
    np.testing.assert_allclose(tx.cpu().numpy(), x1, rtol=1e-15, atol=1e-15)
    np.testing.assert_allclose(tr.cpu().numpy(), r1, rtol=1e-15, atol=1e-15)
    assert abs(rr.value - oracle.dot(r1, r1)) <= 1e-13 * (r1 * r1).sum()
    assert L.mgcg_update_p(ctx, n, beta, _p(tr), _p(tp)) == 0
    torch.cuda.synchronize()
    np.testing.assert_allclose(tp.cpu().numpy(), r1 + beta * p, rtol=1e-15, atol=1e-15)
    # BLAS-1 compat pieces
    assert L.mgcg_axpy(ctx, n, 2.5, _p(tr), _p(tx)) == 0 and L.mgcg_scal(ctx, n, -3.0, _p(tx)) == 0
    torch.cuda.synchronize()
    # the device contracts a + alpha*b into one FMA (as cuBLAS axpy does); numpy rounds the product first: the difference is
    # <= 1 ulp of the larger operand per operation, which cancellation turns into an absolute (not relative) bound
    np.testing.assert_allclose(tx.cpu().numpy(), (x1 + 2.5 * r1) * -3.0, rtol=1e-15, atol=4e-15 * 3.0 * np.abs(x1).max())


def test_spmv_dot_fused(env, oracle, gen):
    torch, L, ctx, _ = env
    for name in ("p3d", "banded"):
        v, rp, c = _matrix(gen, name)
        n = len(rp) - 1
        x = np.random.default_rng(4).standard_normal(n)
        tv, tr, tc, tx = _dev(torch, v), _dev(torch, rp), _dev(torch, c), _dev(torch, x)
        ty = torch.empty(n, dtype=torch.float64, device="cuda:0")
        td = torch.zeros(1, dtype=torch.float64, device="cuda:0")
        outs = []
        for _ in range(2):
            assert L.mgcg_spmv_dot(ctx, n, n, len(v), _p(tv), _p(tr), _p(tc), _p(tx), _p(ty), _p(tx), _p(td), 0) == 0
            torch.cuda.synchronize()
            outs.append(float(td.cpu()[0]))
        y_ref = oracle.spmv(v, rp, c, x)
        assert outs[0] == outs[1]
        assert abs(outs[0] - oracle.dot(x, y_ref)) <= 1e-13 * np.abs(x * y_ref).sum()
        np.testing.assert_allclose(ty.cpu().numpy(), y_ref, rtol=1e-13, atol=1e-13 * np.abs(y_ref).max())


@pytest.mark.parametrize("name", ["banded", "mps", "ragged", "p27"])
def test_spmv_rows_kernel_fused_epilogues(env, oracle, gen, name):
    """The TMA-staged rows kernel (MGCG_SPMV_ROWS) with the fused p.Ap epilogue, on a sub-block view whose row offsets do not start
    at zero and whose arrays end inside the parent's (the sharded path's use), and with every lane count the plan can pick."""
    torch, L, ctx, _ = env
    v, rp, c = _matrix(gen, name)
    n = len(rp) - 1
    x = np.random.default_rng(5).standard_normal(n)
    w = np.random.default_rng(6).standard_normal(n)
    y_ref = oracle.spmv(v, rp, c, x)
    scale = oracle.spmv(np.abs(v), rp, c, np.abs(x)) + 1e-300
    tv, tr, tc, tx, tw = _dev(torch, v), _dev(torch, rp), _dev(torch, c), _dev(torch, x), _dev(torch, w)
    ty = torch.full((n,), np.nan, dtype=torch.float64, device="cuda:0")
    td = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    assert L.mgcg_spmv_dot(ctx, n, n, len(v), _p(tv), _p(tr), _p(tc), _p(tx), _p(ty), _p(tw), _p(td), 3) == 0
    torch.cuda.synchronize()
    y = ty.cpu().numpy()
    assert np.max(np.abs(y - y_ref) / scale) < 1e-14
    assert abs(float(td.item()) - float(np.dot(w, y))) <= 1e-12 * np.abs(w * y).sum()
    # a row sub-block [r0, r1): row offsets keep the parent's numbering, val/col pointers are the parent's
    r0, r1 = n // 3 + 1, (2 * n) // 3
    ty2 = torch.full((r1 - r0,), np.nan, dtype=torch.float64, device="cuda:0")
    assert L.mgcg_spmv(ctx, r1 - r0, n, len(v), _p(tv), C.c_void_p(tr.data_ptr() + 4 * r0), _p(tc), _p(tx), _p(ty2), 1.0, 0.0, 3) == 0
    torch.cuda.synchronize()
    assert np.max(np.abs(ty2.cpu().numpy() - y_ref[r0:r1]) / scale[r0:r1]) < 1e-14
