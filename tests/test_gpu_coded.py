"""Coded columns of the fine operator (csrc/spmv_code.cu): the codes against the oracle's restatement, the SpMV on them bit for
bit against the oracle's storage-order sums and against the int32-column kernel, and whole solves with the layout on and off."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    from conjugategradient_b200 import _lib, generators, solver
    from oracle import oracle
    return torch, _lib, generators, solver, oracle


def _dev(torch, a):
    return torch.as_tensor(a).to("cuda:0")


def _setup(env, v, rp, c, **opt):
    torch, _lib, gen, S, O = env
    m = S.Mgcg(0)
    opt.setdefault("spmv_coding", 0)
    m.setup(_dev(torch, v), _dev(torch, rp), _dev(torch, c), S.default_options(**opt))
    return m


def _few_distances(n, per_row, ndist, seed=3):
    """n rows of `per_row` entries (the last rows: the diagonal only) whose distances col - row take exactly `ndist` values"""
    rng = np.random.default_rng(seed)
    dist = np.arange(ndist, dtype=np.int64)
    body = n - ndist - 1
    cnt = np.full(n, per_row, dtype=np.int64)
    cnt[body:] = 1
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    col = np.empty(rp[-1], dtype=np.int64)
    for j in range(per_row):
        col[rp[:body] + j] = np.arange(body) + dist[(per_row * np.arange(body) + j) % ndist]
    col[rp[body:-1]] = np.arange(body, n)
    v = rng.standard_normal(rp[-1])
    return v, rp.astype(np.int32), col.astype(np.int32)


def test_codes_match_oracle_7pt(env):
    torch, _lib, gen, S, O = env
    v, rp, c = gen.poisson3d7(40, 33, 32)              # 42 240 rows: boundary rows of 4..6 entries, interior rows of 7
    m = _setup(env, v, rp, c)
    got = m.matrix_codes()
    assert got is not None
    want = O.coded_columns(rp, c)
    assert np.array_equal(got[0], want[0]) and list(want[0]) == [-1320, -40, -1, 0, 1, 40, 1320]
    assert np.array_equal(got[1], want[1])
    row_of = np.repeat(np.arange(len(rp) - 1), np.diff(rp))
    assert np.array_equal(row_of + got[0][got[1]], c)                      # the layout restates the columns exactly
    m.close()
    m = _setup(env, v, rp, c, spmv_coding=1)                               # MGCG_CODING_OFF: int32 columns
    assert m.matrix_codes() is None
    m.close()


@pytest.mark.parametrize("case", ["p3d7", "p3d7_values", "dist256"])
def test_coded_spmv_bitwise(env, case):
    torch, _lib, gen, S, O = env
    rng = np.random.default_rng(11)
    if case == "dist256":
        v, rp, c = _few_distances(50000, 5, 256)
    else:
        v, rp, c = gen.poisson3d7(48, 40, 37)
        if case == "p3d7_values":
            v = rng.standard_normal(len(v))
    n = len(rp) - 1
    x = rng.standard_normal(n)
    m = _setup(env, v, rp, c)
    assert m.matrix_codes() is not None
    tx, ty = _dev(torch, x), _dev(torch, np.zeros(n))
    m.spmv(tx, ty)
    y = ty.cpu().numpy()
    assert np.array_equal(y, O.spmv(v, rp, c, x))                          # storage-order sums, bit for bit
    plain = _setup(env, v, rp, c, spmv_coding=1)
    assert plain.matrix_codes() is None
    ty2 = _dev(torch, np.zeros(n))
    plain.spmv(tx, ty2)
    assert np.array_equal(y, ty2.cpu().numpy())
    # alpha / beta epilogue: the coded kernel against the int32-column one
    y0 = rng.standard_normal(n)
    ty3, ty4 = _dev(torch, y0.copy()), _dev(torch, y0.copy())
    m.spmv(tx, ty3, alpha=-2.0, beta=0.5)
    plain.spmv(tx, ty4, alpha=-2.0, beta=0.5)
    assert np.array_equal(ty3.cpu().numpy(), ty4.cpu().numpy())
    assert np.allclose(ty3.cpu().numpy(), -2.0 * y + 0.5 * y0, rtol=1e-14, atol=1e-14)
    m.close(); plain.close()


def test_too_many_distances_stay_csr(env):
    torch, _lib, gen, S, O = env
    v, rp, c = _few_distances(50000, 5, 257)
    assert O.coded_columns(rp, c) is None
    m = _setup(env, v, rp, c)
    assert m.matrix_codes() is None
    n = len(rp) - 1
    x = np.random.default_rng(5).standard_normal(n)
    tx, ty = _dev(torch, x), _dev(torch, np.zeros(n))
    m.spmv(tx, ty)
    assert np.array_equal(ty.cpu().numpy(), O.spmv(v, rp, c, x))
    m.close()


def test_small_and_irregular_matrices_stay_csr(env):
    torch, _lib, gen, S, O = env
    v, rp, c = gen.poisson3d7(20, 20, 20)              # 8 000 rows: below the size where the layout could pay
    m = _setup(env, v, rp, c)
    assert m.matrix_codes() is None
    m.close()
    v, rp, c = gen.mps_particles(34)                   # ~39 k particles: thousands of distinct distances, ROWS kernel
    m = _setup(env, v, rp, c)
    assert m.matrix_codes() is None
    m.close()


@pytest.mark.parametrize("precond", [0, 1])
def test_solve_identical_with_and_without_codes(env, precond):
    torch, _lib, gen, S, O = env
    nx = 40
    v, rp, c = gen.poisson3d7(nx, nx, nx)
    n = nx ** 3
    b = np.random.default_rng(2).standard_normal(n)
    out = []
    for coding in (0, 1):                                # MGCG_CODING_AUTO, MGCG_CODING_OFF
        m = _setup(env, v, rp, c, precond=precond, nx=nx, ny=nx, nz=nx, tolerance=1e-10, norm=_lib.NORM_REL_L2, spmv_coding=coding)
        assert (m.matrix_codes() is not None) == (coding == 0)
        tx = _dev(torch, np.zeros(n))
        res = m.solve(_dev(torch, b), tx)
        assert res.reserved[5] == (4 if coding == 0 else 1)      # TMA stream kernel on coded / on int32 columns
        out.append((res.iteration, tx.cpu().numpy()))
        m.close()
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1], out[1][1])        # same entries, same order, same operands: the same bits


def test_resetup_with_other_matrix(env):
    """the code store is kept across set-ups: a second, larger matrix and then an uncodable one"""
    torch, _lib, gen, S, O = env
    m = S.Mgcg(0)
    rng = np.random.default_rng(9)
    for dims in [(40, 33, 32), (50, 41, 33)]:
        v, rp, c = gen.poisson3d7(*dims)
        m.setup(_dev(torch, v), _dev(torch, rp), _dev(torch, c), S.default_options())
        want = O.coded_columns(rp, c)
        got = m.matrix_codes()
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        x = rng.standard_normal(len(rp) - 1)
        tx, ty = _dev(torch, x), _dev(torch, np.zeros(len(x)))
        m.spmv(tx, ty)
        assert np.array_equal(ty.cpu().numpy(), O.spmv(v, rp, c, x))
    v, rp, c = _few_distances(50000, 5, 300)
    m.setup(_dev(torch, v), _dev(torch, rp), _dev(torch, c), S.default_options())
    assert m.matrix_codes() is None
    m.close()


def test_27pt_operator_runs_on_the_stencil_kernel(env):
    """27 distances in 9 runs of 3: tiles of 64 rows, the x runs staged by TMA; storage-order sums like the oracle's"""
    torch, _lib, gen, S, O = env
    n = 34
    v, rp, c = gen.poisson3d27_varcoef(n)              # 39 304 rows, up to 27 per row, coefficients over six decades
    rows = len(rp) - 1
    m = _setup(env, v, rp, c, precond=1, nx=n, ny=n, nz=n, tolerance=1e-9, norm=_lib.NORM_REL_L2)
    got, want = m.matrix_codes(), O.coded_columns(rp, c)
    assert got is not None and len(want[0]) == 27
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    rng = np.random.default_rng(4)
    x = rng.standard_normal(rows)
    tx, ty = _dev(torch, x), _dev(torch, np.zeros(rows))
    m.spmv(tx, ty)
    assert np.array_equal(ty.cpu().numpy(), O.spmv(v, rp, c, x))
    b = rng.standard_normal(rows)
    sol = _dev(torch, np.zeros(rows))
    res = m.solve(_dev(torch, b), sol)
    assert res.reserved[5] == 4
    plain = _setup(env, v, rp, c, precond=1, nx=n, ny=n, nz=n, tolerance=1e-9, norm=_lib.NORM_REL_L2, spmv_coding=1)
    sol2 = _dev(torch, np.zeros(rows))
    res2 = plain.solve(_dev(torch, b), sol2)
    assert res2.reserved[5] == 3                       # ROWS kernel on int32 columns
    assert abs(res.iteration - res2.iteration) <= 1
    err = float((sol - sol2).abs().max() / sol2.abs().max())
    assert err < 1e-6, err                             # two solves stopped at 1e-9 of an ill-conditioned system
    m.close(); plain.close()


@pytest.mark.parametrize("kind", ["p3d7", "p27"])
def test_stencil_kernel_on_an_x_that_is_only_8_byte_aligned(env, kind):
    """a shard's shifted x view with an odd ghost width is 8-byte but not 16-byte aligned: the runs then start at odd indices"""
    torch, _lib, gen, S, O = env
    v, rp, c = gen.poisson3d7(40, 33, 32) if kind == "p3d7" else gen.poisson3d27_varcoef(34)
    n = len(rp) - 1
    m = _setup(env, v, rp, c)
    assert m.matrix_codes() is not None
    x = np.random.default_rng(8).standard_normal(n)
    big = _dev(torch, np.concatenate([[0.0], x, [0.0, 0.0, 0.0]]))
    tx = big[1:n + 1]
    assert tx.data_ptr() % 16 == 8
    ty = _dev(torch, np.zeros(n))
    m.spmv(tx, ty)
    assert np.array_equal(ty.cpu().numpy(), O.spmv(v, rp, c, x))
    m.close()
