"""Multigrid-preconditioned CG on the GPU against its CPU twin (oracle/mg_ref.c).
The reference has no multigrid so this parity is "unpinned": the twin is the build's own restatement
of the same algorithm; it is cross-checked against the plain-CG oracle and direct solves in tests/test_oracle.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _setup(S, torch, v, rp, c, **kw):
    m = S.Mgcg(0)
    tv, tr, tc = (torch.as_tensor(a).cuda() for a in (v, rp, c))
    m.setup(tv, tr, tc, S.default_options(precond=1, **kw))
    return m


@pytest.mark.parametrize("smoothed", [0, 1])
@pytest.mark.parametrize("dims", [(16, 16, 16), (20, 13, 9), (64, 48, 1)])
def test_hierarchy_bit_exact(oracle, gen, smoothed, dims):
    """Aggregation maps are integers (bit-exact); level operators are sums/products in a fixed order (bit-exact too)."""
    import torch
    from conjugategradient_b200 import solver as S
    nx, ny, nz = dims
    v, rp, c = gen.poisson3d7(nx, ny, nz) if nz > 1 else gen.poisson2d5(nx, ny)
    m = _setup(S, torch, v, rp, c, nx=nx, ny=ny, nz=nz, mg_smoothed=smoothed, mg_coarse_size=64)
    H = oracle.MgRef(v, rp, c, nx, ny, nz, coarse_size=64, smoothed=smoothed)
    assert m.levels() == H.levels() >= 2
    for l in range(m.levels()):
        gv, gr, gc = m.level(l)
        hv, hr, hc = H.level(l)
        assert np.array_equal(gr, hr) and np.array_equal(gc, hc), f"level {l} structure"
        assert np.array_equal(gv, hv), f"level {l} values: max diff {np.abs(gv - hv).max()}"
        if l + 1 < m.levels():
            assert np.array_equal(m.aggregates(l), H.aggregates(l))
    m.close(); H.close()


def test_hierarchy_bit_exact_dense_row_spgemm(oracle, gen, monkeypatch):
    """The block-per-row SpGEMM of the coarsest levels (few, very long rows) gives the same bits as the twin: forced on here
    for every product with <= 2048 rows (at full size it serves products above 16384 terms per row)."""
    import torch
    from conjugategradient_b200 import solver as S
    monkeypatch.setenv("MGCG_SPGEMM_DENSE_MIN", "0")
    nx, ny, nz = 24, 20, 16
    v, rp, c = gen.poisson3d7(nx, ny, nz)
    m = _setup(S, torch, v, rp, c, nx=nx, ny=ny, nz=nz, mg_coarse_size=20)
    H = oracle.MgRef(v, rp, c, nx, ny, nz, coarse_size=20)
    assert m.levels() == H.levels() >= 3
    for l in range(m.levels()):
        gv, gr, gc = m.level(l)
        hv, hr, hc = H.level(l)
        assert np.array_equal(gr, hr) and np.array_equal(gc, hc) and np.array_equal(gv, hv), f"level {l}"
    m.close(); H.close()


def test_hierarchy_no_hint_blocks(oracle, gen):
    import torch
    from conjugategradient_b200 import solver as S
    v, rp, c = gen.mps_particles(12)
    m = _setup(S, torch, v, rp, c, mg_coarse_size=64, mg_aggregation=1)      # MGCG_AGG_BLOCKS: 8 consecutive rows
    H = oracle.MgRef(v, rp, c, coarse_size=64, aggregation=1)
    assert m.levels() == H.levels() >= 2
    assert np.array_equal(m.aggregates(0), np.arange(12 ** 3) // 8)
    for l in range(m.levels()):
        gv, gr, gc = m.level(l)
        hv, hr, hc = H.level(l)
        assert np.array_equal(gr, hr) and np.array_equal(gc, hc) and np.array_equal(gv, hv)
    m.close(); H.close()


@pytest.mark.parametrize("tail", [0, 100000, 2000])     # coarse tail: separate launches (default) / whole cycle in one persistent kernel / from level 1
@pytest.mark.parametrize("smoothed,pre,post", [(0, 1, 1), (1, 1, 1), (1, 2, 2), (1, 3, 0), (1, 2, 1)])
def test_vcycle_matches_twin(oracle, gen, smoothed, pre, post, tail):
    import torch
    from conjugategradient_b200 import solver as S
    n = 20
    v, rp, c = gen.poisson3d7(n)
    m = _setup(S, torch, v, rp, c, nx=n, ny=n, nz=n, mg_smoothed=smoothed, mg_pre_sweeps=pre, mg_post_sweeps=post, mg_coarse_size=100,
               mg_tail_rows=tail)
    H = oracle.MgRef(v, rp, c, n, n, n, coarse_size=100, smoothed=smoothed)
    r = gen.rhs_uniform(n ** 3, 9)
    tr_, tz = torch.as_tensor(r).cuda(), torch.empty(n ** 3, dtype=torch.float64, device="cuda")
    m.mg_apply(tr_, tz)
    z = tz.cpu().numpy()
    zt = H.apply(r, pre, post, 0.85)
    assert np.abs(z - zt).max() <= 1e-13 * np.abs(zt).max()
    m.close(); H.close()


CASES = [("p3d32", 1), ("p3d32", 0), ("p3d_odd", 1), ("p2d128", 1), ("p27", 1), ("mps", 1)]


@pytest.mark.parametrize("name,smoothed", CASES)
def test_mgcg_matches_twin(oracle, gen, name, smoothed):
    import torch
    from conjugategradient_b200 import solver as S
    if name == "p3d32":
        dims = (32, 32, 32); v, rp, c = gen.poisson3d7(32)
    elif name == "p3d_odd":
        dims = (37, 21, 18); v, rp, c = gen.poisson3d7(*dims)
    elif name == "p2d128":
        dims = (128, 128, 1); v, rp, c = gen.poisson2d5(128, 128)
    elif name == "p27":
        dims = (24, 24, 24); v, rp, c = gen.poisson3d27_varcoef(24)
    else:
        dims = (0, 0, 0); v, rp, c = gen.mps_particles(16)
    n = len(rp) - 1
    b = gen.rhs_uniform(n, 2)
    m = _setup(S, torch, v, rp, c, nx=dims[0], ny=dims[1], nz=dims[2], mg_smoothed=smoothed, norm=2, tolerance=1e-8, max_iteration=1000)
    tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
    res = m.solve(tb, tx)
    H = oracle.MgRef(v, rp, c, *dims, smoothed=smoothed)
    xt, itt, rest, st = H.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=1000)
    assert st == 0 and res.levels == H.levels()
    assert abs(res.iteration - itt) <= 1, (res.iteration, itt)
    x = tx.cpu().numpy()
    if res.iteration == itt:
        assert np.abs(x - xt).max() / np.abs(xt).max() <= 1e-10
    # and it solves the system: true relative residual below the tolerance
    import scipy.sparse as sp
    A = sp.csr_matrix((v, c, rp))
    assert np.linalg.norm(b - A @ x) / np.linalg.norm(b) < 2e-8
    # against the unpreconditioned oracle's answer (both at 1e-8 relative residual)
    xo, ito, _, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, norm=oracle.NORM_REL_L2)
    assert res.iteration < ito
    assert np.abs(x - xo).max() / np.abs(xo).max() < 1e-5
    m.close(); H.close()


def test_mgcg_graph_vs_eager_bitwise(gen):
    import torch
    from conjugategradient_b200 import solver as S
    n = 24
    v, rp, c = gen.poisson3d7(n)
    b = gen.rhs_uniform(n ** 3, 4)
    out = []
    for graph in (0, 1, 2):
        m = _setup(S, torch, v, rp, c, nx=n, ny=n, nz=n, norm=2, use_graph=graph, check_interval=4)
        tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n ** 3, dtype=torch.float64, device="cuda")
        res = m.solve(tb, tx)
        res2 = m.solve(tb, tx.zero_())          # second solve reuses the hierarchy (and the graph)
        assert res2.iteration == res.iteration
        out.append((res.iteration, tx.cpu().numpy()))
        m.close()
    assert out[0][0] == out[1][0] == out[2][0] and np.array_equal(out[0][1], out[1][1]) and np.array_equal(out[0][1], out[2][1])


def test_compat_solve_with_preconditioner_switch(oracle, gen):
    """The reference-named `Solve` export can run MGCG via MgcgCompatSetPreconditioner (extension; default off)."""
    from conjugategradient_b200 import _lib, solver as S
    n = 24
    v, rp, c = gen.poisson3d7(n)
    b = gen.rhs_uniform(n ** 3, 0)
    L = _lib.lib()
    cg = S.ConjugateGradientSingleGpu(n ** 3, 7, 0, n ** 3, 1e-8)
    cg.A = S.SparseMatrix(v, rp, c); cg.b[:] = b
    cg.Initialize(); cg.Solve(); cg.Read()
    plain_it, x_plain = cg.Iteration, cg.x.copy()
    try:
        L.MgcgCompatSetPreconditioner(1, n, n, n)
        cg.x[:] = 0
        cg.Initialize(); cg.Solve(); cg.Read()
    finally:
        L.MgcgCompatSetPreconditioner(0, 0, 0, 0)
    cg.close()
    assert cg.Iteration < plain_it / 3
    assert np.abs(cg.x - x_plain).max() / np.abs(x_plain).max() < 1e-6


def _random_spd(n, per_row, seed):
    """Irregular symmetric M-matrix (random graph, random weights over three decades), diagonal first, columns ascending."""
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    i = rng.integers(0, n, n * per_row // 2); j = rng.integers(0, n, n * per_row // 2)
    keep = i != j
    i, j = i[keep], j[keep]
    w = 10.0 ** rng.uniform(-2, 1, len(i))
    W = sp.coo_matrix((w, (i, j)), shape=(n, n)).tocsr()
    W = (W + W.T).tocsr(); W.sum_duplicates(); W.sort_indices()
    d = np.asarray(W.sum(axis=1)).ravel() + 0.01
    cnt = np.diff(W.indptr) + 1
    rp = np.zeros(n + 1, dtype=np.int64); np.cumsum(cnt, out=rp[1:])
    val = np.empty(rp[-1]); col = np.empty(rp[-1], dtype=np.int32)
    val[rp[:-1]] = d; col[rp[:-1]] = np.arange(n)
    off = np.ones(rp[-1], dtype=bool); off[rp[:-1]] = False
    val[off] = -W.data; col[off] = W.indices
    return val, rp.astype(np.int32), col


MIS2_CASES = [("mps12_r21", 0.5), ("mps12_r21", 0.25), ("mps10_r31", 0.5), ("p3d_forced", 0.5), ("p27_forced", 0.25), ("random", 0.5)]


@pytest.mark.parametrize("name,theta", MIS2_CASES)
def test_hierarchy_mis2_bit_exact(oracle, gen, name, theta):
    """Strength filter + MIS-2 aggregation on the device: aggregate maps (integers) and every level operator equal the CPU
    twin's bit for bit — irregular particle matrices (38 and 124 entries per row), grids with the hint ignored, a random graph."""
    import torch
    from conjugategradient_b200 import solver as S
    dims, agg = (0, 0, 0), 0
    if name == "mps12_r21":
        v, rp, c = gen.mps_particles(12, re_over_l0=2.1)
    elif name == "mps10_r31":
        v, rp, c = gen.mps_particles(10, re_over_l0=3.1)
    elif name == "p3d_forced":
        dims, agg = (14, 11, 9), 2; v, rp, c = gen.poisson3d7(*dims)
    elif name == "p27_forced":
        dims, agg = (12, 12, 12), 2; v, rp, c = gen.poisson3d27_varcoef(12)
    else:
        v, rp, c = _random_spd(3000, 12, 5)
    m = _setup(S, torch, v, rp, c, nx=dims[0], ny=dims[1], nz=dims[2], mg_coarse_size=40, mg_aggregation=agg, mg_strength_theta=theta)
    H = oracle.MgRef(v, rp, c, *dims, coarse_size=40, aggregation=agg, theta=theta)
    assert m.levels() == H.levels() >= 2
    for l in range(m.levels()):
        if l + 1 < m.levels():
            ga, ha = m.aggregates(l), H.aggregates(l)
            assert np.array_equal(ga, ha), f"level {l} aggregates differ at {np.flatnonzero(ga != ha)[:5]}"
            sizes = np.bincount(ga)
            assert sizes.min() >= 1 and ga.max() + 1 == len(sizes)
        gv, gr, gc = m.level(l)
        hv, hr, hc = H.level(l)
        assert np.array_equal(gr, hr) and np.array_equal(gc, hc), f"level {l} structure"
        assert np.array_equal(gv, hv), f"level {l} values: max diff {np.abs(gv - hv).max()}"
    m.close(); H.close()


@pytest.mark.parametrize("re,theta", [(2.1, 0.5), (3.1, 0.5)])
def test_mgcg_mis2_particles(oracle, gen, re, theta):
    """MGCG with algebraic aggregation on a particle matrix: iterations within +-1 of the twin, solution to 1e-10, far fewer
    iterations than plain CG, operator complexity bounded."""
    import torch
    from conjugategradient_b200 import solver as S
    v, rp, c = gen.mps_particles(20, re_over_l0=re)
    n = len(rp) - 1
    b = gen.rhs_uniform(n, 2)
    m = _setup(S, torch, v, rp, c, norm=2, tolerance=1e-8, max_iteration=500, mg_strength_theta=theta)
    tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
    res = m.solve(tb, tx)
    H = oracle.MgRef(v, rp, c, theta=theta)
    xt, itt, _, st = H.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=500)
    assert st == 0 and abs(res.iteration - itt) <= 1, (res.iteration, itt)
    x = tx.cpu().numpy()
    if res.iteration == itt:
        assert np.abs(x - xt).max() / np.abs(xt).max() <= 1e-10
    xo, ito, _, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, norm=oracle.NORM_REL_L2)
    assert res.iteration * 3 < ito
    assert res.operator_complexity < 1.3
    m.close(); H.close()


def test_level_size_overflow_is_refused():
    """the entry count of a level operator is summed in 64 bits; a product past 2^31 - 1 entries is refused instead
    of wrapping the int32 scan (2^20 rows x 4096 = 2^32 entries; nothing of that size is allocated)."""
    from conjugategradient_b200 import _lib
    import ctypes as C
    L = _lib.lib()
    ctx = L.mgcg_create(0)
    nnz = C.c_longlong(0)
    assert L.mgcg_test_scan_counts(ctx, 1 << 20, 1000, C.byref(nnz)) == _lib.OK and nnz.value == 1000 << 20
    assert L.mgcg_test_scan_counts(ctx, 1 << 20, 2047, C.byref(nnz)) == _lib.OK and nnz.value == 2047 << 20
    assert L.mgcg_test_scan_counts(ctx, 1 << 20, 2048, C.byref(nnz)) == _lib.ERR_UNSUPPORTED
    assert L.mgcg_test_scan_counts(ctx, 1 << 20, 4096, C.byref(nnz)) == _lib.ERR_UNSUPPORTED      # wraps to 0 in 32 bits
    assert b"int32" in L.mgcg_last_error()
    L.mgcg_destroy(ctx)


@pytest.mark.parametrize("smoothed,pre,post", [(1, 1, 1), (0, 1, 1), (1, 2, 2), (0, 2, 1)])
def test_vcycle_rbgs_matches_twin(oracle, gen, smoothed, pre, post):
    """MGCG_SMOOTHER_RBGS: symmetric red-black Gauss-Seidel (colour-masked in-place half sweeps of the STREAM kernels) on the levels
    whose operator is two-colourable by cell parity — level 0 of a 7-point grid; every level under plain aggregation — weighted
    Jacobi on the others.  One V-cycle equals the CPU twin's, and M is symmetric (pre = red, black; post = black, red)."""
    import torch
    from conjugategradient_b200 import solver as S
    n = 20
    v, rp, c = gen.poisson3d7(n)
    m = _setup(S, torch, v, rp, c, nx=n, ny=n, nz=n, mg_smoothed=smoothed, mg_pre_sweeps=pre, mg_post_sweeps=post, mg_coarse_size=100,
               mg_smoother=1)
    H = oracle.MgRef(v, rp, c, n, n, n, coarse_size=100, smoothed=smoothed)
    flags = H.set_smoother(1)
    assert flags[0] and (smoothed == 1 or all(flags[:-1]))
    rng = np.random.default_rng(9)
    u, w = rng.standard_normal(n ** 3), rng.standard_normal(n ** 3)
    tz = torch.empty(n ** 3, dtype=torch.float64, device="cuda")
    m.mg_apply(torch.as_tensor(u).cuda(), tz); zu = tz.cpu().numpy()
    m.mg_apply(torch.as_tensor(w).cuda(), tz); zw = tz.cpu().numpy()
    zt = H.apply(u, pre, post, 0.85)
    assert np.abs(zu - zt).max() <= 1e-13 * np.abs(zt).max()
    if pre == post:
        assert abs(zu @ w - u @ zw) <= 1e-12 * abs(zu @ w)          # M = M^T
    m.close(); H.close()


@pytest.mark.parametrize("name", ["p3d32", "p3d_odd", "p2d128"])
def test_mgcg_rbgs_matches_twin(oracle, gen, name):
    import torch
    from conjugategradient_b200 import solver as S
    if name == "p3d32":
        dims = (32, 32, 32); v, rp, c = gen.poisson3d7(32)
    elif name == "p3d_odd":
        dims = (37, 21, 18); v, rp, c = gen.poisson3d7(*dims)
    else:
        dims = (128, 128, 1); v, rp, c = gen.poisson2d5(128, 128)
    n = len(rp) - 1
    b = gen.rhs_uniform(n, 2)
    out = []
    for graph in (2, 0):
        m = _setup(S, torch, v, rp, c, nx=dims[0], ny=dims[1], nz=dims[2], norm=2, tolerance=1e-8, max_iteration=1000, mg_smoother=1, use_graph=graph)
        tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
        res = m.solve(tb, tx)
        out.append((res.iteration, tx.cpu().numpy()))
        m.close()
    assert out[0][0] == out[1][0] and np.array_equal(out[0][1], out[1][1])       # graph replay == eager launches, bit for bit
    H = oracle.MgRef(v, rp, c, *dims)
    assert H.set_smoother(1)[0]
    xt, itt, _, st = H.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=1000)
    Hj = oracle.MgRef(v, rp, c, *dims)
    _, itj, _, _ = Hj.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=1000)
    it, x = out[0]
    assert st == 0 and abs(it - itt) <= 1, (it, itt)
    assert itt <= itj                                     # never worse than weighted Jacobi in iterations
    if it == itt:
        assert np.abs(x - xt).max() / np.abs(xt).max() <= 1e-10
    import scipy.sparse as sp
    A = sp.csr_matrix((v, c, rp))
    assert np.linalg.norm(b - A @ x) / np.linalg.norm(b) < 2e-8
    H.close(); Hj.close()


@pytest.mark.parametrize("degree,pre,post", [(1, 1, 1), (2, 1, 1), (3, 1, 1), (2, 2, 2), (2, 2, 1)])
def test_vcycle_chebyshev_matches_twin(oracle, gen, degree, pre, post):
    """MGCG_SMOOTHER_CHEBYSHEV (opt-in): degree-k Chebyshev polynomial in D^-1 A on [lmax/ratio, lmax] with the Gershgorin
    lmax — one V-cycle equals the CPU twin's, the eigenvalue bounds are bit-identical, and M stays symmetric for pre == post."""
    import torch
    from conjugategradient_b200 import solver as S
    nx, ny, nz = 20, 17, 12
    v, rp, c = gen.poisson3d27_varcoef(nx, ny, nz)[:3]
    n = nx * ny * nz
    m = _setup(S, torch, v, rp, c, nx=nx, ny=ny, nz=nz, mg_pre_sweeps=pre, mg_post_sweeps=post, mg_coarse_size=100,
               mg_smoother=2, mg_cheby_degree=degree, mg_cheby_ratio=5.0)
    H = oracle.MgRef(v, rp, c, nx, ny, nz, coarse_size=100)
    lmax = H.set_chebyshev(degree, 5.0)
    assert m.levels() == H.levels() >= 3 and lmax[0] > 1.0
    rng = np.random.default_rng(9)
    u, w = rng.standard_normal(n), rng.standard_normal(n)
    tz = torch.empty(n, dtype=torch.float64, device="cuda")
    m.mg_apply(torch.as_tensor(u).cuda(), tz); zu = tz.cpu().numpy()
    m.mg_apply(torch.as_tensor(w).cuda(), tz); zw = tz.cpu().numpy()
    zt = H.apply(u, pre, post, 0.85)
    assert np.abs(zu - zt).max() <= 1e-12 * np.abs(zt).max()
    if pre == post:
        assert abs(zu @ w - u @ zw) <= 1e-11 * abs(zu @ w)          # M = M^T
    m.close(); H.close()


@pytest.mark.parametrize("name", ["p3d32", "v27_24"])
def test_mgcg_chebyshev_matches_twin(oracle, gen, name):
    """PCG with the Chebyshev-smoothed V-cycle: same iteration count (+-1) and solution as the twin, graph replay == eager launches
    bit for bit, and fewer iterations than weighted Jacobi (what the mode is for)."""
    import torch
    from conjugategradient_b200 import solver as S
    if name == "p3d32":
        dims = (32, 32, 32); v, rp, c = gen.poisson3d7(32)
    else:
        dims = (24, 24, 24); v, rp, c = gen.poisson3d27_varcoef(24)[:3]
    n = len(rp) - 1
    b = gen.rhs_uniform(n, 2)
    out = []
    for graph in (2, 0):
        m = _setup(S, torch, v, rp, c, nx=dims[0], ny=dims[1], nz=dims[2], norm=2, tolerance=1e-8, max_iteration=1000, mg_smoother=2, use_graph=graph)
        tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
        res = m.solve(tb, tx)
        out.append((res.iteration, tx.cpu().numpy()))
        m.close()
    assert out[0][0] == out[1][0] and np.array_equal(out[0][1], out[1][1])
    H = oracle.MgRef(v, rp, c, *dims)
    H.set_chebyshev(2, 6.0)                                # the library defaults
    xt, itt, _, st = H.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=1000)
    Hj = oracle.MgRef(v, rp, c, *dims)
    _, itj, _, _ = Hj.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=1000)
    it, x = out[0]
    assert st == 0 and abs(it - itt) <= 1, (it, itt)
    assert itt < itj
    if it == itt:
        assert np.abs(x - xt).max() / np.abs(xt).max() <= 1e-10
    import scipy.sparse as sp
    A = sp.csr_matrix((v, c, rp))
    assert np.linalg.norm(b - A @ x) / np.linalg.norm(b) < 2e-8
    H.close(); Hj.close()


@pytest.mark.parametrize("name", ["p3d32", "v27_24"])
def test_mgcg_single_reduce_matches_classic(oracle, gen, name):
    """MGCG with the single-reduce recurrence: same iteration count (+-1) as the classic loop and the CPU twin, same solution to
    1e-9, while-graph == eager bit for bit."""
    import torch
    from conjugategradient_b200 import solver as S, _lib
    if name == "p3d32":
        dims = (32, 32, 32); v, rp, c = gen.poisson3d7(32)
    else:
        dims = (24, 24, 24); v, rp, c = gen.poisson3d27_varcoef(24)[:3]
    n = len(rp) - 1
    b = gen.rhs_uniform(n, 2)
    out = {}
    for variant, graph in ((0, 2), (1, 2), (1, 0)):
        m = _setup(S, torch, v, rp, c, nx=dims[0], ny=dims[1], nz=dims[2], norm=2, tolerance=1e-10, max_iteration=1000, cg_variant=variant, use_graph=graph)
        tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
        res = m.solve(tb, tx)
        assert res.status == 0
        out[(variant, graph)] = (res.iteration, tx.cpu().numpy())
        m.close()
    assert out[(1, 2)][0] == out[(1, 0)][0] and np.array_equal(out[(1, 2)][1], out[(1, 0)][1])
    assert abs(out[(1, 2)][0] - out[(0, 2)][0]) <= 1
    assert np.abs(out[(1, 2)][1] - out[(0, 2)][1]).max() / np.abs(out[(0, 2)][1]).max() <= 1e-9


@pytest.mark.parametrize("dims,coarse", [((20, 17, 12), 300), ((24, 24, 24), 600), ((33, 33, 33), 1400)])
def test_blocked_dense_inverse_bitwise(oracle, gen, monkeypatch, dims, coarse):
    """The coarsest level's dense inverse by BLOCKED Gauss-Jordan (16 pivots per pair of launches) is bit-identical to the
    one-launch-per-pivot elimination it replaces (and hence to the CPU twin's): same V-cycle output bit for bit, for coarsest levels
    that are / are not multiples of the block (90, 216, 1331 rows)."""
    import torch
    from conjugategradient_b200 import solver as S
    nx, ny, nz = dims
    n = nx * ny * nz
    v, rp, c = gen.poisson3d7(nx, ny, nz)
    r = gen.rhs_uniform(n, 4)
    out = []
    for unblocked in (False, True):
        if unblocked:
            monkeypatch.setenv("MGCG_GJ_UNBLOCKED", "1")
        else:
            monkeypatch.delenv("MGCG_GJ_UNBLOCKED", raising=False)
        m = _setup(S, torch, v, rp, c, nx=nx, ny=ny, nz=nz, mg_coarse_size=coarse)
        sizes = [m.level(l)[1].shape[0] - 1 for l in range(m.levels())]
        tz = torch.empty(n, dtype=torch.float64, device="cuda")
        m.mg_apply(torch.as_tensor(r).cuda(), tz)
        out.append((sizes, tz.cpu().numpy()))
        m.close()
    assert out[0][0] == out[1][0] and out[0][0][-1] <= coarse
    assert np.array_equal(out[0][1], out[1][1])
    if n <= 20000:
        H = oracle.MgRef(v, rp, c, nx, ny, nz, coarse_size=coarse)
        zt = H.apply(r, 1, 1, 0.85)
        assert np.abs(out[0][1] - zt).max() <= 1e-13 * np.abs(zt).max()
        H.close()
