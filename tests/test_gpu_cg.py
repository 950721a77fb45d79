"""CG parity on the GPU against the CPU oracle (ConjugateGradientCpu.cs restatement), through the reference's own
entry points: the MgcgMain flow (construct -> Initialize -> Solve -> Read -> compare, MgcgMain.cs:47-167).
Bars (BASELINE.json north_star): fp64 relative solution error <= 1e-10, iteration count within +-1, index maps bit-exact."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_ERR = 1e-10   # north_star tolerance for floating point
IT_TOL = 1


def _relerr(x, xref):
    return np.abs(x - xref).max() / np.abs(xref).max()


def _case(gen, name):
    if name == "tridiag10":
        v, rp, c, b = gen.tridiag121(10)
        return v, rp, c, b, np.zeros(10), 0
    if name == "banded2000":
        v, rp, c, b, x0 = gen.banded_sin(2000, 160)
        return v, rp, c, b, x0, 0
    if name == "banded_min200":
        v, rp, c, b, x0 = gen.banded_sin(5000, 160)
        return v, rp, c, b, x0, 200                      # MIN_ITERATION = 200 (MgcgMain.cs:25)
    if name == "p2d_ones":
        v, rp, c = gen.poisson2d5(128, 128)
        return v, rp, c, np.ones(128 * 128), np.zeros(128 * 128), 0
    if name == "p2d_rand":
        v, rp, c = gen.poisson2d5(128, 128)
        return v, rp, c, gen.rhs_uniform(128 * 128, 0), np.zeros(128 * 128), 0
    if name == "p3d":
        v, rp, c = gen.poisson3d7(24)
        return v, rp, c, gen.rhs_uniform(24 ** 3, 0), np.zeros(24 ** 3), 0
    raise KeyError(name)


@pytest.mark.parametrize("name", ["tridiag10", "banded2000", "banded_min200", "p2d_ones", "p2d_rand", "p3d"])
def test_single_gpu_flow_matches_oracle(oracle, gen, name):
    from conjugategradient_b200 import solver as S
    v, rp, c, b, x0, min_it = _case(gen, name)
    n = len(b)
    cg = S.ConjugateGradientSingleGpu(n, 160, min_it, n, 1e-8)
    cg.A = S.SparseMatrix(v, rp, c); cg.b[:] = b; cg.x[:] = x0
    cg.Initialize(); cg.Solve(); cg.Read(); cg.close()
    xo, ito, reso, st = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8, min_iteration=min_it, max_iteration=n)
    assert st == oracle.OK
    assert abs(cg.Iteration - ito) <= IT_TOL, (cg.Iteration, ito)
    if cg.Iteration == ito:
        assert _relerr(cg.x, xo) <= REL_ERR
    else:   # one step apart: both are within the stop tolerance of the exact solution
        assert _relerr(cg.x, xo) <= 1e-6
    # the reference harness's own acceptance test (MgcgMain.cs:135): |x_cpu - x_gpu| / x_cpu <= 1 %
    big = np.abs(xo) > 1e-6 * np.abs(xo).max()
    assert np.all(np.abs(cg.x - xo)[big] / np.abs(xo)[big] <= 0.01)


def test_known_iteration_counts(oracle, gen):
    """SURVEY.md §8c derived known answers: banded N=2000 -> index 26; 2-D 128^2 b=1 -> 266."""
    from conjugategradient_b200 import solver as S
    v, rp, c, b, x0 = gen.banded_sin(2000, 160)
    cg = S.ConjugateGradientSingleGpu(2000, 160, 0, 2000, 1e-8)
    cg.A = S.SparseMatrix(v, rp, c); cg.b[:] = b; cg.x[:] = x0
    cg.Initialize(); cg.Solve(); cg.close()
    assert cg.Iteration == 26
    v, rp, c = gen.poisson2d5(128, 128)
    cg = S.ConjugateGradientSingleGpu(128 * 128, 5, 0, 128 * 128, 1e-8)
    cg.A = S.SparseMatrix(v, rp, c); cg.b[:] = 1.0
    cg.Initialize(); cg.Solve(); cg.close()
    assert abs(cg.Iteration - 266) <= IT_TOL


def test_solvegpu_entry(oracle, gen):
    """SimpleConjugateGradient.cu SolveGpu: x0 forced to 0, strict min-iteration, iteration = #performed."""
    from conjugategradient_b200 import solver as S
    v, rp, c, b = gen.tridiag121(10)
    x, it, res = S.SolveGpu(v, rp, c, b, 1e-8, 0, 10)
    xo, ito, reso, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, min_strict=True)
    assert it == ito + 1 == 10
    assert _relerr(x, xo) <= REL_ERR
    v, rp, c, b = gen.tridiag121(300)
    x, it, res = S.SolveGpu(v, rp, c, b, 1e-8, 0, 300)
    xo, ito, reso, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, min_strict=True)
    assert abs(it - 1 - ito) <= IT_TOL


@pytest.mark.parametrize("norm", [0, 1, 2])
@pytest.mark.parametrize("strict", [0, 1])
def test_stop_rules_native(oracle, gen, norm, strict):
    import torch
    from conjugategradient_b200 import solver as S
    v, rp, c = gen.poisson2d5(48, 40)
    n = 48 * 40
    b = gen.rhs_uniform(n, 3)
    tol = 1e-6
    m = S.Mgcg(0)
    tv, tr, tc = (torch.as_tensor(a).cuda() for a in (v, rp, c))
    m.setup(tv, tr, tc, S.default_options(tolerance=tol, norm=norm, min_strict=strict, min_iteration=3, max_iteration=n))
    tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
    res = m.solve(tb, tx)
    xo, ito, reso, st = oracle.cg_solve(v, rp, c, b, tol=tol, norm=norm, min_strict=bool(strict), min_iteration=3, max_iteration=n)
    assert abs(res.iteration - ito) <= IT_TOL and res.iterations_done == res.iteration + 1
    if res.iteration == ito:
        assert abs(res.residual - reso) <= 1e-6 * reso
        assert _relerr(tx.cpu().numpy(), xo) <= REL_ERR
    m.close()


def test_min_max_iteration_and_errors(oracle, gen):
    import torch
    from conjugategradient_b200 import _lib, solver as S
    v, rp, c, b = gen.tridiag121(50)
    tv, tr, tc, tb = (torch.as_tensor(a).cuda() for a in (v, rp, c, b))
    m = S.Mgcg(0)
    for strict, want in ((0, 60), (1, 61)):
        m.setup(tv, tr, tc, S.default_options(min_iteration=60, max_iteration=1000, min_strict=strict))
        tx = torch.zeros(50, dtype=torch.float64, device="cuda")
        assert m.solve(tb, tx).iteration == want
    # Iteration > MaxIteration: the C# throws (ConjugateGradient.cs:70-74) -> status NOT_CONVERGED at index max+1
    m.setup(tv, tr, tc, S.default_options(tolerance=1e-300, max_iteration=5))
    tx = torch.zeros(50, dtype=torch.float64, device="cuda")
    res = m.solve(tb, tx, raise_on_fail=False)
    assert res.status == _lib.ERR_NOT_CONVERGED and res.iteration == 6
    with pytest.raises(S.MgcgError):
        m.solve(tb, tx)
    # indefinite matrix -> breakdown status instead of the reference's silent NaN
    vi = v.copy(); vi[rp[:-1]] = -2.0
    tvi = torch.as_tensor(vi).cuda()
    m.setup(tvi, tr, tc, S.default_options(max_iteration=100))
    tx = torch.zeros(50, dtype=torch.float64, device="cuda")
    assert m.solve(tb, tx, raise_on_fail=False).status == _lib.ERR_BREAKDOWN
    # b == 0 with x0 == 0: already solved, stays finite
    m.setup(tv, tr, tc, S.default_options())
    tz = torch.zeros(50, dtype=torch.float64, device="cuda")
    res = m.solve(tz, tx.zero_())
    assert res.iteration == 0 and res.residual == 0.0 and torch.all(tx == 0)
    # invalid arguments
    assert m.L.mgcg_setup(m.ctx, 0, 0, None, None, None, None) == _lib.ERR_INVALID
    m.close()


def test_graph_and_eager_agree_bitwise(gen):
    import torch
    from conjugategradient_b200 import solver as S
    v, rp, c = gen.poisson3d7(20)
    n = 20 ** 3
    b = gen.rhs_uniform(n, 1)
    tv, tr, tc, tb = (torch.as_tensor(a).cuda() for a in (v, rp, c, b))
    xs = []
    for graph, interval in ((0, 1), (1, 8), (1, 3), (0, 16), (2, 8)):   # 2: device-side while loop, one launch
        m = S.Mgcg(0)
        m.setup(tv, tr, tc, S.default_options(use_graph=graph, check_interval=interval))
        tx = torch.zeros(n, dtype=torch.float64, device="cuda")
        res = m.solve(tb, tx)
        if graph == 2:      # no speculative iterations: exactly iterations x (3 kernels + the loop-condition kernel) + the 2 init kernels
            assert res.kernel_launches == 4 * res.iterations_done + 2, res.kernel_launches
            res = m.solve(tb, tx.zero_())    # replay
        xs.append((res.iteration, tx.cpu().numpy()))
        m.close()
    for it, x in xs[1:]:
        assert it == xs[0][0] and np.array_equal(x, xs[0][1])   # deterministic: speculation past convergence is a no-op


@pytest.mark.parametrize("devices", [1, 2, 3, 4])
def test_parallel_flow_matches_oracle(oracle, gen, devices):
    """ConjugateGradientParallelGpu flow (Initialize / SyncP / Solve0-3 / Read) with `devices` row blocks; blocks are
    mapped round-robin onto the GPUs present.  Oracle: ref_cg_solve_parallel (same partition, same summation order)."""
    from conjugategradient_b200 import solver as S
    v, rp, c, b, x0 = gen.banded_sin(3001, 160)
    n = 3001
    cg = S.ConjugateGradientParallelGpu(n, 160, 0, n, 1e-8, devices=devices)
    cg.A = S.SparseMatrix(v, rp, c); cg.b[:] = b; cg.x[:] = x0
    cg.Initialize()
    # partition + halo maps are integers: bit-exact
    assert cg.offsetsForDevice.tolist() == oracle.partition_offsets(n, devices).tolist()
    for d in range(devices):
        cnt = cg.CountForDevice(d)
        lro, eo, ec, mn, mx = oracle.initialize_block(rp, c, int(cg.offsetsForDevice[d]), cnt)
        assert (cg.minJ[d], cg.maxJ[d]) == (mn, mx)
        got = np.empty(cnt + 1, dtype=np.int32)
        cg.L.SetDevice(d % cg.physical)
        cg.L.CopyToArray_Int(cg.v[d]["rp"], got.ctypes.data_as(C.c_void_p), cnt + 1, 0, 0)
        assert np.array_equal(got, lro)                                          # rebased row offsets, Mgcg.cu:73
        assert cg._halo(d)[2:] == oracle.halo_counts(d, devices, int(cg.offsetsForDevice[d]), cnt, mn, mx)
    cg.Solve(); cg.Read(); cg.close()
    xo, ito, reso, st = oracle.cg_solve_parallel(v, rp, c, b, devices, x0, tol=1e-8)
    assert abs(cg.Iteration - ito) <= IT_TOL
    assert _relerr(cg.x, xo) <= REL_ERR


def test_solve_host_end_to_end(oracle, gen):
    from conjugategradient_b200 import solver as S
    v, rp, c, b, x0 = gen.banded_sin(4000, 160)
    m = S.Mgcg(0)
    x = x0.copy()
    res = m.solve_host(v, rp, c, b, x, S.default_options(max_iteration=4000))
    xo, ito, _, _ = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8)
    assert abs(res.iteration - ito) <= IT_TOL and _relerr(x, xo) <= REL_ERR
    m.close()


@pytest.mark.parametrize("name", ["banded2000", "p2d_rand", "p3d"])
@pytest.mark.parametrize("graph", [2, 0])
def test_single_reduce_cg_matches_oracle(oracle, gen, name, graph):
    """MGCG_CG_SINGLE_REDUCE (Chronopoulos-Gear; opt-in): the recurrence that finishes an iteration's sums at one
    point reaches the reference's tolerance at the reference's iteration (+-1) with the same solution to 1e-10 — the iterates are
    the classic ones in exact arithmetic."""
    import torch
    from conjugategradient_b200 import solver as S, _lib
    v, rp, c, b, x0, min_it = _case(gen, name)
    n = len(b)
    m = S.Mgcg(0)
    tv, tr, tc = (torch.as_tensor(a).cuda() for a in (v, rp, c))
    m.setup(tv, tr, tc, S.default_options(tolerance=1e-8, min_iteration=min_it, max_iteration=n, cg_variant=_lib.CG_SINGLE_REDUCE, use_graph=graph))
    tb, tx = torch.as_tensor(b).cuda(), torch.as_tensor(x0).cuda().clone()
    res = m.solve(tb, tx)
    xo, ito, reso, st = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8, min_iteration=min_it, max_iteration=n)
    assert st == 0 and res.status == 0
    assert abs(res.iteration - ito) <= IT_TOL, (res.iteration, ito)
    if res.iteration == ito:
        assert _relerr(tx.cpu().numpy(), xo) <= 1e-9
    import scipy.sparse as sp
    A = sp.csr_matrix((v, c, rp))
    assert np.linalg.norm(b - A @ tx.cpu().numpy()) < 2e-8 * max(1.0, np.linalg.norm(b))
    m.close()
