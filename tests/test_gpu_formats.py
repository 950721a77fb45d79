"""The reference's other two input surfaces on the GPU (include/mgcg_formats.h), against the oracle:
ELL in (OpenCL variant: SparseMatrix.cs, ConjugateGradientSingleGpu.cs:213-296, max|r| stop) and the
Write / Solve / Read / Iteration object (ViennaCL variant: ComputerGpu.hpp:68-103, relative stop, strict minimum).
Bars: index work bit-exact; fp64 relative solution error <= 1e-10, iteration count within +-1."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_ERR = 1e-10
IT_TOL = 1


def _relerr(x, xref):
    return np.abs(x - xref).max() / np.abs(xref).max()


def _ell_cases(gen):
    from conjugategradient_b200.formats import SparseMatrixEll
    v, rp, c, b, x0 = gen.banded_sin(700, 160)
    yield "banded", SparseMatrixEll.from_csr(v, rp, c, 160)               # ragged: 80..159 entries per row, width 160
    v, rp, c = gen.poisson3d7(9, 7, 5)
    yield "p3d_wide", SparseMatrixEll.from_csr(v, rp, c, 11)              # unused slots at the end of every row
    d = SparseMatrixEll(50, 4)                                            # diagonal only, one long row
    for i in range(50):
        d[i, i] = 2.0 + i
    d[7, 3] = -0.5; d[7, 9] = -0.25; d[7, 0] = -0.125
    yield "diag", d


def test_ell_to_csr_on_device_bit_exact(oracle, gen):
    import torch
    from conjugategradient_b200 import _lib, solver as S
    L = _lib.lib()
    m = S.Mgcg(0)
    for name, A in _ell_cases(gen):
        n, M = A.RowCount, A.MaxNonzeroCountPerRow
        te, tc, tn = torch.as_tensor(A.Elements).cuda(), torch.as_tensor(A.ColumnIndeces).cuda(), torch.as_tensor(A.NonzeroCounts).cuda()
        nnz = L.mgcg_ell_nnz(m.ctx, n, M, S._p(tn))
        assert nnz == int(A.NonzeroCounts.sum()), name
        val = torch.empty(nnz, dtype=torch.float64, device="cuda")
        col = torch.empty(nnz, dtype=torch.int32, device="cuda")
        rp = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        _lib.check(L.mgcg_ell_to_csr(m.ctx, n, M, S._p(te), S._p(tc), S._p(tn), S._p(val), S._p(rp), S._p(col)), "mgcg_ell_to_csr")
        torch.cuda.synchronize()
        ov, orp, oc = oracle.ell_to_csr(A.Elements, A.ColumnIndeces, A.NonzeroCounts, M)
        assert np.array_equal(rp.cpu().numpy(), orp) and np.array_equal(col.cpu().numpy(), oc) and np.array_equal(val.cpu().numpy(), ov), name
    m.close()


def test_ell_bad_counts_fail_loudly(gen):
    import torch
    from conjugategradient_b200 import _lib, solver as S
    L = _lib.lib()
    m = S.Mgcg(0)
    cnt = torch.tensor([1, 2, 0, 1], dtype=torch.int32, device="cuda")     # a row without its diagonal slot
    assert L.mgcg_ell_nnz(m.ctx, 4, 3, S._p(cnt)) < 0 and "outside [1, 3]" in _lib.last_error()
    cnt = torch.tensor([1, 4, 1, 1], dtype=torch.int32, device="cuda")     # more entries than the row can hold
    assert L.mgcg_ell_nnz(m.ctx, 4, 3, S._p(cnt)) < 0
    m.close()


@pytest.mark.parametrize("name", ["banded", "p3d", "p2d_min"])
def test_ell_solver_matches_oracle(oracle, gen, name):
    """construct -> fill A, x, b -> Initialize -> Solve, like MgcgCLMain; the OpenCL variant stops on max|r_i|."""
    from conjugategradient_b200.formats import ConjugateGradientEll, SparseMatrixEll
    min_it = 0
    if name == "banded":
        v, rp, c, b, x0 = gen.banded_sin(2000, 160); width = 160
    elif name == "p3d":
        v, rp, c = gen.poisson3d7(20); b = gen.rhs_uniform(20 ** 3, 1); x0 = np.zeros(20 ** 3); width = 7
    else:
        v, rp, c = gen.poisson2d5(64, 64); b = np.ones(64 * 64); x0 = np.zeros(64 * 64); width = 8; min_it = 150
    n = len(b)
    cg = ConjugateGradientEll(n, width, min_it, n, 1e-8)
    cg.A = SparseMatrixEll.from_csr(v, rp, c, width); cg.b[:] = b; cg.x[:] = x0
    cg.Initialize(); cg.Solve(); cg.close()
    xo, ito, reso, st = oracle.cg_solve_ell(cg.A.Elements, cg.A.ColumnIndeces, cg.A.NonzeroCounts, width, b, x0, tol=1e-8,
                                            min_iteration=min_it, max_iteration=n)
    assert st == oracle.OK
    assert abs(cg.Iteration - ito) <= IT_TOL, (cg.Iteration, ito)
    assert cg.Residual < 1e-8
    assert _relerr(cg.x, xo) <= (REL_ERR if cg.Iteration == ito else 1e-6)


def test_ell_solver_throws_past_max_iteration(gen):
    from conjugategradient_b200.formats import ConjugateGradientEll, SparseMatrixEll
    v, rp, c = gen.poisson2d5(48, 48)
    cg = ConjugateGradientEll(48 * 48, 5, 0, 5, 1e-12)
    cg.A = SparseMatrixEll.from_csr(v, rp, c, 5); cg.b[:] = 1.0
    with pytest.raises(RuntimeError):
        cg.Solve()
    cg.close()


@pytest.mark.parametrize("name,min_it", [("banded", 0), ("p3d", 0), ("p3d", 60)])
def test_computer_four_calls_match_oracle(oracle, gen, name, min_it):
    """ComputerGpu(n); Write; Solve(residual, min, max); Read; Iteration — relative stop rr/rr0 < residual^2, `min < iteration`."""
    from conjugategradient_b200.formats import ComputerGpu
    if name == "banded":
        v, rp, c, b, x0 = gen.banded_sin(1500, 160)
    else:
        v, rp, c = gen.poisson3d7(18); b = gen.rhs_uniform(18 ** 3, 3); x0 = 0.01 * gen.rhs_uniform(18 ** 3, 4)
    n = len(b)
    k = ComputerGpu(n)
    k.Write(v, rp.astype(np.uint32), c.astype(np.uint32), x0, b)
    k.Solve(1e-9, min_it, n)
    x = np.empty(n)
    k.Read(x)
    xo, ito, reso, st = oracle.cg_solve(v, rp, c, b, x0, tol=1e-9, min_iteration=min_it, max_iteration=n,
                                        norm=oracle.NORM_REL_L2, min_strict=True)
    assert st == oracle.OK and ito > min_it
    assert abs(k.Iteration() - (ito + 1)) <= IT_TOL, (k.Iteration(), ito + 1)      # iterations performed (ComputerGpu.cpp:59,78)
    assert _relerr(x, xo) <= (REL_ERR if k.Iteration() == ito + 1 else 1e-6)
    # the object is reusable: Write again, same answer
    k.Write(v, rp.astype(np.uint32), c.astype(np.uint32), x0, b)
    k.Solve(1e-9, min_it, n)
    x2 = np.empty(n)
    k.Read(x2)
    assert np.array_equal(x, x2)
    k.close()


def test_computer_with_multigrid_and_errors(oracle, gen):
    from conjugategradient_b200 import _lib
    from conjugategradient_b200.formats import ComputerGpu
    n1 = 24
    v, rp, c = gen.poisson3d7(n1); b = gen.rhs_uniform(n1 ** 3, 7)
    k = ComputerGpu(n1 ** 3)
    with pytest.raises(_lib.MgcgError):
        k.Solve(1e-8, 0, 100)                                             # nothing written yet
    k.Write(v, rp.astype(np.uint32), c.astype(np.uint32), np.zeros(n1 ** 3), b)
    k.Solve(1e-9, 0, 0)                                                   # max <= 0: unbounded, like the reference
    plain = k.Iteration()
    x_plain = np.empty(n1 ** 3); k.Read(x_plain)
    k.SetPreconditioner(_lib.PRECOND_MG, n1, n1, n1)
    k.Write(v, rp.astype(np.uint32), c.astype(np.uint32), np.zeros(n1 ** 3), b)
    k.Solve(1e-9, 0, 0)
    x_mg = np.empty(n1 ** 3); k.Read(x_mg)
    assert k.Iteration() < plain // 2 and k.Result().levels >= 2
    assert _relerr(x_mg, x_plain) <= 1e-6          # both stopped at a relative residual of 1e-9, cond ~ 2.5e2
    bad = rp.astype(np.uint32).copy(); bad[-1] += 1
    with pytest.raises(_lib.MgcgError):
        k.Write(v, bad, c.astype(np.uint32), np.zeros(n1 ** 3), b)
    k.close()
