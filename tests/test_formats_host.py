"""Host side of the reference's other two input surfaces (conjugategradient_b200/formats.py, include/mgcg_formats.h):
the ELL container of the OpenCL variant and its CPU restatement in the oracle.  No GPU needed."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ell_container_follows_the_reference_rules():
    """SparseMatrix.cs (HandmadeCL): slot 0 is the diagonal, new entries are appended, a full row throws."""
    from conjugategradient_b200.formats import SparseMatrixEll
    m = SparseMatrixEll(4, 3)
    assert m.NonzeroCounts.tolist() == [1, 1, 1, 1]                       # Clear(): every row keeps its diagonal (:52-63)
    assert m.ColumnIndeces[::3].tolist() == [0, 1, 2, 3] and not m.Elements.any()
    m[1, 3] = -2.0                                                        # appended behind the diagonal (:136-153)
    m[1, 1] = 5.0                                                         # the diagonal always lands in slot 0 (:82-91)
    m[1, 0] = -1.0
    assert m.NonzeroCounts[1] == 3
    assert m.ColumnIndeces[3:6].tolist() == [1, 3, 0] and m.Elements[3:6].tolist() == [5.0, -2.0, -1.0]
    m[1, 3] = -4.0                                                        # an existing entry is overwritten in place
    assert m.NonzeroCounts[1] == 3 and m[1, 3] == -4.0 and m[1, 2] == 0.0 and m[1, 1] == 5.0
    with pytest.raises(IndexError):
        m[1, 2] = 1.0                                                     # the row is full (:138-141)
    m.Clear()
    assert m.NonzeroCounts.tolist() == [1, 1, 1, 1] and m[1, 1] == 0.0


def test_ell_from_csr_and_back(gen):
    from conjugategradient_b200.formats import SparseMatrixEll
    v, rp, c = gen.poisson3d7(6, 5, 4)                                    # rows store the diagonal first, like the reference generator
    m = SparseMatrixEll.from_csr(v, rp, c)
    assert m.MaxNonzeroCountPerRow == 7 and m.NonzeroCounts.min() == 4 and m.NonzeroCounts.max() == 7
    v2, rp2, c2 = m.to_csr()
    assert np.array_equal(v2, v) and np.array_equal(rp2, rp) and np.array_equal(c2, c)
    # entry by entry, as a caller of the reference would fill it
    e = SparseMatrixEll(len(rp) - 1, 7)
    for i in range(len(rp) - 1):
        for k in range(rp[i], rp[i + 1]):
            e[i, int(c[k])] = v[k]
    assert np.array_equal(e.NonzeroCounts, m.NonzeroCounts)
    for i in range(len(rp) - 1):
        s = slice(i * 7, i * 7 + int(m.NonzeroCounts[i]))
        assert np.array_equal(e.Elements[s], m.Elements[s]) and np.array_equal(e.ColumnIndeces[s], m.ColumnIndeces[s])
    with pytest.raises(IndexError):
        SparseMatrixEll.from_csr(v, rp, c, 6)


def test_oracle_ell_multiply_is_the_csr_multiply_in_slot_order(oracle, gen):
    from conjugategradient_b200.formats import SparseMatrixEll
    v, rp, c, b, x0 = gen.banded_sin(300, 160)
    m = SparseMatrixEll.from_csr(v, rp, c, 160)
    cv, crp, cc = oracle.ell_to_csr(m.Elements, m.ColumnIndeces, m.NonzeroCounts, 160)
    hv, hrp, hc = m.to_csr()
    assert np.array_equal(cv, hv) and np.array_equal(crp, hrp) and np.array_equal(cc, hc)
    x = gen.rhs_uniform(300, 5)
    assert np.array_equal(oracle.ell_multiply(m.Elements, m.ColumnIndeces, m.NonzeroCounts, 160, x), oracle.spmv(cv, crp, cc, x))


def test_oracle_ell_solver_known_answer(oracle, gen):
    """Pinned like the CSR oracle: the tridiagonal [1 2 1] N=10 system against the committed direct solve."""
    from conjugategradient_b200.formats import SparseMatrixEll
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "tridiag10.json")))
    v, rp, c, b = gen.tridiag121(10)
    m = SparseMatrixEll.from_csr(v, rp, c, 3)
    x, it, res, st = oracle.cg_solve_ell(m.Elements, m.ColumnIndeces, m.NonzeroCounts, 3, b, tol=1e-8)
    assert st == oracle.OK and res < 1e-8
    A = np.zeros((10, 10))
    for i in range(10):
        A[i, c[rp[i]:rp[i + 1]]] = v[rp[i]:rp[i + 1]]
    np.testing.assert_allclose(x, np.linalg.solve(A, b), rtol=1e-9)
    np.testing.assert_allclose(x, gold["x"], rtol=1e-6)       # the 7-digit values recorded in SURVEY.md §8c


def test_oracle_ell_solver_equals_the_linf_csr_loop(oracle, gen):
    """The OpenCL variant's GPU loop and its CPU class (norm = ABS_LINF in ref_cg_solve) are the same recurrence."""
    from conjugategradient_b200.formats import SparseMatrixEll
    v, rp, c, b, x0 = gen.banded_sin(400, 160)
    m = SparseMatrixEll.from_csr(v, rp, c, 160)
    x1, it1, r1, s1 = oracle.cg_solve_ell(m.Elements, m.ColumnIndeces, m.NonzeroCounts, 160, b, x0, tol=1e-8)
    cv, crp, cc = m.to_csr()
    x2, it2, r2, s2 = oracle.cg_solve(cv, crp, cc, b, x0, tol=1e-8, norm=oracle.NORM_ABS_LINF)
    assert (it1, s1) == (it2, s2) and r1 == r2 and np.array_equal(x1, x2)
