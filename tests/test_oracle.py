"""CPU oracle (oracle/cg_ref.c) against the known answers derivable from the reference's own inputs.
The reference ships no golden vectors; these are its generators + a direct solve (the R/CG.R:34 idea)."""
import json
import os

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spl

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_tridiag10_known_answer(oracle):
    # SimpleConjugateGradientCpu.cpp's N=10 input, solved to 1e-8 with the strict min-iteration rule of the .cu
    v, rp, c, b = oracle.gen_tridiag121(10)
    x, it, res, st = oracle.cg_solve(v, rp, c, b, tol=1e-8, min_strict=True)
    assert st == oracle.OK and it == 9
    A = sp.csr_matrix((v, c, rp)).toarray()
    np.testing.assert_allclose(x, np.linalg.solve(A, b), rtol=1e-12)
    gold = json.load(open(os.path.join(GOLD, "tridiag10.json")))
    np.testing.assert_allclose(x, gold["x"], rtol=1e-6)       # the 7-digit values recorded in SURVEY.md §8c
    assert it == gold["iteration_index"]


def test_tridiag_int_wrap(oracle, gen):
    # b[i] = (int32)(i*i) * 0.5 wraps negative from i = 46341 (SimpleConjugateGradient.cu:196)
    v, rp, c, b = oracle.gen_tridiag121(46400)
    assert b[46340] == 46340 * 46340 * 0.5 and b[46341] < 0
    v2, rp2, c2, b2 = gen.tridiag121(46400)
    assert np.array_equal(v, v2) and np.array_equal(rp, rp2) and np.array_equal(c, c2) and np.array_equal(b, b2)


def test_banded_sin_known_answer(oracle, gen):
    v, rp, c, b, x0 = oracle.gen_banded_sin(2000, 160)
    g = gen.banded_sin(2000, 160)
    for a, bb in zip((v, rp, c, b, x0), g):
        assert np.array_equal(a, bb)
    assert np.all(c[rp[:-1]] == np.arange(2000))               # diagonal stored first (MgcgMain.cs:59-60)
    x, it, res, st = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8)
    assert st == oracle.OK and it == 26                        # SURVEY.md §8c case (2)
    A = sp.csr_matrix((v, c, rp)).tocsc()
    np.testing.assert_allclose(x, spl.spsolve(A, b), rtol=1e-7, atol=1e-10)   # |r| < 1e-8, lambda_min ~ 46
    # MIN_ITERATION = 200 pads the loop (MgcgMain.cs:25): the index is exactly 200
    x2, it2, res2, st2 = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8, min_iteration=200)
    assert it2 == 200 and np.isfinite(res2)


def test_poisson2d_c1_iterations(oracle, gen):
    v, rp, c = gen.poisson2d5(128, 128)
    n = 128 * 128
    x, it, res, st = oracle.cg_solve(v, rp, c, np.ones(n), tol=1e-8)
    assert it == 266                                           # SURVEY.md §8c case (3)
    x, it, res, st = oracle.cg_solve(v, rp, c, gen.rhs_uniform(n, 0), tol=1e-8)
    assert it == 442
    x, it, res, st = oracle.cg_solve(v, rp, c, gen.rhs_uniform(n, 0), tol=1e-8, norm=oracle.NORM_REL_L2)
    assert it == 382


def test_stop_rules(oracle):
    v, rp, c, b = oracle.gen_tridiag121(50)
    _, it_ns, _, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, min_iteration=60, max_iteration=1000)
    _, it_s, _, _ = oracle.cg_solve(v, rp, c, b, tol=1e-8, min_iteration=60, max_iteration=1000, min_strict=True)
    assert it_ns == 60 and it_s == 61                          # `min <= it` vs `min < it`
    _, it, _, st = oracle.cg_solve(v, rp, c, b, tol=1e-300, max_iteration=5)
    assert st == oracle.MAX_ITERATION_EXCEEDED and it == 6     # C# throws when Iteration > Max
    xi, iti, resi, _ = oracle.cg_solve(v, rp, c, b, tol=1e-6, norm=oracle.NORM_ABS_LINF)
    A = sp.csr_matrix((v, c, rp))
    assert abs(np.abs(b - A @ xi).max() - resi) < 1e-9 and resi < 1e-6


def test_partition_and_halo_maps(oracle):
    # ConjugateGradientParallelGpu.cs:271-277,397-398 and Mgcg.cu:68-84
    off = oracle.partition_offsets(207402, 4)
    assert off.tolist() == [0, 51850, 103700, 155550, 207402]
    assert oracle.partition_offsets(10, 3).tolist() == [0, 3, 6, 10]
    assert oracle.partition_offsets(5, 8).tolist() == [0, 0, 0, 0, 0, 0, 0, 0, 5]
    v, rp, c, b, x0 = oracle.gen_banded_sin(1000, 160)
    off = oracle.partition_offsets(1000, 4)
    for d in range(4):
        cnt = int(off[d + 1] - off[d])
        lro, eo, ec, mn, mx = oracle.initialize_block(rp, c, int(off[d]), cnt)
        assert lro[0] == 0 and lro[-1] == ec and eo == rp[off[d]]
        assert mn == max(0, off[d] - 79) and mx == min(999, off[d + 1] - 1 + 79)
        lc, nc = oracle.halo_counts(d, 4, int(off[d]), cnt, mn, mx)
        assert lc == (79 if d > 0 else 0) and nc == (79 if d < 3 else 0)


def test_parallel_restatement_matches_sequential(oracle):
    v, rp, c, b, x0 = oracle.gen_banded_sin(3000, 160)
    x1, it1, r1, _ = oracle.cg_solve(v, rp, c, b, x0, tol=1e-8)
    for D in (1, 2, 3, 8):
        xd, itd, rd, st = oracle.cg_solve_parallel(v, rp, c, b, D, x0, tol=1e-8)
        assert st == oracle.OK and itd == it1
        np.testing.assert_allclose(xd, x1, rtol=1e-11, atol=1e-13)


def test_mg_twin_against_direct_and_plain_cg(oracle, gen):
    # multigrid has no reference (parity unpinned): the twin must at least solve the same system as the plain oracle
    n = 16
    v, rp, c = gen.poisson3d7(n)
    b = gen.rhs_uniform(n ** 3, 0)
    xd = spl.spsolve(sp.csr_matrix((v, c, rp)).tocsc(), b)
    xc, itc, _, _ = oracle.cg_solve(v, rp, c, b, tol=1e-10)
    for sm in (0, 1):
        H = oracle.MgRef(v, rp, c, n, n, n, smoothed=sm)
        assert H.levels() == 2
        x, it, res, st = H.pcg(b, tol=1e-10)
        assert st == 0 and it < itc / 2
        np.testing.assert_allclose(x, xd, rtol=1e-8, atol=1e-10)
        # M is symmetric: <M u, w> == <u, M w>
        u, w = gen.rhs_uniform(n ** 3, 5), gen.rhs_uniform(n ** 3, 6)
        assert abs(H.apply(u) @ w - u @ H.apply(w)) < 1e-10 * abs(H.apply(u) @ w)
        H.close()


def test_threaded_port_matches_sequential(oracle, gen):
    """cg_ref_mt.c (bench CPU baseline: the reference's commented-out Parallel.For loops over OpenMP) vs cg_ref.c."""
    v, rp, c = gen.poisson3d7(20)
    b = gen.rhs_uniform(20 ** 3, 0)
    x1, i1, r1, s1 = oracle.cg_solve(v, rp, c, b, tol=1e-8, norm=oracle.NORM_REL_L2)
    xs, is_, rs, ss = oracle.cg_solve_mt(v, rp, c, b, tol=1e-8, norm=oracle.NORM_REL_L2, threads=1)
    assert (is_, ss) == (i1, s1) and np.array_equal(xs, x1)          # one thread: bit-identical
    for t in (3, 0):
        x2, i2, r2, s2 = oracle.cg_solve_mt(v, rp, c, b, tol=1e-8, norm=oracle.NORM_REL_L2, threads=t)
        assert s2 == s1 and abs(i2 - i1) <= 1
        assert np.abs(x2 - x1).max() / np.abs(x1).max() <= 1e-10
    # bounded sample: max_iteration stops the loop like the C# throw
    _, it, _, st = oracle.cg_solve_mt(v, rp, c, b, tol=0.0, max_iteration=8, threads=2)
    assert st == oracle.MAX_ITERATION_EXCEEDED and it == 9


def test_coded_columns_restates_the_columns():
    """the checker of the coded-column layout (csrc/spmv_code.cu): stencil matrices have a handful of distances col - row, a shard's
    local rows shift them by a constant, an irregular matrix has too many"""
    from conjugategradient_b200 import generators as gen
    from oracle import oracle as O
    v, rp, c = gen.poisson3d7(9, 7, 5)
    d, code = O.coded_columns(rp, c)
    assert list(d) == [-63, -9, -1, 0, 1, 9, 63] and code.dtype == np.uint8
    row_of = np.repeat(np.arange(len(rp) - 1), np.diff(rp))
    assert np.array_equal(row_of + d[code], c)
    v, rp, c = gen.poisson3d7(9, 7, 6, zrange=(2, 4))           # rows of a z-slab shard, global columns
    d2, code2 = O.coded_columns(rp, c)
    assert list(d2) == [x + 2 * 63 for x in d]
    rng = np.random.default_rng(0)
    rp = np.arange(0, 4 * 1000 + 1, 4, dtype=np.int32)
    assert O.coded_columns(rp, rng.integers(0, 1000, 4000).astype(np.int32)) is None


def test_stencil_run_addressing_model():
    """The index algebra of the STENCIL kernel (csrc/kernels_spmv.cu k_spmv_stencil), restated in numpy: for a tile of R rows starting
    at the even row r0, run g of the distances is staged from the index s0 = ((r0 + first_g - xpar) & ~1) + xpar (16-byte aligned
    source when x itself is misaligned by xpar doubles), and row r0 + t finds the operand of an entry of distance d at
    s0 + t + (d - first_g) + ((first_g - xpar) & 1) — which must be its column, within a run of R + 8 doubles."""
    from conjugategradient_b200 import generators as gen
    from oracle import oracle as O
    for (v, rp, c), nruns in ((gen.poisson3d7(12, 10, 9), 5), (gen.poisson3d27_varcoef(10), 9)):
        dist, code = O.coded_columns(rp, c)
        firsts, run_of = O.coded_runs(dist)
        assert len(firsts) == nruns
        row_of = np.repeat(np.arange(len(rp) - 1), np.diff(rp))
        for R in (256, 64):
            for xpar in (0, 1):
                r0 = (row_of // R) * R
                t = row_of - r0
                first = firsts[run_of[code]]
                s0 = ((r0 + first - xpar) & ~1) + xpar
                pos = t + (dist[code].astype(np.int64) - first) + ((first - xpar) & 1)
                assert np.array_equal(s0 + pos, c)
                assert pos.min() >= 0 and pos.max() < R + 8
                assert np.all((s0 + xpar) % 2 == 0)            # (x address / 8 + s0) even: the bulk copy's source is 16-byte aligned
