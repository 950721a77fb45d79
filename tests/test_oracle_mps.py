"""oracle/mps_ref.c — the CPU statement of the particle-matrix assembly that is to move onto the device,
cross-checked against an independent kd-tree neighbour search.  No GPU needed."""
import numpy as np
import pytest


def _cloud(n_side, jitter, seed):
    rng = np.random.default_rng(seed)
    g = np.arange(n_side)
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    pos = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1) + rng.uniform(-jitter, jitter, size=(n_side ** 3, 3))
    return pos[rng.permutation(len(pos))]            # any input order


@pytest.mark.parametrize("n_side,re", [(6, 2.1), (9, 2.1), (7, 3.1)])
def test_cell_list_assembly_matches_kdtree(oracle, n_side, re):
    from scipy.spatial import cKDTree
    pos = _cloud(n_side, 0.2, n_side)
    val, rp, col, order = oracle.mps_assemble(pos, re, shift_abs=0.25, max_per_row=160)
    n = len(pos)
    assert sorted(order.tolist()) == list(range(n))
    p = pos[order]
    pairs = cKDTree(p).query_pairs(re, output_type="ndarray")
    i = np.concatenate([pairs[:, 0], pairs[:, 1]]); j = np.concatenate([pairs[:, 1], pairs[:, 0]])
    d = p[i] - p[j]
    r = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])
    keep = r < re                                     # the tree's own test may differ in the last bit at r == re
    i, j, r = i[keep], j[keep], r[keep]
    o = np.lexsort((j, i)); i, j, r = i[o], j[o], r[o]
    cnt = np.bincount(i, minlength=n) + 1
    assert np.array_equal(rp, np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32))
    off = np.ones(len(val), dtype=bool); off[rp[:-1]] = False
    assert np.array_equal(col[rp[:-1]], np.arange(n))                      # diagonal first
    assert np.array_equal(col[off], j)                                     # then ascending columns, no per-row sort needed
    assert np.array_equal(val[off], -(re / r - 1.0))                       # same roundings, entry for entry
    diag = np.zeros(n)
    for k in range(len(i)):                                                # sequential, in column order
        diag[i[k]] += re / r[k] - 1.0
    assert np.array_equal(val[rp[:-1]], diag + 0.25)
    # symmetric, weakly diagonally dominant with a positive shift: SPD
    import scipy.sparse as sp
    A = sp.csr_matrix((val, col, rp), shape=(n, n))
    assert abs(A - A.T).max() == 0.0


def test_row_limit_is_reported(oracle):
    pos = _cloud(6, 0.2, 1)
    with pytest.raises(ValueError):
        oracle.mps_assemble(pos, 2.1, max_per_row=8)                       # the reference's container throws on a full row
