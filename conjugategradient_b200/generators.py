"""Synthetic input generators for the MGCG hot path (host side, numpy).

The reference ships only two generators (tridiagonal [1 2 1] and a banded |sin(i+j)| matrix); the Poisson and
particle matrices named by BASELINE.json do not exist in it and are produced here.
All return CSR as (elements fp64[nnz], row_offsets int32[n+1], column_indeces int32[nnz]) with the
DIAGONAL STORED FIRST in every row, like the reference generators (MgcgMain.cs:59-60,
SimpleConjugateGradient.cu:168-170), 0-based, lexicographic idx = x + nx*(y + ny*z).
"""
import numpy as np


def tridiag121(n):
    """SimpleConjugateGradient.cu:163-197: [1 2 1], b[i] = (int32)(i*i) * 0.5 (wraps for i >= 46341)."""
    i = np.arange(n, dtype=np.int64)
    cnt = np.full(n, 3, dtype=np.int64)
    cnt[0] -= 1
    cnt[-1] -= 1 if n > 1 else 0
    if n == 1:
        cnt[0] = 1
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    val = np.ones(rp[-1])
    col = np.empty(rp[-1], dtype=np.int32)
    val[rp[:-1]] = 2.0
    col[rp[:-1]] = i
    has_l = i > 0
    col[rp[:-1][has_l] + 1] = i[has_l] - 1
    has_r = i < n - 1
    col[rp[:-1][has_r] + 1 + has_l[has_r]] = i[has_r] + 1
    ii = (i * i).astype(np.uint32).astype(np.int32)  # 32-bit wrap like the reference's int arithmetic
    b = ii.astype(np.float64) * 0.5
    return val, rp.astype(np.int32), col, b


def banded_sin(n, max_nonzero=160):
    """MgcgMain.cs:51-104: a_ij = |sin(i+j)|, j in [max(0,i-M/2+1), min(n,i+M/2)), diag = sum offdiag (stored first);
    b = 10 cos(i), x0 = i/100."""
    h = max_nonzero // 2
    i = np.arange(n, dtype=np.int64)
    j0 = np.maximum(0, i - h + 1)
    j1 = np.minimum(n, i + h)
    cnt = j1 - j0  # includes the diagonal slot (j==i is skipped, diagonal is stored first)
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    nnz = int(rp[-1])
    row = np.repeat(i, cnt)
    k = np.arange(nnz, dtype=np.int64) - rp[:-1][row]  # slot within the row
    # slot 0 = diagonal; slots 1.. = ascending j without i
    jj = j0[row] + (k - 1)
    jj = np.where(jj >= row, jj + 1, jj)
    col = np.where(k == 0, row, jj).astype(np.int32)
    val = np.abs(np.sin((row + jj).astype(np.float64)))
    val[k == 0] = 0.0
    # diagonal = sequential sum of the off-diagonals in storage order (MgcgMain.cs:79)
    diag = np.zeros(n)
    maxc = int(cnt.max())
    for s in range(1, maxc):
        m = cnt > s
        diag[m] += val[rp[:-1][m] + s]
    val[rp[:-1]] = diag
    b = np.cos(i.astype(np.float64)) * 10
    x0 = i.astype(np.float64) / 100
    return val, rp.astype(np.int32), col, b, x0


def _stencil_csr(dims, offsets, coef_fn, zrange=None):
    """Generic structured stencil -> CSR with diagonal first. offsets: list of (dx,dy,dz) excluding (0,0,0).
    coef_fn(mask_idx, off_index) -> off-diagonal values for the rows in mask.
    zrange=(z0, z1): only the rows of planes z0 <= z < z1 (a shard), column indices stay global, row offsets start at 0."""
    nx, ny, nz = dims
    if zrange is None:
        idx = np.arange(nx * ny * nz, dtype=np.int64)
    else:
        idx = np.arange(zrange[0] * nx * ny, zrange[1] * nx * ny, dtype=np.int64)
    n = len(idx)
    x = idx % nx
    y = (idx // nx) % ny
    z = idx // (nx * ny)
    valid = []
    for (dx, dy, dz) in offsets:
        valid.append((x + dx >= 0) & (x + dx < nx) & (y + dy >= 0) & (y + dy < ny) & (z + dz >= 0) & (z + dz < nz))
    cnt = 1 + np.sum(valid, axis=0).astype(np.int64)
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    nnz = int(rp[-1])
    val = np.zeros(nnz)
    col = np.empty(nnz, dtype=np.int32)
    col[rp[:-1]] = idx
    slot = np.ones(n, dtype=np.int64)
    for o, (dx, dy, dz) in enumerate(offsets):
        m = valid[o]
        pos = rp[:-1][m] + slot[m]
        col[pos] = idx[m] + dx + nx * (dy + ny * dz)
        val[pos] = coef_fn(idx[m], o)
        slot[m] += 1
    return val, rp, col, n


def poisson2d5(nx, ny):
    """C1: 2-D 5-point Poisson, diag 4, off -1, homogeneous Dirichlet eliminated."""
    offs = [(-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0)]
    val, rp, col, n = _stencil_csr((nx, ny, 1), offs, lambda idx, o: -1.0)
    val[rp[:-1]] = 4.0
    return val, rp.astype(np.int32), col


def poisson3d7(nx, ny=None, nz=None, zrange=None):
    """C2/C4: 3-D 7-point Poisson, diag 6, off -1, homogeneous Dirichlet eliminated.  zrange: see _stencil_csr."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    offs = [(-1, 0, 0), (1, 0, 0), (0, -1, 0), (0, 1, 0), (0, 0, -1), (0, 0, 1)]
    val, rp, col, n = _stencil_csr((nx, ny, nz), offs, lambda idx, o: -1.0, zrange)
    val[rp[:-1]] = 6.0
    assert rp[-1] < 2 ** 31
    return val, rp.astype(np.int32), col


def poisson3d27_varcoef(nx, ny=None, nz=None, seed=3, block=8, contrast_decades=3.0, aniso=(1.0, 1.0, 1e-2)):
    """C5: 27-point variable-coefficient operator.  Edge weight between nodes i and j =
    harmonic-free symmetric mean sqrt(k_i k_j) * w(d), k = 10^U(-c,c) piecewise constant on block^3 cells,
    w(d) = anisotropy-weighted inverse squared distance class; diag = sum |offdiag| + Dirichlet closure
    (rows next to the boundary keep the missing neighbours' weight on the diagonal) => SPD M-matrix."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    rng = np.random.default_rng(seed)
    bx, by, bz = -(-nx // block), -(-ny // block), -(-nz // block)
    kblk = 10.0 ** rng.uniform(-contrast_decades, contrast_decades, size=bx * by * bz)
    n = nx * ny * nz
    idx = np.arange(n, dtype=np.int64)
    x = idx % nx
    y = (idx // nx) % ny
    z = idx // (nx * ny)

    def kappa(xx, yy, zz):
        xx = np.clip(xx, 0, nx - 1); yy = np.clip(yy, 0, ny - 1); zz = np.clip(zz, 0, nz - 1)
        return kblk[(xx // block) + bx * ((yy // block) + by * (zz // block))]

    offs = [(dx, dy, dz) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1) if (dx, dy, dz) != (0, 0, 0)]

    def w(dx, dy, dz):
        d2 = dx * dx + dy * dy + dz * dz  # 1 face, 2 edge, 3 corner
        a = (abs(dx) * aniso[0] + abs(dy) * aniso[1] + abs(dz) * aniso[2]) / d2
        return a / d2

    k_self = kappa(x, y, z)
    diag = np.zeros(n)
    for (dx, dy, dz) in offs:  # full weight incl. eliminated Dirichlet neighbours
        diag += np.sqrt(k_self * kappa(x + dx, y + dy, z + dz)) * w(dx, dy, dz)

    def coef(ids, o):
        dx, dy, dz = offs[o]
        xi, yi, zi = ids % nx, (ids // nx) % ny, ids // (nx * ny)
        return -np.sqrt(kappa(xi, yi, zi) * kappa(xi + dx, yi + dy, zi + dz)) * w(dx, dy, dz)

    val, rp, col, n = _stencil_csr((nx, ny, nz), offs, coef)
    val[rp[:-1]] = diag
    assert rp[-1] < 2 ** 31, "global nnz exceeds int32: generate per shard"
    return val, rp.astype(np.int32), col


def mps_particles(n_side, re_over_l0=2.1, jitter=0.2, seed=1, shift=1e-3):
    """C3: MPS-style particle pressure-Poisson matrix.  n_side^3 particles on a unit lattice jittered by
    U(-jitter,jitter)*l0, Morton (cell) sorted; neighbours r < re; A_ij = -w(r), w = re/r - 1;
    A_ii = sum w + shift*mean(diag) (stand-in for the free-surface Dirichlet condition) => SPD.
    Rows: diagonal first, then neighbours in ascending column order."""
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(seed)
    g = np.arange(n_side)
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    cell = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1)
    pos = cell + rng.uniform(-jitter, jitter, size=cell.shape)

    def part1by2(v):
        v = v.astype(np.uint64) & np.uint64(0x1FFFFF)
        v = (v | (v << np.uint64(32))) & np.uint64(0x1F00000000FFFF)
        v = (v | (v << np.uint64(16))) & np.uint64(0x1F0000FF0000FF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
        v = (v | (v << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
        return v

    key = part1by2(cell[:, 0]) | (part1by2(cell[:, 1]) << np.uint64(1)) | (part1by2(cell[:, 2]) << np.uint64(2))
    order = np.argsort(key, kind="stable")
    pos = pos[order]
    n = pos.shape[0]
    tree = cKDTree(pos)
    pairs = tree.query_pairs(re_over_l0, output_type="ndarray")
    i = np.concatenate([pairs[:, 0], pairs[:, 1]])
    j = np.concatenate([pairs[:, 1], pairs[:, 0]])
    r = np.linalg.norm(pos[i] - pos[j], axis=1)
    w = re_over_l0 / r - 1.0
    o = np.lexsort((j, i))
    i, j, w = i[o], j[o], w[o]
    cnt = np.bincount(i, minlength=n).astype(np.int64) + 1
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(cnt, out=rp[1:])
    nnz = int(rp[-1])
    val = np.empty(nnz)
    col = np.empty(nnz, dtype=np.int32)
    starts = rp[:-1]
    within = np.arange(len(i), dtype=np.int64) - (starts[i] - np.arange(n, dtype=np.int64)[i])
    # position of off-diagonal e (sorted by row) = rp[row] + 1 + rank within row
    rank = np.arange(len(i), dtype=np.int64) - np.concatenate([[0], np.cumsum(cnt - 1)])[:-1][i]
    del within
    posn = starts[i] + 1 + rank
    val[posn] = -w
    col[posn] = j
    diag = np.bincount(i, weights=w, minlength=n)
    diag = diag + shift * diag.mean()
    val[starts] = diag
    col[starts] = np.arange(n, dtype=np.int32)
    assert nnz < 2 ** 31
    return val, rp.astype(np.int32), col


def rhs_uniform(n, seed=0):
    """b ~ U(-1,1), numpy default_rng(seed)."""
    return np.random.default_rng(seed).uniform(-1.0, 1.0, n)
