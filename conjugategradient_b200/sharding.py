"""Host-side shard arithmetic of the one-process-per-GPU path (no device work here).

The layout mirrors csrc/dist.cuh: rank r owns the contiguous row block [off[r], off[r+1]) with GLOBAL column indices and
row offsets rebased to 0 (Mgcg.cu:73); a sharded vector is stored as [lo ghost | own | hi ghost] where the ghost width is
the matrix's global half bandwidth — the generalisation of the reference's minJ/maxJ rule
(Mgcg.cu:83-84, ConjugateGradientParallelGpu.cs:397-398)."""
import numpy as np


def partition_offsets(count, world):
    """The reference's row-block rule, bit for bit (ConjugateGradientParallelGpu.cs:271-277)."""
    off = [0] * (world + 1)
    for i in range(1, world):
        off[i] = off[i - 1] + int(np.floor(count / world))
    off[world] = count
    return off


def plane_partition(n_side, world):
    """z-plane aligned row blocks of an n^3 grid: equals partition_offsets(n^3, world) whenever world divides n."""
    planes = [n_side // world] * world
    planes[-1] += n_side - sum(planes)
    z = np.concatenate([[0], np.cumsum(planes)]).astype(np.int64)
    return [int(v) for v in z], [int(v) * n_side * n_side for v in z]


def shard_csr(val, rowptr, col, r0, r1):
    """Rows [r0, r1) of a global CSR: values, rebased row offsets (int32), global columns."""
    e0, e1 = int(rowptr[r0]), int(rowptr[r1])
    return val[e0:e1], (rowptr[r0:r1 + 1] - rowptr[r0]).astype(np.int32), col[e0:e1]


def half_bandwidth(rowptr_local, col, r0):
    """max |col - row| over the rows of a shard (the level's halo width is the max of this over the ranks)."""
    rows = np.repeat(np.arange(len(rowptr_local) - 1, dtype=np.int64) + r0, np.diff(rowptr_local))
    return int(np.abs(col.astype(np.int64) - rows).max()) if len(col) else 0


def ghost_widths(rank, world, halo):
    return (halo if rank > 0 else 0), (halo if rank < world - 1 else 0)


def shifted_view(ext, off, lo, n_glob):
    """A length-n_glob view whose entry [global j] is ext[j - (off - lo)] — what the kernels get by pointer shifting."""
    full = np.full(n_glob, np.nan)
    full[off - lo: off - lo + len(ext)] = ext
    return full


def rank_ordered_sum(partials):
    """Per-rank partial dots added in rank order (ConjugateGradientParallelGpu.cs:463 resultsDot.Sum())."""
    s = 0.0
    for t in partials:
        s += float(t)
    return s
