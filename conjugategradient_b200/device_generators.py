"""Stencil matrices assembled directly in device memory through the C ABI (mgcg_generate_*), one z-slab at a time.

Same matrices, bit for bit, as the host generators in generators.py (tests/test_gpu_generators.py); used where the host
copy would be too large or too slow to build: the 512^3 benches and the 27-point 512^3 shards.
"""
import ctypes as C

import numpy as np

from . import _lib
from .solver import _p, check


def poisson3d7(ctx, nx, ny=None, nz=None, zrange=None, device=None):
    """-> (elements fp64, row_offsets int32, column_indeces int32) torch tensors on `device` for planes zrange of the grid."""
    import torch
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    z0, z1 = (0, nz) if zrange is None else zrange
    L = _lib.lib()
    nnz = L.mgcg_generate_poisson3d7_nnz(nx, ny, nz, z0, z1)
    if nnz < 0:
        raise ValueError("bad grid / slab")
    rows = nx * ny * (z1 - z0)
    val = torch.empty(nnz, dtype=torch.float64, device=device)
    col = torch.empty(nnz, dtype=torch.int32, device=device)
    rp = torch.empty(rows + 1, dtype=torch.int32, device=device)
    check(L.mgcg_generate_poisson3d7(ctx, nx, ny, nz, z0, z1, _p(val), _p(rp), _p(col)), "mgcg_generate_poisson3d7")
    return val, rp, col


def varcoef_cells(nx, ny, nz, seed=3, block=8, contrast_decades=3.0):
    """Cell coefficients of generators.poisson3d27_varcoef: k = 10^U(-c, c) on block^3 cells (host, small)."""
    rng = np.random.default_rng(seed)
    bx, by, bz = -(-nx // block), -(-ny // block), -(-nz // block)
    return 10.0 ** rng.uniform(-contrast_decades, contrast_decades, size=bx * by * bz)


def poisson3d27_varcoef(ctx, nx, ny=None, nz=None, zrange=None, seed=3, block=8, contrast_decades=3.0, aniso=(1.0, 1.0, 1e-2), device=None):
    import torch
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    z0, z1 = (0, nz) if zrange is None else zrange
    L = _lib.lib()
    nnz = L.mgcg_generate_poisson3d27_nnz(nx, ny, nz, z0, z1)
    if nnz < 0:
        raise ValueError("bad grid / slab")
    rows = nx * ny * (z1 - z0)
    kblk = torch.as_tensor(varcoef_cells(nx, ny, nz, seed, block, contrast_decades)).to(device)
    val = torch.empty(nnz, dtype=torch.float64, device=device)
    col = torch.empty(nnz, dtype=torch.int32, device=device)
    rp = torch.empty(rows + 1, dtype=torch.int32, device=device)
    an = (C.c_double * 3)(*aniso)
    check(L.mgcg_generate_poisson3d27(ctx, nx, ny, nz, z0, z1, block, _p(kblk), an, _p(val), _p(rp), _p(col)), "mgcg_generate_poisson3d27")
    return val, rp, col


def mps_particles(ctx, pos, re, shift_abs=0.0, max_per_row=160):
    """Particle (MPS) pressure-Poisson matrix from positions `pos` (torch CUDA tensor n x 3, fp64, any order), assembled on the
    device by a cell-list neighbour search (mgcg_mps_order / _count / _fill; semantics of oracle/mps_ref.c).
    -> (elements, row_offsets, column_indeces, order, pos_sorted); order[k] = input index of the particle numbered k."""
    import torch
    L = _lib.lib()
    pos = pos.contiguous()
    n = pos.shape[0]
    order = torch.empty(n, dtype=torch.int32, device=pos.device)
    sorted_pos = torch.empty_like(pos)
    cells = L.mgcg_mps_order(ctx, n, _p(pos), re, _p(order), _p(sorted_pos))
    if cells <= 0:
        raise _lib.MgcgError(int(-cells), "mgcg_mps_order")
    rp = torch.empty(n + 1, dtype=torch.int32, device=pos.device)
    nnz = L.mgcg_mps_count(ctx, n, _p(sorted_pos), re, max_per_row, _p(rp))
    if nnz < 0:
        raise _lib.MgcgError(int(-nnz), "mgcg_mps_count")
    val = torch.empty(nnz, dtype=torch.float64, device=pos.device)
    col = torch.empty(nnz, dtype=torch.int32, device=pos.device)
    check(L.mgcg_mps_fill(ctx, n, _p(sorted_pos), re, shift_abs, _p(rp), _p(val), _p(col)), "mgcg_mps_fill")
    return val, rp, col, order, sorted_pos
