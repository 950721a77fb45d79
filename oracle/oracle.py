"""ctypes loader for the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
c_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")

NORM_ABS_L2, NORM_ABS_LINF, NORM_REL_L2 = 0, 1, 2
OK, MAX_ITERATION_EXCEEDED = 0, 3


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("cg_ref.c", "mg_ref.c", "cg_ref_mt.c", "mps_ref.c") if os.path.exists(os.path.join(_HERE, f))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.ref_spmv.argtypes = [C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp]
        L.ref_dot.argtypes = [C.c_int, c_dp, c_dp]; L.ref_dot.restype = C.c_double
        L.ref_max_absolute.argtypes = [C.c_int, c_dp]; L.ref_max_absolute.restype = C.c_double
        L.ref_set_added.argtypes = [C.c_int, c_dp, c_dp, c_dp, C.c_double]
        L.ref_cg_solve.argtypes = [C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp, c_dp, c_dp, c_dp,
                                   C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.ref_cg_solve.restype = C.c_int
        L.ref_partition_offsets.argtypes = [C.c_int, C.c_int, c_ip]
        L.ref_initialize_block.argtypes = [c_ip, c_ip, C.c_int, C.c_int, c_ip] + [C.POINTER(C.c_int)] * 4
        L.ref_halo_counts.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int)] * 2
        L.ref_cg_solve_parallel.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp,
                                            C.c_double, C.c_int, C.c_int,
                                            C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.ref_cg_solve_parallel.restype = C.c_int
        L.ref_gen_tridiag121.argtypes = [C.c_int, c_dp, c_ip, c_ip, c_dp]; L.ref_gen_tridiag121.restype = C.c_int
        L.ref_gen_banded_sin.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp]
        L.ref_gen_banded_sin.restype = C.c_int
        L.ref_cg_solve_mt.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp, c_dp, c_dp, c_dp,
                                      C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.ref_cg_solve_mt.restype = C.c_int
        L.ref_mt_max_threads.restype = C.c_int
        L.ref_mps_cell_order.argtypes = [C.c_int, c_dp, C.c_double, c_ip, C.c_void_p, C.c_int]; L.ref_mps_cell_order.restype = C.c_int
        L.ref_mps_assemble.argtypes = [C.c_int, c_dp, C.c_double, C.c_double, C.c_int, c_ip, C.c_void_p, C.c_void_p]
        L.ref_mps_assemble.restype = C.c_longlong
        L.ref_ell_multiply.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp]
        L.ref_ell_to_csr.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_ip, c_ip]; L.ref_ell_to_csr.restype = C.c_int
        L.ref_cg_solve_ell.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, c_dp, c_dp, c_dp, c_dp, c_dp,
                                       C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.ref_cg_solve_ell.restype = C.c_int
        _LIB = L
    return _LIB


def spmv(val, rowptr, col, x):
    n = len(rowptr) - 1
    y = np.empty(n)
    lib().ref_spmv(n, val, rowptr, col, x, y)
    return y


def dot(a, b):
    return lib().ref_dot(len(a), a, b)


def coded_columns(rowptr, col, max_codes=256):
    """Checker for the library's coded-column layout (csrc/spmv_code.cu; the reference has only int32 CSR columns, Mgcg.cu:10-19):
    the sorted distinct distances col - row and each entry's rank among them, or None when there are more than `max_codes`."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    row_of = np.repeat(np.arange(len(rowptr) - 1, dtype=np.int64), np.diff(rowptr))
    d = np.asarray(col, dtype=np.int64) - row_of
    dist = np.unique(d)
    if len(dist) > max_codes:
        return None
    return dist.astype(np.int32), np.searchsorted(dist, d).astype(np.uint8)


def coded_runs(dist, span=6, max_runs=12):
    """Checker for the run grouping of the coded distances (csrc/spmv_code.cu k_code_runs): a run takes sorted distances while they
    stay within `span` of its first one.  -> (first distance of every run, run index of every distance) or None beyond `max_runs`."""
    firsts, run_of = [], []
    for d in (int(v) for v in dist):
        if not firsts or d - firsts[-1] > span:
            firsts.append(d)
        run_of.append(len(firsts) - 1)
    if len(firsts) > max_runs:
        return None
    return np.array(firsts, dtype=np.int64), np.array(run_of, dtype=np.int64)


def cg_solve(val, rowptr, col, b, x0=None, tol=1e-8, min_iteration=0, max_iteration=None,
             norm=NORM_ABS_L2, min_strict=False):
    """ConjugateGradientCpu.Solve restatement. Returns (x, iteration_index, residual, status)."""
    n = len(rowptr) - 1
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
    r, p, Ap = np.empty(n), np.empty(n), np.empty(n)
    it, res = C.c_int(0), C.c_double(0)
    st = lib().ref_cg_solve(n, val, rowptr, col, x, np.ascontiguousarray(b, dtype=np.float64), r, p, Ap, tol,
                            min_iteration, n if max_iteration is None else max_iteration,
                            norm, int(min_strict), C.byref(it), C.byref(res))
    return x, it.value, res.value, st


def cg_solve_mt(val, rowptr, col, b, x0=None, tol=1e-8, min_iteration=0, max_iteration=None, norm=NORM_ABS_L2, threads=0):
    """cg_ref_mt.c: the same CG with the reference's commented-out Parallel.For loops run over OpenMP threads."""
    n = len(rowptr) - 1
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
    r, p, Ap = np.empty(n), np.empty(n), np.empty(n)
    it, res = C.c_int(0), C.c_double(0)
    st = lib().ref_cg_solve_mt(threads, n, val, rowptr, col, x, np.ascontiguousarray(b, dtype=np.float64), r, p, Ap, tol,
                               min_iteration, n if max_iteration is None else max_iteration, norm, C.byref(it), C.byref(res))
    return x, it.value, res.value, st


def ell_multiply(elements, columns, counts, max_nonzero, x):
    """SparseMatrix.Multiply of the OpenCL variant (HandmadeCL/MgcgCL/SparseMatrix.cs:202-224)."""
    n = len(counts)
    y = np.empty(n)
    lib().ref_ell_multiply(n, max_nonzero, elements, columns, counts, np.ascontiguousarray(x, dtype=np.float64), y)
    return y


def ell_to_csr(elements, columns, counts, max_nonzero):
    n = len(counts)
    total = int(np.sum(counts))
    val, rp, col = np.empty(total), np.empty(n + 1, dtype=np.int32), np.empty(total, dtype=np.int32)
    assert lib().ref_ell_to_csr(n, max_nonzero, elements, columns, counts, val, rp, col) == total
    return val, rp, col


def cg_solve_ell(elements, columns, counts, max_nonzero, b, x0=None, tol=1e-8, min_iteration=0, max_iteration=None):
    """ConjugateGradientSingleGpu.Solve of the OpenCL variant (ELL matrix, max|r| stop). Returns (x, iteration_index, residual, status)."""
    n = len(counts)
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
    r, p, Ap = np.empty(n), np.empty(n), np.empty(n)
    it, res = C.c_int(0), C.c_double(0)
    st = lib().ref_cg_solve_ell(n, max_nonzero, elements, columns, counts, x, np.ascontiguousarray(b, dtype=np.float64), r, p, Ap, tol,
                                min_iteration, n if max_iteration is None else max_iteration, C.byref(it), C.byref(res))
    return x, it.value, res.value, st


def mps_assemble(pos, re, shift_abs=0.0, max_per_row=160):
    """oracle/mps_ref.c: cell-list neighbour search -> CSR (diagonal first, neighbours ascending).  pos: (n, 3) in any order.
    Returns (val, rowptr, col, order) with order[k] = input index of the particle numbered k."""
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    n = pos.shape[0]
    order = np.empty(n, dtype=np.int32)
    assert lib().ref_mps_cell_order(n, pos.reshape(-1), re, order, None, 0) > 0
    p = np.ascontiguousarray(pos[order]).reshape(-1)
    rp = np.empty(n + 1, dtype=np.int32)
    nnz = lib().ref_mps_assemble(n, p, re, shift_abs, max_per_row, rp, None, None)
    if nnz < 0:
        raise ValueError(f"ref_mps_assemble: {nnz}")
    val, col = np.empty(nnz), np.empty(nnz, dtype=np.int32)
    assert lib().ref_mps_assemble(n, p, re, shift_abs, max_per_row, rp, val.ctypes.data_as(C.c_void_p), col.ctypes.data_as(C.c_void_p)) == nnz
    return val, rp, col, order


def max_threads():
    return lib().ref_mt_max_threads()


def cg_solve_parallel(val, rowptr, col, b, devices, x0=None, tol=1e-8, min_iteration=0, max_iteration=None):
    n = len(rowptr) - 1
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
    it, res = C.c_int(0), C.c_double(0)
    st = lib().ref_cg_solve_parallel(n, devices, val, rowptr, col, x, np.ascontiguousarray(b, dtype=np.float64), tol,
                                     min_iteration, n if max_iteration is None else max_iteration,
                                     C.byref(it), C.byref(res))
    return x, it.value, res.value, st


def partition_offsets(count, devices):
    off = np.zeros(devices + 1, dtype=np.int32)
    lib().ref_partition_offsets(count, devices, off)
    return off


def initialize_block(rowptr, col, offset, count):
    lro = np.zeros(count + 1, dtype=np.int32)
    eo, ec, mn, mx = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    lib().ref_initialize_block(rowptr, col, offset, count, lro, C.byref(eo), C.byref(ec), C.byref(mn), C.byref(mx))
    return lro, eo.value, ec.value, mn.value, mx.value


def halo_counts(dev, devices, offset, count, min_j, max_j):
    lc, nc = C.c_int(), C.c_int()
    lib().ref_halo_counts(dev, devices, offset, count, min_j, max_j, C.byref(lc), C.byref(nc))
    return lc.value, nc.value


def gen_tridiag121(n):
    val, col, rp, b = np.zeros(3 * n), np.zeros(3 * n, dtype=np.int32), np.zeros(n + 1, dtype=np.int32), np.zeros(n)
    nnz = lib().ref_gen_tridiag121(n, val, col, rp, b)
    return val[:nnz].copy(), rp, col[:nnz].copy(), b


def gen_banded_sin(n, max_nonzero=160):
    val, col = np.zeros(n * max_nonzero), np.zeros(n * max_nonzero, dtype=np.int32)
    rp, b, x0 = np.zeros(n + 1, dtype=np.int32), np.zeros(n), np.zeros(n)
    nnz = lib().ref_gen_banded_sin(n, max_nonzero, val, col, rp, b, x0)
    return val[:nnz].copy(), rp, col[:nnz].copy(), b, x0


# ---- multigrid CPU twin (oracle/mg_ref.c; parity unpinned: the reference has no multigrid) ------------------
class MgRef:
    def __init__(self, val, rowptr, col, nx=0, ny=0, nz=0, max_levels=12, coarse_size=1400, smoothed=1, prolong_omega=2.0 / 3.0,
                 smoothed_levels=0, agg_size=3, agg_size_coarse=2, aggregation=0, theta=0.5):
        L = lib()
        L.mg_ref_build.restype = C.c_void_p
        L.mg_ref_build.argtypes = [C.c_int, C.c_int, c_dp, c_ip, c_ip, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double]
        L.mg_ref_free.argtypes = [C.c_void_p]
        L.mg_ref_levels.argtypes = [C.c_void_p]
        L.mg_ref_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.mg_ref_level_get.argtypes = [C.c_void_p, C.c_int, c_dp, c_ip, c_ip]
        L.mg_ref_level_agg.argtypes = [C.c_void_p, C.c_int, c_ip]
        L.mg_ref_apply.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, c_dp, c_dp]
        L.mg_ref_pcg.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, c_dp, c_dp, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.POINTER(C.c_int), C.POINTER(C.c_double)]
        self._keep = (np.ascontiguousarray(val, dtype=np.float64), np.ascontiguousarray(rowptr, dtype=np.int32),
                      np.ascontiguousarray(col, dtype=np.int32))
        self.n = len(rowptr) - 1
        self.h = L.mg_ref_build(self.n, len(val), *self._keep, nx, ny, nz, max_levels, coarse_size, smoothed, prolong_omega, smoothed_levels, agg_size, agg_size_coarse, aggregation, theta)

    def levels(self):
        return lib().mg_ref_levels(self.h)

    def set_smoother(self, smoother):
        """0: weighted Jacobi everywhere; 1: red-black Gauss-Seidel on the two-colourable grid levels (returns their flags)."""
        L = lib()
        L.mg_ref_set_smoother.argtypes = [C.c_void_p, C.c_int]
        L.mg_ref_level_rb.argtypes = [C.c_void_p, C.c_int]
        L.mg_ref_set_smoother(self.h, smoother)
        return [bool(L.mg_ref_level_rb(self.h, l)) for l in range(self.levels())]

    def set_chebyshev(self, degree, ratio=4.0):
        """Chebyshev polynomial smoother of `degree` in D^-1 A on [lmax/ratio, lmax], lmax = Gershgorin bound (0: off).  Returns lmax per level."""
        L = lib()
        L.mg_ref_set_chebyshev.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.mg_ref_level_lmax.argtypes = [C.c_void_p, C.c_int]
        L.mg_ref_level_lmax.restype = C.c_double
        L.mg_ref_set_chebyshev(self.h, degree, ratio)
        return [L.mg_ref_level_lmax(self.h, l) for l in range(self.levels())]

    def level(self, l):
        rows, nnz = C.c_int(), C.c_int()
        lib().mg_ref_level_size(self.h, l, C.byref(rows), C.byref(nnz))
        val, rp, col = np.empty(nnz.value), np.empty(rows.value + 1, dtype=np.int32), np.empty(nnz.value, dtype=np.int32)
        lib().mg_ref_level_get(self.h, l, val, rp, col)
        return val, rp, col

    def aggregates(self, l):
        rows, nnz = C.c_int(), C.c_int()
        lib().mg_ref_level_size(self.h, l, C.byref(rows), C.byref(nnz))
        agg = np.empty(rows.value, dtype=np.int32)
        lib().mg_ref_level_agg(self.h, l, agg)
        return agg

    def apply(self, r, pre=1, post=1, omega=0.85):
        z = np.empty(self.n)
        lib().mg_ref_apply(self.h, pre, post, omega, np.ascontiguousarray(r, dtype=np.float64), z)
        return z

    def pcg(self, b, x0=None, tol=1e-8, min_iteration=0, max_iteration=None, norm=NORM_ABS_L2, min_strict=False, pre=1, post=1, omega=0.85):
        x = np.zeros(self.n) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
        it, res = C.c_int(0), C.c_double(0)
        st = lib().mg_ref_pcg(self.h, pre, post, omega, np.ascontiguousarray(b, dtype=np.float64), x, tol, min_iteration,
                              self.n if max_iteration is None else max_iteration, norm, int(min_strict), C.byref(it), C.byref(res))
        return x, it.value, res.value, st

    def close(self):
        if self.h:
            lib().mg_ref_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
