"""Host-side mirror of the reference's other two solver surfaces, above include/mgcg_formats.h.

* `SparseMatrixEll` + `ConjugateGradientEll`: the ELL container and solver object of the hand-written OpenCL variant
  (Mgcg/HandmadeCL/MgcgCL/SparseMatrix.cs:13-160, ConjugateGradient.cs, ConjugateGradientSingleGpu.cs:213-296).
* `ComputerGpu`: the Write / Solve / Read / Iteration object of the ViennaCL variant (Mgcg/ViennaCL/Mgcg/ComputerGpu.hpp:68-103).

Names and argument order follow the reference.  Nothing is computed on the host: containers only hold the arrays.
"""
import ctypes as C
import numpy as np

from . import _lib
from ._lib import Options, Result, MgcgError, check
from .solver import _p, default_options


class SparseMatrixEll:
    """SparseMatrix.cs:8-160 — Elements[i*Max + k], ColumnIndeces[i*Max + k], NonzeroCounts[i]; slot 0 is the diagonal."""

    def __init__(self, rowCount, maxNonzeroCountPerRow):
        self.RowCount = int(rowCount)
        self.MaxNonzeroCountPerRow = int(maxNonzeroCountPerRow)
        self.Elements = np.zeros(self.RowCount * self.MaxNonzeroCountPerRow, dtype=np.float64)
        self.ColumnIndeces = np.zeros(self.RowCount * self.MaxNonzeroCountPerRow, dtype=np.int32)
        self.NonzeroCounts = np.zeros(self.RowCount, dtype=np.int32)
        self.Clear()

    def Clear(self):
        """SparseMatrix.cs:52-63: every row keeps one entry, its (zero) diagonal."""
        self.NonzeroCounts[:] = 1
        self.Elements[::self.MaxNonzeroCountPerRow] = 0.0
        self.ColumnIndeces[::self.MaxNonzeroCountPerRow] = np.arange(self.RowCount, dtype=np.int32)

    def _local(self, i, j):          # GetLocalIndex, SparseMatrix.cs:165-181: slots 1.. of row i
        first = i * self.MaxNonzeroCountPerRow
        hit = np.nonzero(self.ColumnIndeces[first + 1:first + int(self.NonzeroCounts[i])] == j)[0]
        return int(hit[0]) + 1 if hit.size else -1

    def __getitem__(self, ij):
        i, j = ij
        if i == j:
            return float(self.Elements[i * self.MaxNonzeroCountPerRow])
        k = self._local(i, j)
        return float(self.Elements[i * self.MaxNonzeroCountPerRow + k]) if k >= 0 else 0.0

    def __setitem__(self, ij, value):
        i, j = ij
        first = i * self.MaxNonzeroCountPerRow
        if i == j:                                  # SparseMatrix.cs:82-91
            self.Elements[first] = value
            self.ColumnIndeces[first] = i
            return
        k = self._local(i, j)
        if k < 0:                                   # a new entry goes to the end of the row, SparseMatrix.cs:136-153
            if self.NonzeroCounts[i] == self.MaxNonzeroCountPerRow:
                raise IndexError(f"row {i}: more than {self.MaxNonzeroCountPerRow} entries")
            k = int(self.NonzeroCounts[i])
            self.ColumnIndeces[first + k] = j
            self.NonzeroCounts[i] += 1
        self.Elements[first + k] = value

    @classmethod
    def from_csr(cls, val, rowptr, col, maxNonzeroCountPerRow=None):
        """The matrix a caller would have built entry by entry from a CSR listing: the diagonal in slot 0 (zero when the row
        stores none), the other entries behind it in storage order.  Vectorised; used to make large ELL inputs."""
        val, rowptr, col = np.asarray(val, np.float64), np.asarray(rowptr, np.int64), np.asarray(col, np.int64)
        n = len(rowptr) - 1
        rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr))
        is_diag = col == rows
        off = ~is_diag
        cnt_off = np.bincount(rows[off], minlength=n)
        width = int(cnt_off.max()) + 1 if n else 1
        m = cls(n, width if maxNonzeroCountPerRow is None else maxNonzeroCountPerRow)
        if width > m.MaxNonzeroCountPerRow:
            raise IndexError(f"a row has {width} entries, more than {m.MaxNonzeroCountPerRow}")
        M = m.MaxNonzeroCountPerRow
        m.Elements[rows[is_diag] * M] = val[is_diag]
        start = np.concatenate([[0], np.cumsum(cnt_off)])[:-1]
        slot = np.arange(int(off.sum()), dtype=np.int64) - np.repeat(start, cnt_off) + 1      # storage order within the row
        m.Elements[rows[off] * M + slot] = val[off]
        m.ColumnIndeces[rows[off] * M + slot] = col[off]
        m.NonzeroCounts[:] = cnt_off + 1
        return m

    def to_csr(self):
        """CSR arrays with the slot order kept (what mgcg_ell_to_csr produces on the device) — index bookkeeping only."""
        M, cnt = self.MaxNonzeroCountPerRow, self.NonzeroCounts.astype(np.int64)
        rowptr = np.concatenate([[0], np.cumsum(cnt)])
        rows = np.repeat(np.arange(self.RowCount, dtype=np.int64), cnt)
        src = rows * M + (np.arange(int(rowptr[-1]), dtype=np.int64) - np.repeat(rowptr[:-1], cnt))
        return self.Elements[src].copy(), rowptr.astype(np.int32), self.ColumnIndeces[src].copy()


class ConjugateGradientEll:
    """The OpenCL variant's solver object: A (ELL), x, b members; Initialize() / Solve(); Iteration, Residual.
    ConjugateGradient.cs (HandmadeCL) ctor + IsConverged; ConjugateGradientSingleGpu.cs:213-296.  Stop rule max|r_i| < allowable."""

    def __init__(self, count, maxNonZeroCount, minIteration, maxIteration, allowableResidual, device=0):
        self.Count = count
        self.A = SparseMatrixEll(count, maxNonZeroCount)
        self.x = np.zeros(count)
        self.b = np.zeros(count)
        self.MinIteration, self.MaxIteration, self.AllowableResidual = minIteration, maxIteration, allowableResidual
        self.Iteration, self.Residual = 0, 0.0
        self.options = default_options(norm=_lib.NORM_ABS_LINF)
        self.result = None
        self.L = _lib.lib()
        self.ctx = self.L.mgcg_create(device)
        if not self.ctx:
            raise RuntimeError("mgcg_create failed: " + _lib.last_error())

    def Initialize(self):
        """The reference uploads here (ConjugateGradientSingleGpu.cs:213-220); the C entry point uploads inside Solve."""

    def Solve(self):
        o = self.options
        o.tolerance, o.min_iteration, o.max_iteration = self.AllowableResidual, self.MinIteration, self.MaxIteration
        res = Result()
        x = np.ascontiguousarray(self.x, dtype=np.float64)
        s = self.L.mgcg_solve_host_ell(self.ctx, self.Count, self.A.MaxNonzeroCountPerRow, _p(self.A.Elements), _p(self.A.ColumnIndeces),
                                       _p(self.A.NonzeroCounts), _p(np.ascontiguousarray(self.b, dtype=np.float64)), _p(x),
                                       C.byref(o), C.byref(res))
        self.result = res
        self.Iteration, self.Residual = res.iteration, res.residual
        if s == _lib.ERR_NOT_CONVERGED:             # ConjugateGradient.cs (HandmadeCL) IsConverged: throws past MaxIteration
            raise RuntimeError("pressure equation did not converge (Iteration > MaxIteration)")
        check(s, "mgcg_solve_host_ell")
        self.x = x

    def close(self):
        if self.ctx:
            self.L.mgcg_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ComputerGpu:
    """ComputerGpu.hpp:68-103: ComputerGpu(n); Write(elements, rowOffsets, columnIndeces, x, b); Solve(residual, minIteration,
    maxIteration); Read(x); Iteration().  CSR indices are unsigned like the reference's (ComputerGpu.hpp:47)."""

    def __init__(self, n, device=0):
        self.L = _lib.lib()
        self.n = n
        self.h = self.L.mgcg_computer_create(device, n)
        if not self.h:
            raise RuntimeError("mgcg_computer_create failed: " + _lib.last_error())

    def SetPreconditioner(self, precond, nx=0, ny=0, nz=0):
        check(self.L.mgcg_computer_set_preconditioner(self.h, precond, nx, ny, nz), "mgcg_computer_set_preconditioner")

    def Write(self, elements, rowOffsets, columnIndeces, x, b):
        e = np.ascontiguousarray(elements, dtype=np.float64)
        rp = np.ascontiguousarray(rowOffsets, dtype=np.uint32)
        ci = np.ascontiguousarray(columnIndeces, dtype=np.uint32)
        check(self.L.mgcg_computer_write(self.h, _p(e), _p(rp), _p(ci), len(e), _p(np.ascontiguousarray(x, dtype=np.float64)),
                                         _p(np.ascontiguousarray(b, dtype=np.float64))), "mgcg_computer_write")

    def Solve(self, residual, minIteration, maxIteration):
        check(self.L.mgcg_computer_solve(self.h, residual, minIteration, maxIteration), "mgcg_computer_solve")

    def Read(self, x):
        check(self.L.mgcg_computer_read(self.h, _p(x)), "mgcg_computer_read")

    def Iteration(self):
        return self.L.mgcg_computer_iteration(self.h)

    def Result(self):
        r = Result()
        check(self.L.mgcg_computer_result(self.h, C.byref(r)), "mgcg_computer_result")
        return r

    def close(self):
        if self.h:
            self.L.mgcg_computer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
