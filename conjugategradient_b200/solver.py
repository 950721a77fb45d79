"""Host-side mirror of the reference's solver classes, above the C ABI of libMgcgGpu.so.

Class and member names follow the reference (Mgcg/cuBlas/Mgcg/*.cs) so the parity tests read like its harness
(MgcgMain.cs:41-178): construct -> set A, b, x -> Initialize() -> Solve() -> Read() -> compare.
`Mgcg` is the native (non-compat) API on torch CUDA tensors: torch only owns device memory and streams here.
"""
import ctypes as C
import numpy as np

from . import _lib
from ._lib import Options, Result, MgcgError, check


def _p(a):
    """void* of a numpy array or a torch tensor."""
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(a.data_ptr())


def default_options(**kw):
    o = Options()
    _lib.lib().mgcg_options_default(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(f"mgcg_options has no field {k}")
        setattr(o, k, v)
    return o


class SparseMatrix:
    """CSR container — SparseMatrix.cs:8-101 (Elements / ColumnIndeces / RowOffsets)."""

    def __init__(self, Elements, RowOffsets, ColumnIndeces):
        self.Elements = np.ascontiguousarray(Elements, dtype=np.float64)
        self.RowOffsets = np.ascontiguousarray(RowOffsets, dtype=np.int32)
        self.ColumnIndeces = np.ascontiguousarray(ColumnIndeces, dtype=np.int32)


class ConjugateGradient:
    """ConjugateGradient.cs:6-84 + LinerEquations.cs:6-47."""

    def __init__(self, count, maxNonZeroCount, minIteration, maxIteration, allowableResidual):
        self.Count = count
        self.MaxNonZeroCount = maxNonZeroCount
        self.MinIteration = minIteration
        self.MaxIteration = maxIteration
        self.AllowableResidual = allowableResidual
        self.A = None
        self.x = np.zeros(count)
        self.b = np.zeros(count)
        self.Iteration = 0
        self.Residual = 0.0


class ConjugateGradientSingleGpu(ConjugateGradient):
    """ConjugateGradientSingleGpu.cs:98-178 on the compat exports (CreateBlas.., Create_Double.., Solve)."""

    def __init__(self, count, maxNonZeroCount, minIteration, maxIteration, allowableResidual):
        super().__init__(count, maxNonZeroCount, minIteration, maxIteration, allowableResidual)
        L = self.L = _lib.lib()
        self.cublas, self.cusparse, self.matDescr = L.CreateBlas(), L.CreateSparse(), L.CreateMatDescr()
        self._vecs = {}

    def Initialize(self):
        L, A = self.L, self.A
        nnz = int(A.RowOffsets[self.Count])
        v = self._vecs
        for k in list(v):
            (L.Delete_Int if k in ("col", "rp") else L.Delete_Double)(v.pop(k))
        # the reference allocates count*maxNonZeroCount (cs:107-108); the actual non-zero count is enough
        v["A"], v["col"], v["rp"] = L.Create_Double(nnz), L.Create_Int(nnz), L.Create_Int(self.Count + 1)
        for k in ("x", "b", "Ap", "p", "r"):
            v[k] = L.Create_Double(self.Count)
        L.CopyFromArray_Double(v["A"], _p(A.Elements), nnz, 0, 0)
        L.CopyFromArray_Int(v["col"], _p(A.ColumnIndeces), nnz, 0, 0)
        L.CopyFromArray_Int(v["rp"], _p(A.RowOffsets), self.Count + 1, 0, 0)
        L.CopyFromArray_Double(v["b"], _p(np.ascontiguousarray(self.b)), self.Count, 0, 0)
        L.CopyFromArray_Double(v["x"], _p(np.ascontiguousarray(self.x)), self.Count, 0, 0)

    def Solve(self):
        L, v = self.L, self._vecs
        nnz = int(self.A.RowOffsets[self.Count])
        it, res = C.c_int(0), C.c_double(0)
        L.Solve(self.cublas, self.cusparse, self.matDescr, v["A"], v["rp"], v["col"], v["x"], v["b"], v["Ap"], v["p"], v["r"],
                nnz, self.Count, self.AllowableResidual, self.MinIteration, self.MaxIteration, C.byref(it), C.byref(res))
        self.Iteration = it.value - 1            # cs:168
        self.Residual = res.value
        if self.Iteration > self.MaxIteration:   # ConjugateGradient.cs:70-74
            raise RuntimeError("pressure equation did not converge (Iteration > MaxIteration)")

    def Read(self):
        self.x = np.empty(self.Count)
        self.L.CopyToArray_Double(self._vecs["x"], _p(self.x), self.Count, 0, 0)

    def close(self):
        L, v = self.L, self._vecs
        for k in list(v):
            (L.Delete_Int if k in ("col", "rp") else L.Delete_Double)(v.pop(k))
        if self.cublas:
            L.DestroyBlas(self.cublas); L.DestroySparse(self.cusparse); L.DestroyMatDescr(self.matDescr)
            self.cublas = None


class ConjugateGradientParallelGpu(ConjugateGradient):
    """ConjugateGradientParallelGpu.cs:264-594: row-block split over `deviceCount` devices, host-staged halo exchange of p
    (SyncP), split-phase Solve0..3 and host sums in device order — all through the compat exports.
    `devices` lets a test map several logical blocks onto the GPUs that exist (block d runs on device d % GetDeviceCount())."""

    def __init__(self, count, maxNonZeroCount, minIteration, maxIteration, allowableResidual, devices=None):
        super().__init__(count, maxNonZeroCount, minIteration, maxIteration, allowableResidual)
        L = self.L = _lib.lib()
        self.physical = L.GetDeviceCount()
        if self.physical <= 0:
            raise RuntimeError("no CUDA device (no CPU fallback)")
        self.deviceCount = devices if devices else self.physical
        off = np.zeros(self.deviceCount + 1, dtype=np.int32)
        check(L.mgcg_partition_offsets(count, self.deviceCount, _p(off)), "mgcg_partition_offsets")   # cs:271-277
        self.offsetsForDevice = off
        self.minJ = [0] * self.deviceCount
        self.maxJ = [0] * self.deviceCount
        self.bufferHost = np.zeros(count)
        self.h, self.v = [], []
        for d in range(self.deviceCount):
            L.SetDevice(d % self.physical)
            self.h.append((L.CreateBlas(), L.CreateSparse(), L.CreateMatDescr()))
            self.v.append({})

    def CountForDevice(self, d):
        return int(self.offsetsForDevice[d + 1] - self.offsetsForDevice[d]) if 0 <= d < self.deviceCount else 0

    def Initialize(self):
        L, A = self.L, self.A
        for d in range(self.deviceCount):
            L.SetDevice(d % self.physical)
            cnt, off = self.CountForDevice(d), int(self.offsetsForDevice[d])
            eoff = int(A.RowOffsets[off])
            ecnt = int(A.RowOffsets[self.offsetsForDevice[d + 1]]) - eoff
            v = self.v[d]
            v["A"], v["col"], v["rp"] = L.Create_Double(ecnt), L.Create_Int(ecnt), L.Create_Int(cnt + 1)
            for k in ("x", "b", "Ap", "r"):
                v[k] = L.Create_Double(cnt)
            v["p"] = L.Create_Double(self.Count)      # global-indexed full-length p (cs:321)
            v["ecnt"] = ecnt
            mn, mx = C.c_int(0), C.c_int(0)
            L.Initialize(_p(A.Elements), _p(A.RowOffsets), _p(A.ColumnIndeces), _p(np.ascontiguousarray(self.x)),
                         _p(np.ascontiguousarray(self.b)), v["A"], v["rp"], v["col"], v["x"], v["b"], v["p"],
                         C.byref(mn), C.byref(mx), self.Count, cnt, off, ecnt, eoff)
            self.minJ[d], self.maxJ[d] = mn.value, mx.value

    def _halo(self, d):
        cnt, off = self.CountForDevice(d), int(self.offsetsForDevice[d])
        last = off - self.minJ[d] if d > 0 else 0                                   # cs:397
        nxt = self.maxJ[d] - cnt - off + 1 if d < self.deviceCount - 1 else 0       # cs:398
        return cnt, off, last, nxt

    def SyncP(self):
        L = self.L
        for d in range(self.deviceCount):
            L.SetDevice(d % self.physical)
            cnt, off, last, nxt = self._halo(d)
            L.P2Host(self.v[d]["p"], _p(self.bufferHost), cnt, off, last, nxt)
        for d in range(self.deviceCount):
            L.SetDevice(d % self.physical)
            cnt, off, last, nxt = self._halo(d)
            L.P2Device(self.v[d]["p"], _p(self.bufferHost), cnt, off, last, nxt)

    def Solve(self):
        L, D = self.L, self.deviceCount
        dots = [0.0] * D
        self.SyncP()
        for d in range(D):
            L.SetDevice(d % self.physical)
            v, (bl, sp, md) = self.v[d], self.h[d]
            dots[d] = L.Solve0(bl, sp, md, v["A"], v["rp"], v["col"], v["x"], v["b"], v["Ap"], v["p"], v["r"],
                               self.Count, self.CountForDevice(d), int(self.offsetsForDevice[d]), v["ecnt"])
        rr = 0.0
        for t in dots:
            rr += t                                                                 # cs:463 resultsDot.Sum()
        self.Iteration = 0
        while True:
            self.SyncP()
            for d in range(D):
                L.SetDevice(d % self.physical)
                v, (bl, sp, md) = self.v[d], self.h[d]
                dots[d] = L.Solve1(bl, sp, md, v["A"], v["rp"], v["col"], v["Ap"], v["p"],
                                   self.Count, self.CountForDevice(d), int(self.offsetsForDevice[d]), v["ecnt"])
            pAp = 0.0
            for t in dots:
                pAp += t
            alpha = rr / pAp
            for d in range(D):
                L.SetDevice(d % self.physical)
                v, (bl, sp, md) = self.v[d], self.h[d]
                dots[d] = L.Solve2(bl, alpha, v["x"], v["Ap"], v["p"], v["r"], self.CountForDevice(d), int(self.offsetsForDevice[d]))
            rrNew = 0.0
            for t in dots:
                rrNew += t
            self.Residual = float(np.sqrt(rrNew))
            if self.Iteration >= self.MinIteration:                                 # ConjugateGradient.cs:64-77
                if self.Iteration > self.MaxIteration:
                    raise RuntimeError("pressure equation did not converge (Iteration > MaxIteration)")
                if self.Residual < self.AllowableResidual:
                    break
            beta = rrNew / rr
            for d in range(D):
                L.SetDevice(d % self.physical)
                v, (bl, sp, md) = self.v[d], self.h[d]
                L.Solve3(bl, beta, v["p"], v["r"], self.CountForDevice(d), int(self.offsetsForDevice[d]))
            rr = rrNew
            self.Iteration += 1

    def Read(self):
        self.x = np.empty(self.Count)
        for d in range(self.deviceCount):
            self.L.SetDevice(d % self.physical)
            self.L.CopyToArray_Double(self.v[d]["x"], _p(self.x), self.CountForDevice(d), 0, int(self.offsetsForDevice[d]))  # cs:581

    def close(self):
        L = self.L
        for d in range(self.deviceCount):
            L.SetDevice(d % self.physical)
            v = self.v[d]
            for k in list(v):
                if k == "ecnt":
                    v.pop(k); continue
                (L.Delete_Int if k in ("col", "rp") else L.Delete_Double)(v.pop(k))
            bl, sp, md = self.h[d]
            L.DestroyBlas(bl); L.DestroySparse(sp); L.DestroyMatDescr(md)
        self.h = []
        L.SetDevice(0)


def SolveGpu(elements, rowOffsets, columnIndeces, b, allowableResidual, minIteration, maxIteration, device="cuda:0"):
    """SimpleConjugateGradient.cu:12-126 through SolveGpu_c: x0 = 0, strict min-iteration.  Returns (x, iteration, residual)."""
    import torch
    L = _lib.lib()
    dev = torch.device(device)
    L.SetDevice(dev.index or 0)
    tv = torch.as_tensor(elements, dtype=torch.float64).to(dev)
    tr = torch.as_tensor(rowOffsets, dtype=torch.int32).to(dev)
    tc = torch.as_tensor(columnIndeces, dtype=torch.int32).to(dev)
    tb = torch.as_tensor(b, dtype=torch.float64).to(dev)
    tx = torch.empty_like(tb)
    torch.cuda.synchronize(dev)
    it, res = C.c_int(0), C.c_double(0)
    L.SolveGpu_c(_p(tv), _p(tr), _p(tc), _p(tx), _p(tb), tv.numel(), tb.numel(), allowableResidual, minIteration, maxIteration,
                 C.byref(it), C.byref(res))
    return tx.cpu().numpy(), it.value, res.value


class Mgcg:
    """Native API on device memory: setup once (SpMV plan + multigrid hierarchy on the device), solve many times."""

    def __init__(self, device=0, use_torch_stream=False):
        self.L = _lib.lib()
        self.ctx = self.L.mgcg_create(device)
        if not self.ctx:
            raise RuntimeError("mgcg_create failed: " + _lib.last_error())
        self.device = device
        self._keep = None
        if use_torch_stream:
            import torch
            check(self.L.mgcg_set_stream(self.ctx, C.c_void_p(torch.cuda.current_stream(device).cuda_stream)), "mgcg_set_stream")

    def setup(self, val, rowptr, col, options=None):
        """val/rowptr/col: torch CUDA tensors (fp64/int32/int32); kept alive by this object."""
        self._keep = (val, rowptr, col)
        self.n, self.nnz = rowptr.numel() - 1, val.numel()
        self.options = options if options is not None else default_options()
        check(self.L.mgcg_setup(self.ctx, self.n, self.nnz, _p(val), _p(rowptr), _p(col), C.byref(self.options)), "mgcg_setup")

    def solve(self, b, x, raise_on_fail=True):
        res = Result()
        s = self.L.mgcg_solve(self.ctx, _p(b), _p(x), C.byref(res))
        if s != _lib.OK and raise_on_fail:
            raise MgcgError(s, "mgcg_solve")
        return res

    def solve_host(self, val, rowptr, col, b, x, options=None, raise_on_fail=True):
        """numpy (ideally pinned) host arrays in, x updated in place: upload + setup + solve + download."""
        res = Result()
        o = options if options is not None else default_options()
        s = self.L.mgcg_solve_host(self.ctx, len(rowptr) - 1, len(val), _p(val), _p(rowptr), _p(col), _p(b), _p(x), C.byref(o), C.byref(res))
        if s != _lib.OK and raise_on_fail:
            raise MgcgError(s, "mgcg_solve_host")
        return res

    def levels(self):
        return self.L.mgcg_mg_levels(self.ctx)

    def level(self, l):
        rows, nnz = C.c_int(0), C.c_int(0)
        check(self.L.mgcg_mg_level_size(self.ctx, l, C.byref(rows), C.byref(nnz)), "mgcg_mg_level_size")
        val, rp, col = np.empty(nnz.value), np.empty(rows.value + 1, dtype=np.int32), np.empty(nnz.value, dtype=np.int32)
        check(self.L.mgcg_mg_level_get(self.ctx, l, _p(val), _p(rp), _p(col)), "mgcg_mg_level_get")
        return val, rp, col

    def aggregates(self, l):
        rows, nnz = C.c_int(0), C.c_int(0)
        check(self.L.mgcg_mg_level_size(self.ctx, l, C.byref(rows), C.byref(nnz)), "mgcg_mg_level_size")
        agg = np.empty(rows.value, dtype=np.int32)
        check(self.L.mgcg_mg_level_aggregates(self.ctx, l, _p(agg)), "mgcg_mg_level_aggregates")
        return agg

    def mg_apply(self, r, z):
        check(self.L.mgcg_mg_apply(self.ctx, _p(r), _p(z)), "mgcg_mg_apply")

    def matrix_codes(self):
        """Coded columns of the set-up matrix (csrc/spmv_code.cu): (dict, codes) with col[k] == row(k) + dict[codes[k]], or None when
        the matrix is not coded."""
        nd = C.c_int(0)
        check(self.L.mgcg_matrix_codes(self.ctx, C.byref(nd), None, None), "mgcg_matrix_codes")
        if nd.value == 0:
            return None
        d, codes = np.zeros(256, dtype=np.int32), np.empty(self.nnz, dtype=np.uint8)
        check(self.L.mgcg_matrix_codes(self.ctx, C.byref(nd), _p(d), _p(codes)), "mgcg_matrix_codes")
        return d[:nd.value].copy(), codes

    def spmv(self, x, y, alpha=1.0, beta=0.0):
        """y = alpha A x + beta y with the set-up matrix (runs on its coded columns when it has them)."""
        val, rowptr, col = self._keep
        check(self.L.mgcg_spmv(self.ctx, self.n, self.n, self.nnz, _p(val), _p(rowptr), _p(col), _p(x), _p(y), alpha, beta,
                               self.options.spmv_kernel), "mgcg_spmv")

    def close(self):
        if self.ctx:
            self.L.mgcg_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MgcgDist:
    """One-process-per-GPU sharded solve (mgcg_comm_init / mgcg_setup_dist / mgcg_solve_dist): the native replacement of
    ConjugateGradientParallelGpu (cs:264-594).  Rank r owns rows [row_offset, row_offset + local_count) with GLOBAL column
    indices and row offsets rebased to 0 (Mgcg.cu:73).  The 128-byte communicator id is created on rank 0 (`unique_id`)
    and handed to the other ranks out of band (e.g. torch.distributed.broadcast)."""

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        check(_lib.lib().mgcg_comm_unique_id(buf), "mgcg_comm_unique_id")
        return buf.raw

    def __init__(self, device, rank, world, id_bytes, use_torch_stream=False):
        self.L = _lib.lib()
        self.ctx = self.L.mgcg_create(device)
        if not self.ctx:
            raise RuntimeError("mgcg_create failed: " + _lib.last_error())
        self.rank, self.world = rank, world
        if use_torch_stream:
            import torch
            check(self.L.mgcg_set_stream(self.ctx, C.c_void_p(torch.cuda.current_stream(device).cuda_stream)), "mgcg_set_stream")
        check(self.L.mgcg_comm_init(self.ctx, C.c_char_p(id_bytes), rank, world), "mgcg_comm_init")
        self._keep = None

    def setup(self, global_count, row_offset, val, rowptr, col, options=None):
        self._keep = (val, rowptr, col)
        self.options = options if options is not None else default_options()
        check(self.L.mgcg_setup_dist(self.ctx, global_count, row_offset, rowptr.numel() - 1, val.numel(), _p(val), _p(rowptr), _p(col),
                                     C.byref(self.options)), "mgcg_setup_dist")

    def solve(self, b_local, x_local, raise_on_fail=True):
        res = Result()
        s = self.L.mgcg_solve_dist(self.ctx, _p(b_local), _p(x_local), C.byref(res))
        if s != _lib.OK and raise_on_fail:
            raise MgcgError(s, "mgcg_solve_dist")
        return res

    def solve_host(self, global_count, row_offset, val, rowptr, col, b_local, x_local, options=None, raise_on_fail=True):
        """numpy (ideally pinned) host arrays of this rank's row block in, x_local updated in place: one C-ABI call
        (mgcg_solve_host_dist) doing upload + setup + solve + download."""
        res = Result()
        o = options if options is not None else default_options()
        s = self.L.mgcg_solve_host_dist(self.ctx, global_count, row_offset, len(rowptr) - 1, len(val), _p(val), _p(rowptr), _p(col),
                                        _p(b_local), _p(x_local), C.byref(o), C.byref(res))
        if s != _lib.OK and raise_on_fail:
            raise MgcgError(s, "mgcg_solve_host_dist")
        return res

    def close(self):
        if self.ctx:
            self.L.mgcg_comm_finalize(self.ctx)     # collective: unmap peers, rendezvous, free (every rank closes together)
            self.L.mgcg_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
