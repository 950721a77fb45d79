"""ctypes binding of conjugategradient_b200/lib/libMgcgGpu.so (the C ABI of include/mgcg.h + mgcg_compat.h).

There is NO CPU fallback: if the shared library is missing the import of any product module raises.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libMgcgGpu.so")
CSRC = os.path.join(_HERE, "csrc")

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_CONVERGED, ERR_BREAKDOWN, ERR_COMM, ERR_UNSUPPORTED = range(7)
NORM_ABS_L2, NORM_ABS_LINF, NORM_REL_L2 = 0, 1, 2
PRECOND_NONE, PRECOND_MG = 0, 1
SPMV_AUTO, SPMV_STREAM, SPMV_VECTOR, SPMV_ROWS = 0, 1, 2, 3
TRANSPORT_AUTO, TRANSPORT_NCCL, TRANSPORT_PEER = 0, 1, 2
AGG_AUTO, AGG_BLOCKS, AGG_MIS2 = 0, 1, 2
SMOOTHER_JACOBI, SMOOTHER_RBGS, SMOOTHER_CHEBYSHEV = 0, 1, 2
CG_CLASSIC, CG_SINGLE_REDUCE = 0, 1
CODING_AUTO, CODING_OFF = 0, 1


class Options(C.Structure):
    _fields_ = [
        ("tolerance", C.c_double), ("min_iteration", C.c_int), ("max_iteration", C.c_int), ("norm", C.c_int),
        ("min_strict", C.c_int), ("precond", C.c_int), ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int),
        ("mg_max_levels", C.c_int), ("mg_coarse_size", C.c_int), ("mg_pre_sweeps", C.c_int), ("mg_post_sweeps", C.c_int),
        ("mg_omega", C.c_double), ("mg_smoother", C.c_int), ("mg_smoothed", C.c_int), ("mg_prolong_omega", C.c_double),
        ("spmv_kernel", C.c_int), ("check_interval", C.c_int), ("use_graph", C.c_int), ("profile", C.c_int),
        ("mg_smoothed_levels", C.c_int), ("mg_aggregate_size", C.c_int), ("mg_aggregate_size_coarse", C.c_int),
        ("dist_replicate_below", C.c_int), ("dist_transport", C.c_int), ("mg_tail_rows", C.c_int), ("mg_aggregation", C.c_int),
        ("mg_cheby_degree", C.c_int), ("mg_strength_theta", C.c_double), ("mg_cheby_ratio", C.c_double), ("cg_variant", C.c_int),
        ("spmv_coding", C.c_int), ("reserved", C.c_int * 2),
    ]


class Result(C.Structure):
    _fields_ = [
        ("status", C.c_int), ("iteration", C.c_int), ("iterations_done", C.c_int), ("residual", C.c_double),
        ("rr", C.c_double), ("rr0", C.c_double), ("setup_ms", C.c_float), ("solve_ms", C.c_float), ("levels", C.c_int),
        ("kernel_launches", C.c_int), ("spmv_ms_total", C.c_float), ("spmv_launches", C.c_int),
        ("operator_complexity", C.c_double), ("reserved", C.c_int * 6),
    ]


# every symbol include/mgcg.h declares
NATIVE_SYMBOLS = [
    "mgcg_version", "mgcg_last_error", "mgcg_options_default", "mgcg_create", "mgcg_destroy", "mgcg_set_stream",
    "mgcg_setup", "mgcg_solve", "mgcg_solve_host", "mgcg_spmv", "mgcg_spmv_dot", "mgcg_dot", "mgcg_max_abs",
    "mgcg_axpy", "mgcg_scal", "mgcg_copy", "mgcg_update_xr", "mgcg_update_p", "mgcg_mg_levels", "mgcg_mg_level_size",
    "mgcg_mg_level_get", "mgcg_mg_level_aggregates", "mgcg_mg_apply", "mgcg_partition_offsets",
    "mgcg_comm_unique_id", "mgcg_comm_init", "mgcg_setup_dist", "mgcg_solve_dist", "mgcg_solve_host_dist", "mgcg_comm_finalize", "mgcg_dist_microbench",
    "mgcg_generate_poisson3d7_nnz", "mgcg_generate_poisson3d7", "mgcg_generate_poisson3d27_nnz", "mgcg_generate_poisson3d27",
    "mgcg_mps_order", "mgcg_mps_count", "mgcg_mps_fill", "mgcg_test_scan_counts", "mgcg_matrix_codes", "mgcg_dot_static", "mgcg_test_pool_stats",
]
# every symbol include/mgcg_formats.h declares (ELL input, Write/Solve/Read/Iteration object)
FORMAT_SYMBOLS = [
    "mgcg_ell_nnz", "mgcg_ell_to_csr", "mgcg_solve_host_ell", "mgcg_computer_create", "mgcg_computer_destroy",
    "mgcg_computer_set_preconditioner", "mgcg_computer_write", "mgcg_computer_solve", "mgcg_computer_read",
    "mgcg_computer_iteration", "mgcg_computer_result",
]
# the reference's 32 MgcgGpu.dll exports (+ the C entry of SolveGpu and the preconditioner switch)
COMPAT_SYMBOLS = [
    "GetDeviceCount", "SetDevice", "CreateBlas", "DestroyBlas", "CreateSparse", "DestroySparse", "CreateMatDescr",
    "DestroyMatDescr", "Create_Double", "CopyToArray_Double", "CopyFromArray_Double", "Delete_Double", "ToRawPtr_Double",
    "CopyFromDevice_Double", "Create_Int", "CopyToArray_Int", "CopyFromArray_Int", "Delete_Int", "ToRawPtr_Int",
    "CsrMV", "Axpy", "Dot", "Scal", "Copy", "Initialize", "P2Host", "P2Device", "Solve0", "Solve1", "Solve2", "Solve3",
    "Solve", "SolveGpu_c", "MgcgCompatSetPreconditioner",
]

_lib = None


def build(verbose=False):
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building libMgcgGpu.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(this package has no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, ip, dp, cd, ci = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_int
    L.mgcg_version.restype = ci
    L.mgcg_last_error.restype = C.c_char_p
    L.mgcg_options_default.argtypes = [C.POINTER(Options)]
    L.mgcg_create.argtypes = [ci]; L.mgcg_create.restype = vp
    L.mgcg_destroy.argtypes = [vp]; L.mgcg_destroy.restype = None
    L.mgcg_set_stream.argtypes = [vp, vp]
    L.mgcg_setup.argtypes = [vp, ci, ci, vp, vp, vp, C.POINTER(Options)]
    L.mgcg_solve.argtypes = [vp, vp, vp, C.POINTER(Result)]
    L.mgcg_solve_host.argtypes = [vp, ci, ci, vp, vp, vp, vp, vp, C.POINTER(Options), C.POINTER(Result)]
    L.mgcg_spmv.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, vp, cd, cd, ci]
    L.mgcg_matrix_codes.argtypes = [vp, C.POINTER(ci), vp, vp]
    L.mgcg_spmv_dot.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp, vp, vp, vp, ci]
    L.mgcg_dot.argtypes = [vp, ci, vp, vp, vp, dp]
    L.mgcg_max_abs.argtypes = [vp, ci, vp, vp, dp]
    L.mgcg_dot_static.argtypes = [ci, ci, vp, vp, vp, vp]
    L.mgcg_test_pool_stats.argtypes = [ci, ci, vp]
    L.mgcg_axpy.argtypes = [vp, ci, cd, vp, vp]
    L.mgcg_scal.argtypes = [vp, ci, cd, vp]
    L.mgcg_copy.argtypes = [vp, ci, vp, vp]
    L.mgcg_update_xr.argtypes = [vp, ci, cd, vp, vp, vp, vp, vp, dp]
    L.mgcg_update_p.argtypes = [vp, ci, cd, vp, vp]
    L.mgcg_mg_levels.argtypes = [vp]
    L.mgcg_mg_level_size.argtypes = [vp, ci, ip, ip]
    L.mgcg_mg_level_get.argtypes = [vp, ci, vp, vp, vp]
    L.mgcg_mg_level_aggregates.argtypes = [vp, ci, vp]
    L.mgcg_mg_apply.argtypes = [vp, vp, vp]
    L.mgcg_partition_offsets.argtypes = [ci, ci, vp]
    L.mgcg_comm_unique_id.argtypes = [vp]
    L.mgcg_comm_init.argtypes = [vp, C.c_char_p, ci, ci]
    L.mgcg_setup_dist.argtypes = [vp, ci, ci, ci, ci, vp, vp, vp, C.POINTER(Options)]
    L.mgcg_solve_dist.argtypes = [vp, vp, vp, C.POINTER(Result)]
    L.mgcg_solve_host_dist.argtypes = [vp, ci, ci, ci, ci, vp, vp, vp, vp, vp, C.POINTER(Options), C.POINTER(Result)]
    L.mgcg_dist_microbench.argtypes = [vp, ci, dp]
    L.mgcg_comm_finalize.argtypes = [vp]
    L.mgcg_generate_poisson3d7_nnz.argtypes = [ci] * 5; L.mgcg_generate_poisson3d7_nnz.restype = C.c_longlong
    L.mgcg_generate_poisson3d7.argtypes = [vp] + [ci] * 5 + [vp, vp, vp]
    L.mgcg_generate_poisson3d27_nnz.argtypes = [ci] * 5; L.mgcg_generate_poisson3d27_nnz.restype = C.c_longlong
    L.mgcg_generate_poisson3d27.argtypes = [vp] + [ci] * 6 + [vp, dp, vp, vp, vp]
    # compat exports
    L.mgcg_mps_order.argtypes = [vp, ci, vp, cd, vp, vp]; L.mgcg_mps_order.restype = C.c_longlong
    L.mgcg_mps_count.argtypes = [vp, ci, vp, cd, ci, vp]; L.mgcg_mps_count.restype = C.c_longlong
    L.mgcg_mps_fill.argtypes = [vp, ci, vp, cd, cd, vp, vp, vp]
    L.mgcg_test_scan_counts.argtypes = [vp, ci, ci, C.POINTER(C.c_longlong)]
    L.mgcg_ell_nnz.argtypes = [vp, ci, ci, vp]; L.mgcg_ell_nnz.restype = C.c_longlong
    L.mgcg_ell_to_csr.argtypes = [vp, ci, ci, vp, vp, vp, vp, vp, vp]
    L.mgcg_solve_host_ell.argtypes = [vp, ci, ci, vp, vp, vp, vp, vp, C.POINTER(Options), C.POINTER(Result)]
    L.mgcg_computer_create.argtypes = [ci, ci]; L.mgcg_computer_create.restype = vp
    L.mgcg_computer_destroy.argtypes = [vp]; L.mgcg_computer_destroy.restype = None
    L.mgcg_computer_set_preconditioner.argtypes = [vp, ci, ci, ci, ci]
    L.mgcg_computer_write.argtypes = [vp, vp, vp, vp, ci, vp, vp]
    L.mgcg_computer_solve.argtypes = [vp, cd, ci, ci]
    L.mgcg_computer_read.argtypes = [vp, vp]
    L.mgcg_computer_iteration.argtypes = [vp]
    L.mgcg_computer_result.argtypes = [vp, C.POINTER(Result)]
    L.GetDeviceCount.restype = ci
    L.SetDevice.argtypes = [ci]; L.SetDevice.restype = None
    for n in ("CreateBlas", "CreateSparse", "CreateMatDescr"):
        getattr(L, n).restype = vp
    for n in ("DestroyBlas", "DestroySparse", "DestroyMatDescr", "Delete_Double", "Delete_Int"):
        getattr(L, n).argtypes = [vp]; getattr(L, n).restype = None
    L.Create_Double.argtypes = [ci]; L.Create_Double.restype = vp
    L.Create_Int.argtypes = [ci]; L.Create_Int.restype = vp
    for n in ("CopyToArray_Double", "CopyFromArray_Double", "CopyToArray_Int", "CopyFromArray_Int", "CopyFromDevice_Double"):
        getattr(L, n).argtypes = [vp, vp, ci, ci, ci]; getattr(L, n).restype = None
    L.ToRawPtr_Double.argtypes = [vp]; L.ToRawPtr_Double.restype = vp
    L.ToRawPtr_Int.argtypes = [vp]; L.ToRawPtr_Int.restype = vp
    L.CsrMV.argtypes = [vp, vp, vp, vp, vp, vp, vp, ci, ci, ci, cd, cd]; L.CsrMV.restype = None
    L.Axpy.argtypes = [vp, vp, vp, ci, cd]; L.Axpy.restype = None
    L.Dot.argtypes = [vp, vp, vp, ci]; L.Dot.restype = cd
    L.Scal.argtypes = [vp, vp, cd, ci]; L.Scal.restype = None
    L.Copy.argtypes = [vp, vp, vp, ci, ci, ci]; L.Copy.restype = None
    L.Initialize.argtypes = [vp] * 11 + [ip, ip] + [ci] * 5; L.Initialize.restype = None
    L.P2Host.argtypes = [vp, vp, ci, ci, ci, ci]; L.P2Host.restype = None
    L.P2Device.argtypes = [vp, vp, ci, ci, ci, ci]; L.P2Device.restype = None
    L.Solve0.argtypes = [vp] * 11 + [ci] * 4; L.Solve0.restype = cd
    L.Solve1.argtypes = [vp] * 8 + [ci] * 4; L.Solve1.restype = cd
    L.Solve2.argtypes = [vp, cd, vp, vp, vp, vp, ci, ci]; L.Solve2.restype = cd
    L.Solve3.argtypes = [vp, cd, vp, vp, ci, ci]; L.Solve3.restype = None
    L.Solve.argtypes = [vp] * 11 + [ci, ci, cd, ci, ci, ip, dp]; L.Solve.restype = None
    L.SolveGpu_c.argtypes = [vp] * 5 + [ci, ci, cd, ci, ci, ip, dp]; L.SolveGpu_c.restype = None
    L.MgcgCompatSetPreconditioner.argtypes = [ci] * 4; L.MgcgCompatSetPreconditioner.restype = None
    _lib = L
    return L


def last_error():
    return lib().mgcg_last_error().decode("utf-8", "replace")


class MgcgError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        super().__init__(f"{where}: status {status}: {last_error()}")


def check(status, where):
    if status != OK:
        raise MgcgError(status, where)
