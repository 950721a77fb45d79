"""The C-ABI library loads on a CPU-only box and exports every symbol include/*.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    from conjugategradient_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.lib()


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^[A-Za-z_][\w \*]*?\b(\w+)\s*\(", src, flags=re.M)
    return [n for n in names if n not in ("defined",)]


def test_native_symbols_exported(L):
    from conjugategradient_b200 import _lib
    decl = _declared("mgcg.h")
    assert set(decl) == set(_lib.NATIVE_SYMBOLS), set(decl) ^ set(_lib.NATIVE_SYMBOLS)
    for n in decl:
        assert hasattr(L, n), n


def test_reference_export_names(L):
    # the 32 exports of MgcgGpu.dll + SolveGpu_c + the preconditioner switch
    from conjugategradient_b200 import _lib
    decl = _declared("mgcg_compat.h")
    assert set(decl) == set(_lib.COMPAT_SYMBOLS), set(decl) ^ set(_lib.COMPAT_SYMBOLS)
    assert len(_lib.COMPAT_SYMBOLS) == 32 + 2
    for n in decl:
        assert hasattr(L, n), n


def test_format_symbols_exported(L):
    # ELL input + the four-call object of the reference's other variants (include/mgcg_formats.h)
    from conjugategradient_b200 import _lib
    decl = _declared("mgcg_formats.h")
    assert set(decl) == set(_lib.FORMAT_SYMBOLS), set(decl) ^ set(_lib.FORMAT_SYMBOLS)
    for n in decl:
        assert hasattr(L, n), n


def test_version_and_partition_rule(L):
    import numpy as np
    assert L.mgcg_version() == 200
    off = np.zeros(5, dtype=np.int32)
    assert L.mgcg_partition_offsets(207402, 4, off.ctypes.data_as(ctypes.c_void_p)) == 0
    assert off.tolist() == [0, 51850, 103700, 155550, 207402]   # ConjugateGradientParallelGpu.cs:271-277


def test_fails_loudly_without_gpu(L):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from conjugategradient_b200 import _lib
    assert not L.mgcg_create(0)
    assert "no CPU fallback" in _lib.last_error()


def test_options_struct_layout(L):
    from conjugategradient_b200.solver import default_options
    o = default_options()
    assert o.tolerance == 1e-8 and o.norm == 0 and o.precond == 0 and o.mg_pre_sweeps == 1 and abs(o.mg_omega - 0.85) < 1e-15
    assert o.check_interval == 8 and o.use_graph == 2 and abs(o.mg_prolong_omega - 2 / 3) < 1e-15


def test_generator_sizes_host_side():
    from conjugategradient_b200 import _lib
    L = _lib.lib()
    assert L.mgcg_generate_poisson3d7_nnz(4, 4, 4, 2, 2) == -1 and L.mgcg_generate_poisson3d7_nnz(4, 4, 4, 0, 5) == -1
    assert L.mgcg_generate_poisson3d7_nnz(256, 256, 256, 0, 256) == 117047296          # SURVEY.md §8 size table, C4
    assert L.mgcg_generate_poisson3d27_nnz(512, 512, 512, 0, 512) == 3609741304        # C5: exceeds int32 globally
