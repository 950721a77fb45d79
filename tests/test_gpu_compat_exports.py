"""Every BLAS-level and vector export of the reference's MgcgGpu.dll, called through its own name and argument order and
checked against the CPU oracle / exact expectations (these seven were exported but never called).
Reference semantics: Mgcg/cuBlas/MgcgGpu/Mgcg.cu:10-54 (CsrMV, Axpy, Dot, Scal, Copy), Vector_Double.cu:7-45, Vector_Int.cu:9-40.
Argument orders that are easy to transpose: Dot(cublas, y, x, count); Scal(cublas, x, alpha, count);
Copy(cublas, y, x, count, yOffset, xOffset) == cublasDcopy(count, x + xOffset, 1, y + yOffset, 1)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from conjugategradient_b200 import _lib
    L = _lib.lib()
    L.SetDevice(0)
    h = (L.CreateBlas(), L.CreateSparse(), L.CreateMatDescr())
    yield L, h
    L.DestroyBlas(h[0]); L.DestroySparse(h[1]); L.DestroyMatDescr(h[2])


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _dvec(L, a):
    v = L.Create_Double(len(a))
    L.CopyFromArray_Double(v, _p(np.ascontiguousarray(a, dtype=np.float64)), len(a), 0, 0)
    return v


def _ivec(L, a):
    v = L.Create_Int(len(a))
    L.CopyFromArray_Int(v, _p(np.ascontiguousarray(a, dtype=np.int32)), len(a), 0, 0)
    return v


def _read(L, v, n):
    out = np.empty(n)
    L.CopyToArray_Double(v, _p(out), n, 0, 0)
    return out


def test_create_is_zero_and_array_copies_honour_offsets(env):
    L, _ = env
    v = L.Create_Double(10)
    assert np.array_equal(_read(L, v, 10), np.zeros(10))            # thrust::device_vector(n) value-initialises
    src = np.arange(100, 120, dtype=np.float64)
    L.CopyFromArray_Double(v, _p(src), 4, 3, 5)                      # (dst, src, count, sourceOffset, destinationOffset)
    got = _read(L, v, 10)
    assert np.array_equal(got, [0, 0, 0, 0, 0, 103, 104, 105, 106, 0])
    out = np.full(12, -1.0)
    L.CopyToArray_Double(v, _p(out), 3, 5, 7)                        # (src, dst, count, sourceOffset, destinationOffset)
    assert np.array_equal(out, [-1] * 7 + [103, 104, 105] + [-1] * 2)
    L.Delete_Double(v)
    iv = L.Create_Int(6)
    isrc = np.arange(50, 60, dtype=np.int32)
    L.CopyFromArray_Int(iv, _p(isrc), 3, 2, 1)
    iout = np.full(8, -7, dtype=np.int32)
    L.CopyToArray_Int(iv, _p(iout), 6, 0, 2)
    assert iout.tolist() == [-7, -7, 0, 52, 53, 54, 0, 0]
    L.Delete_Int(iv)


def test_to_raw_ptr_and_copy_from_device(env):
    """ToRawPtr_* return the device address of element 0 (usable by the raw-pointer exports); CopyFromDevice_Double copies
    `count` ELEMENTS device-to-device with element offsets.  (The reference passes `count` to cudaMemcpy as a BYTE count,
    Vector_Double.cu:44 — an evident slip in an export its C# side never binds; the element meaning of every sibling is kept.)"""
    import torch
    L, _ = env
    a = np.random.default_rng(0).standard_normal(64)
    va, vb = _dvec(L, a), L.Create_Double(64)
    pa, pb = L.ToRawPtr_Double(va), L.ToRawPtr_Double(vb)
    assert pa and pb and pa != pb
    L.CopyFromDevice_Double(pa, pb, 10, 20, 30)                      # (source, destination, count, sourceOffset, destinationOffset)
    got = _read(L, vb, 64)
    exp = np.zeros(64); exp[30:40] = a[20:30]
    assert np.array_equal(got, exp)
    # the raw pointer really is the vector's storage: write through it with torch's allocator-independent memcpy
    t = torch.arange(64, dtype=torch.float64, device="cuda:0")
    L.CopyFromDevice_Double(C.c_void_p(t.data_ptr()), pb, 64, 0, 0)
    assert np.array_equal(_read(L, vb, 64), np.arange(64.0))
    # ToRawPtr_Int: the address is the int vector's storage — torch reads it back through a raw-pointer view
    iv = _ivec(L, np.arange(5, dtype=np.int32) * 3)
    pi = L.ToRawPtr_Int(iv)
    assert pi
    class _Raw:                                                      # zero-copy view of the raw pointer (CUDA array interface)
        __cuda_array_interface__ = {"shape": (5,), "typestr": "<i4", "data": (int(pi), False), "version": 2}
    assert torch.as_tensor(_Raw(), device="cuda:0").cpu().tolist() == [0, 3, 6, 9, 12]
    L.Delete_Double(va); L.Delete_Double(vb); L.Delete_Int(iv)


def test_axpy_scal_dot_copy_argument_order(env, oracle):
    L, (blas, _, _) = env
    rng = np.random.default_rng(3)
    n = 5000
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    vx, vy = _dvec(L, x), _dvec(L, y)
    px, py = L.ToRawPtr_Double(vx), L.ToRawPtr_Double(vy)
    L.Axpy(blas, py, px, n, 0.37)                                    # y += alpha x          (Mgcg.cu:22-27)
    y1 = _read(L, vy, n)
    # one fused multiply-add per element (as cublasDaxpy): within half an ulp of the exact value, x untouched
    assert np.all(np.abs(y1 - (y + 0.37 * x)) <= 2.3e-16 * (np.abs(y) + np.abs(0.37 * x))) and np.array_equal(_read(L, vx, n), x)
    L.Scal(blas, px, -2.5, n)                                        # x *= alpha            (Mgcg.cu:41-46)
    x1 = x * -2.5
    assert np.array_equal(_read(L, vx, n), x1)
    d = L.Dot(blas, py, px, n)                                       # sum_i x_i y_i         (Mgcg.cu:30-38)
    ref = oracle.dot(x1, y1)                                         # LongVector.Dot: sequential left-to-right
    assert abs(d - ref) <= 1e-13 * np.abs(x1 * y1).sum()
    # sub-range calls through pointer arithmetic, as the C# side does with offsets (count < length)
    L.Scal(blas, C.c_void_p(px + 8 * 100), 2.0, 50)
    x2 = x1.copy(); x2[100:150] *= 2.0
    assert np.array_equal(_read(L, vx, n), x2)
    # Copy(y, x, count, yOffset, xOffset): y[yOffset + i] = x[xOffset + i]                   (Mgcg.cu:49-54)
    L.Copy(blas, py, px, 40, 7, 1000)
    y2 = y1.copy(); y2[7:47] = x2[1000:1040]
    assert np.array_equal(_read(L, vy, n), y2) and np.array_equal(_read(L, vx, n), x2)
    L.Copy(blas, py, px, 0, 3, 4)                                    # empty copy is a no-op
    assert np.array_equal(_read(L, vy, n), y2)
    L.Delete_Double(vx); L.Delete_Double(vy)


@pytest.mark.parametrize("shape", ["square", "rectangular", "band160"])
def test_csrmv_export(env, oracle, gen, shape):
    """CsrMV(cusparse, descr, y, elements, rowOffsets, columnIndeces, x, elementsCount, rowCount, columnCount, alpha, beta):
    y = alpha*A*x + beta*y, rectangular allowed (Mgcg.cu:10-19)."""
    L, (_, sparse, descr) = env
    rng = np.random.default_rng(11)
    if shape == "square":
        v, rp, c = gen.poisson3d7(19, 12, 7)
        rows = cols = len(rp) - 1
    elif shape == "band160":
        v, rp, c, _, _ = gen.banded_sin(4000, 160)                   # the reference's own matrix (MgcgMain.cs:51-104)
        rows = cols = 4000
    else:
        rows, cols = 700, 1900
        cnt = rng.integers(0, 30, rows)
        rp = np.zeros(rows + 1, dtype=np.int32); np.cumsum(cnt, out=rp[1:])
        c = rng.integers(0, cols, rp[-1]).astype(np.int32)
        v = rng.standard_normal(rp[-1])
    x, y0 = rng.standard_normal(cols), rng.standard_normal(rows)
    va, vr, vc, vx, vy = _dvec(L, v), _ivec(L, rp), _ivec(L, c), _dvec(L, x), _dvec(L, y0)
    args = (sparse, descr, L.ToRawPtr_Double(vy), L.ToRawPtr_Double(va), L.ToRawPtr_Int(vr), L.ToRawPtr_Int(vc), L.ToRawPtr_Double(vx),
            len(v), rows, cols)
    # oracle SpMV works on a rows x cols CSR as well (it only walks row offsets)
    ax = np.empty(rows)
    oracle.lib().ref_spmv(rows, np.ascontiguousarray(v), np.ascontiguousarray(rp, dtype=np.int32), np.ascontiguousarray(c), np.ascontiguousarray(x), ax)
    scale = np.empty(rows)
    oracle.lib().ref_spmv(rows, np.abs(v), np.ascontiguousarray(rp, dtype=np.int32), np.ascontiguousarray(c), np.abs(x), scale)
    scale += 1e-300
    L.CsrMV(*args, 1.0, 0.0)
    y = _read(L, vy, rows)
    assert np.max(np.abs(y - ax) / scale) < 1e-14
    L.CsrMV(*args, -0.5, 2.0)                                        # y = alpha A x + beta y (y holds A x now)
    y2 = _read(L, vy, rows)
    assert np.max(np.abs(y2 - (-0.5 * ax + 2.0 * y)) / (scale + np.abs(y))) < 1e-14
    assert np.array_equal(_read(L, vx, cols), x)
    for d in (va, vx, vy):
        L.Delete_Double(d)
    L.Delete_Int(vr); L.Delete_Int(vc)
