"""SpMV(+dot) kernels on a 27-point variable-coefficient operator assembled on the device: the WIDE TMA stream kernel (one thread per
row, AUTO's choice for 8..32 per row) against the lanes-per-row ROWS kernel and the plain VECTOR kernel; results compared.
usage: python tools/spmv_27pt.py [n_side=256] [nz=n_side]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from conjugategradient_b200 import solver as S, device_generators as dg, _lib
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
nz = int(sys.argv[2]) if len(sys.argv) > 2 else n
dev = torch.device("cuda:0")
m = S.Mgcg(0, use_torch_stream=True)
tv, tr, tc = dg.poisson3d27_varcoef(m.ctx, n, n, nz, device=dev)
rows, nnz = tr.numel() - 1, tv.numel()
x = torch.rand(rows, dtype=torch.float64, device=dev)
d = torch.zeros(1, dtype=torch.float64, device=dev)
peak, _ = bench.measured_peak()
byt = 12.0 * nnz + 20.0 * rows
print(f"27-point {n}x{n}x{nz}: rows {rows} nnz {nnz} ({nnz / rows:.1f} per row), {byt / 1e9:.2f} GB per launch, peak {peak:.0f} GB/s")
ys = {}
for name, kernel in (("auto (wide TMA stream)", _lib.SPMV_AUTO), ("rows", _lib.SPMV_ROWS), ("vector", _lib.SPMV_VECTOR)):
    y = torch.empty_like(x)
    for _ in range(3):
        S.check(m.L.mgcg_spmv_dot(m.ctx, rows, rows, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), kernel), "spmv")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps):
        m.L.mgcg_spmv_dot(m.ctx, rows, rows, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), kernel)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ys[name] = (y, float(d.item()))
    print(f"  {name:26s} {ms:8.3f} ms  {byt / ms / 1e6:8.0f} GB/s  {byt / ms / 1e6 / peak:5.3f} of peak")
# the set-up matrix: coded columns, STENCIL kernel on 64-row tiles (values, codes and the x runs staged by TMA)
m.setup(tv, tr, tc, S.default_options())
nd = __import__("ctypes").c_int(0)
m.L.mgcg_matrix_codes(m.ctx, __import__("ctypes").byref(nd), None, None)
if nd.value > 0:
    byt9 = 9.0 * nnz + 20.0 * rows
    y = torch.empty_like(x)
    for _ in range(3):
        S.check(m.L.mgcg_spmv_dot(m.ctx, rows, rows, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), 0), "spmv")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        m.L.mgcg_spmv_dot(m.ctx, rows, rows, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ys["stencil (coded, %d distances)" % nd.value] = (y, float(d.item()))
    print(f"  {'stencil (coded columns)':26s} {ms:8.3f} ms  {byt9 / ms / 1e6:8.0f} GB/s of its 9 B/entry layout = {byt9 / ms / 1e6 / peak:5.3f} of peak; CSR-equivalent {byt / ms / 1e6:8.0f} GB/s")
ref = ys["vector"][0]
for k, (y, dot) in ys.items():
    print(f"  {k:26s} max|y - y_vector| / max|y| = {float((y - ref).abs().max() / ref.abs().max()):.2e}  dot {dot:.15e}")
