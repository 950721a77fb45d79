"""Sweep of the rows SpMV kernel's launch shape (ring size, blocks per SM, lanes per row) on the C3 particle matrices, isolated
SpMV+dot launches back to back.  One-off tuning tool: python tools/spmv_sweep.py [n_side] > gpurun_out/spmv_sweep.txt"""
import ctypes as C, itertools, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from conjugategradient_b200 import solver as S, device_generators as dg
import bench

n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 160
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
torch.cuda.set_stream(torch.cuda.Stream(dev))      # torch's events and the library's launches on one stream
m = S.Mgcg(0, use_torch_stream=True)
L = m.L
peak = bench.measured_peak()[0]


def timeit(tv, tr, tc, n, nnz, kernel, reps=10):
    x = torch.rand(n, dtype=torch.float64, device=dev); y = torch.empty_like(x); d = torch.zeros(1, dtype=torch.float64, device=dev)
    for _ in range(2):
        S.check(L.mgcg_spmv_dot(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), kernel), "spmv")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        L.mgcg_spmv_dot(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), kernel)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return ms, (12.0 * nnz + 20.0 * n) / (ms * 1e-3) / 1e9


for re in (2.1, 3.1):
    pos = torch.from_numpy(bench.mps_positions(n_side)).to(dev)
    tv, tr, tc, order, _ = dg.mps_particles(m.ctx, pos, re, shift_abs=1.0, max_per_row=1 << 20)
    del pos, order
    n, nnz = tr.numel() - 1, tv.numel()
    print(f"== MPS {n_side}^3 re={re}: rows {n} nnz {nnz} avg {nnz / n:.1f}", flush=True)
    for k, name in ((2, "vector"), (1, "stream")):
        for v in ("MGCG_ROWS_TOTAL", "MGCG_ROWS_CTAS", "MGCG_ROWS_LANES", "MGCG_ROWS_CARVEOUT"):
            os.environ.pop(v, None)
        ms, gbs = timeit(tv, tr, tc, n, nnz, k)
        print(f"  {name:28s} {ms:8.3f} ms {gbs:7.0f} GB/s  {gbs / peak:.3f}", flush=True)
    res = []
    for total, ctas, lanes in itertools.product((2560, 3584, 4096, 5120, 6144, 8192), (2, 3, 4, 5), (8, 16, 32)):
        if (total * 12 + 1600) * ctas > 226 * 1024:
            continue
        os.environ["MGCG_ROWS_TOTAL"] = str(total); os.environ["MGCG_ROWS_CTAS"] = str(ctas); os.environ["MGCG_ROWS_LANES"] = str(lanes)
        ms, gbs = timeit(tv, tr, tc, n, nnz, 3)
        res.append((gbs, total, ctas, lanes, ms))
        print(f"  rows total={total:5d} ctas={ctas} L={lanes:2d}   {ms:8.3f} ms {gbs:7.0f} GB/s  {gbs / peak:.3f}", flush=True)
    res.sort(reverse=True)
    print("  best:", res[:5], flush=True)
    # carve-out preference on the best shape (attribute is set once per process and instantiation: last step of the loop)
    del tv, tr, tc
    torch.cuda.empty_cache()
