"""Sweep of the rows SpMV kernel's launch shape (lanes per row, stages, stage capacity, blocks per SM, L1 carve-out) on the C3
particle matrices; isolated SpMV+dot launches back to back on the library's stream.
One-off tuning tool: python tools/spmv_sweep.py [n_side] > gpurun_out/spmv_sweep.txt"""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from conjugategradient_b200 import solver as S, device_generators as dg
import bench

n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 160
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
torch.cuda.set_stream(torch.cuda.Stream(dev))      # torch's events and the library's launches on one stream
m = S.Mgcg(0, use_torch_stream=True)
L = m.L
peak = bench.measured_peak()[0]
KNOBS = ("MGCG_ROWS_UNROLL", "MGCG_ROWS_LANES", "MGCG_ROWS_STAGES", "MGCG_ROWS_CAPF", "MGCG_ROWS_CTAS", "MGCG_ROWS_CARVEOUT")


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


def clear():
    for v in KNOBS:
        os.environ.pop(v, None)


for re in (2.1, 3.1):
    pos = torch.from_numpy(bench.mps_positions(n_side)).to(dev)
    tv, tr, tc, order, _ = dg.mps_particles(m.ctx, pos, re, shift_abs=1.0, max_per_row=1 << 20)
    del pos, order
    n, nnz = tr.numel() - 1, tv.numel()
    print(f"== MPS {n_side}^3 re={re}: rows {n} nnz {nnz} avg {nnz / n:.1f}", flush=True)
    clear()
    for k, name in ((2, "vector"), (1, "stream"), (3, "rows (library default shape)")):
        ms, gbs = timeit(tv, tr, tc, n, nnz, k)
        print(f"  {name:34s} {ms:8.3f} ms {gbs:7.0f} GB/s  {gbs / peak:.3f}", flush=True)
    res = []
    lanes_set = (4, 8, 16) if re < 3 else (8, 16, 32)
    quick = os.environ.get("SWEEP_QUICK") is not None
    for unroll, lanes, stages, capf, ctas in itertools.product((4, 8), lanes_set, (2,) if quick else (2, 3, 4), (1.3,) if quick else (1.15, 1.3, 1.6), (4, 5) if quick else (2, 3, 4, 5, 6)):
        clear()
        os.environ.update({"MGCG_ROWS_UNROLL": str(unroll), "MGCG_ROWS_LANES": str(lanes), "MGCG_ROWS_STAGES": str(stages), "MGCG_ROWS_CAPF": str(capf), "MGCG_ROWS_CTAS": str(ctas)})
        ms, gbs = timeit(tv, tr, tc, n, nnz, 3, reps=6)
        res.append((gbs, lanes, stages, capf, ctas, ms, unroll))
        print(f"  rows U={unroll} L={lanes:2d} stages={stages} capf={capf:4.2f} ctas={ctas}   {ms:8.3f} ms {gbs:7.0f} GB/s  {gbs / peak:.3f}", flush=True)
    res.sort(reverse=True)
    print("  best:", [tuple(round(v, 3) if isinstance(v, float) else v for v in r) for r in res[:6]], flush=True)
    gbs, lanes, stages, capf, ctas, _, unroll = res[0]
    for carve in (50, 65, 80, 100):
        clear()
        os.environ.update({"MGCG_ROWS_UNROLL": str(unroll), "MGCG_ROWS_LANES": str(lanes), "MGCG_ROWS_STAGES": str(stages), "MGCG_ROWS_CAPF": str(capf), "MGCG_ROWS_CTAS": str(ctas),
                           "MGCG_ROWS_CARVEOUT": str(carve)})
        ms, gbs = timeit(tv, tr, tc, n, nnz, 3, reps=6)
        print(f"  best shape, carve-out {carve:3d} %   {ms:8.3f} ms {gbs:7.0f} GB/s  {gbs / peak:.3f}", flush=True)
    clear()
    del tv, tr, tc
    torch.cuda.empty_cache()
