"""A handful of isolated SpMV(+dot) launches on a C3 particle matrix, for ncu.  usage: python tools/profile_spmv.py re n_side kernel"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from conjugategradient_b200 import solver as S, device_generators as dg
import bench
re, n_side, kernel = float(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda:0")
m = S.Mgcg(0)
pos = torch.from_numpy(bench.mps_positions(n_side)).to(dev)
tv, tr, tc, order, _ = dg.mps_particles(m.ctx, pos, re, shift_abs=1.0, max_per_row=1 << 20)
n, nnz = tr.numel() - 1, tv.numel()
x = torch.rand(n, dtype=torch.float64, device=dev); y = torch.empty_like(x); d = torch.zeros(1, dtype=torch.float64, device=dev)
for _ in range(4):
    S.check(m.L.mgcg_spmv_dot(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x), S._p(y), S._p(x), S._p(d), kernel), "spmv")
torch.cuda.synchronize()
print("ok", n, nnz)
