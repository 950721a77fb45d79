"""The FIRST host-pointer solve of a fresh process (context creation, module load, every device allocation, hierarchy build, graph
capture included), then two more.  usage: python tools/cold_start.py [n_side=256]   (MGCG_TRACE=1 prints the phases)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
t_imp = time.perf_counter()
import numpy as np, torch
from conjugategradient_b200 import generators as g, solver as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
v, rp, c = g.poisson3d7(n)
b = g.rhs_uniform(n ** 3, 0)
hv, hr, hc, hb = (torch.from_numpy(a).pin_memory() for a in (v, rp, c, b))
hx = [torch.zeros(n ** 3, dtype=torch.float64).pin_memory() for _ in range(3)]
torch.cuda.synchronize()
o = S.default_options(precond=1, norm=2, tolerance=1e-8, max_iteration=5000, nx=n, ny=n, nz=n)
t0 = time.perf_counter()
m = S.Mgcg(0)
t1 = time.perf_counter()
out = []
for i in range(3):
    ta = time.perf_counter()
    r = m.solve_host(hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx[i].numpy(), o)
    out.append((1e3 * (time.perf_counter() - ta), r.iterations_done, r.setup_ms, r.solve_ms))
print(f"{n}^3: context {1e3 * (t1 - t0):.1f} ms; host-pointer solves (ms, iterations, setup ms, solve ms): " + "; ".join(f"{a:.1f} {b_} {c_:.1f} {d:.1f}" for a, b_, c_, d in out))
