"""Sweep multigrid options on the GPU: iterations, setup and solve time (tuning aid, not a test)."""
import sys, itertools, torch
sys.path.insert(0, '/root/repo')
from conjugategradient_b200 import generators as g, solver as S
n = int(sys.argv[1])
v, rp, c = g.poisson3d7(n)
b = g.rhs_uniform(n ** 3, 0)
dev = torch.device('cuda:0')
tv, tr, tc, tb = (torch.as_tensor(a).to(dev) for a in (v, rp, c, b))
m = S.Mgcg(0)
for kw in eval(sys.argv[2]):
    o = S.default_options(precond=1, nx=n, ny=n, nz=n, norm=2, tolerance=1e-8, max_iteration=500, **kw)
    try:
        m.setup(tv, tr, tc, o)
        tx = torch.zeros(n ** 3, dtype=torch.float64, device=dev)
        m.solve(tb, tx); tx.zero_()
        r = m.solve(tb, tx)
        print(kw, 'it', r.iterations_done, 'setup_ms %.1f solve_ms %.2f levels %d opc %.3f launches %d' % (r.setup_ms, r.solve_ms, r.levels, r.operator_complexity, r.kernel_launches), flush=True)
    except Exception as e:
        print(kw, 'FAILED', e, flush=True)
