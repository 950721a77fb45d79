"""One-off parity record at the size the headline is quoted on (BASELINE config C4, 256^3): the CUDA path against the CPU twin
(oracle/mg_ref.c) — hierarchy bit for bit, MGCG iteration index, relative solution error.  The twin needs ~2.5 minutes of one
host core at this size, which is why bench.py cross-checks against a 1e-12 device solve instead and this script is run once per
round with its output committed under profiles/.   usage: python tools/parity_c4.py [n_side] > profiles/rNN_parity_c4.json"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from conjugategradient_b200 import generators as g, solver as S
from oracle import oracle as o

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
v, rp, c = g.poisson3d7(n)
b = g.rhs_uniform(n ** 3, 0)
m = S.Mgcg(0)
tv, tr, tc = (torch.as_tensor(a).cuda() for a in (v, rp, c))
m.setup(tv, tr, tc, S.default_options(precond=1, nx=n, ny=n, nz=n, norm=2, tolerance=1e-8, max_iteration=1000))
tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n ** 3, dtype=torch.float64, device="cuda")
res = m.solve(tb, tx)
x = tx.cpu().numpy()
t0 = time.time()
H = o.MgRef(v, rp, c, n, n, n)
t_build = time.time() - t0
levels = []
ok = m.levels() == H.levels()
for l in range(min(m.levels(), H.levels())):
    gv, gr, gc = m.level(l)
    hv, hr, hc = H.level(l)
    same_s = bool(np.array_equal(gr, hr) and np.array_equal(gc, hc))
    same_v = bool(same_s and np.array_equal(gv, hv))
    same_a = bool(np.array_equal(m.aggregates(l), H.aggregates(l))) if l + 1 < m.levels() else None
    levels.append({"level": l, "rows": int(len(gr) - 1), "nnz": int(len(gv)), "structure_bit_exact": same_s, "values_bit_exact": same_v,
                   "aggregates_bit_exact": same_a})
    ok = ok and same_v and (same_a is not False)
t0 = time.time()
xt, itt, rest, st = H.pcg(b, tol=1e-8, norm=o.NORM_REL_L2, max_iteration=1000)
t_pcg = time.time() - t0
err = float(np.abs(x - xt).max() / np.abs(xt).max())
print(json.dumps({"config": f"C4 3-D 7-point Poisson {n}^3 MGCG fp64 to 1e-8", "device_iteration": res.iteration, "twin_iteration": itt,
                  "x_rel_err_vs_twin": err, "device_residual": res.residual, "twin_residual": rest, "twin_status": st,
                  "hierarchy_bit_exact": bool(ok), "levels": levels, "twin_build_s": round(t_build, 1), "twin_pcg_s": round(t_pcg, 1),
                  "bars": {"x_rel_err": 1e-10, "iteration": "+-1"}, "pass": bool(ok and abs(res.iteration - itt) <= 1 and err <= 1e-10)}))
