"""BASELINE.json's other configs on one GPU (one-off measurements for DESIGN.md, not tests):
C2 7-point 128^3, C3 MPS particles (160^3, r_e = 2.1 l0), C5-like 27-point variable coefficient 256^3 (device generator)."""
import sys, time, json
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from conjugategradient_b200 import generators as g, solver as S, device_generators as dg

dev = torch.device('cuda:0')
which = sys.argv[1:] or ['c2', 'c5', 'c3']


def run(name, tv, tr, tc, tb, **kw):
    n = tr.numel() - 1
    m = S.Mgcg(0)
    o = S.default_options(precond=1, norm=2, tolerance=1e-8, max_iteration=2000, **kw)
    m.setup(tv, tr, tc, o)
    tx = torch.zeros(n, dtype=torch.float64, device=dev)
    m.solve(tb, tx); tx.zero_()
    r = m.solve(tb, tx)
    ty = torch.empty_like(tx)
    S.check(m.L.mgcg_spmv(m.ctx, n, n, tv.numel(), S._p(tv), S._p(tr), S._p(tc), S._p(tx), S._p(ty), 1.0, 0.0, 0), "spmv")
    torch.cuda.synchronize()
    true_rel = float(torch.linalg.vector_norm(tb - ty) / torch.linalg.vector_norm(tb))
    o0 = S.default_options(precond=0, norm=2, tolerance=1e-8, max_iteration=20000)
    m0 = S.Mgcg(0); m0.setup(tv, tr, tc, o0); tx.zero_(); r0 = m0.solve(tb, tx, raise_on_fail=False)
    print(json.dumps({"config": name, "rows": n, "nnz": tv.numel(), "mgcg_iterations": r.iterations_done, "mgcg_solve_ms": round(r.solve_ms, 2),
                      "setup_ms": round(r.setup_ms, 1), "levels": r.levels, "operator_complexity": round(r.operator_complexity, 3),
                      "true_relative_residual": true_rel, "plain_cg_iterations": r0.iterations_done, "plain_cg_solve_ms": round(r0.solve_ms, 2)}), flush=True)
    m.close(); m0.close()


ctxm = S.Mgcg(0)
if 'c2' in which:
    n = 128
    tv, tr, tc = dg.poisson3d7(ctxm.ctx, n, device=dev)
    run("C2 7-point 128^3", tv, tr, tc, torch.from_numpy(g.rhs_uniform(n ** 3, 0)).to(dev), nx=n, ny=n, nz=n)
if 'c5' in which:
    n = 256
    tv, tr, tc = dg.poisson3d27_varcoef(ctxm.ctx, n, device=dev)
    run("C5-like 27-point variable coefficient 256^3 (contrast 1e6, anisotropy 1e-2)", tv, tr, tc, torch.from_numpy(g.rhs_uniform(n ** 3, 0)).to(dev), nx=n, ny=n, nz=n)
if 'c3' in which:
    t = time.time()
    v, rp, c = g.mps_particles(int(sys.argv[-1]) if sys.argv[-1].isdigit() else 160)
    print("mps generated in %.1f s: rows %d nnz %d" % (time.time() - t, len(rp) - 1, len(v)), flush=True)
    tv, tr, tc = (torch.as_tensor(a).to(dev) for a in (v, rp, c))
    run("C3 MPS particles, r_e = 2.1 l0", tv, tr, tc, torch.from_numpy(g.rhs_uniform(len(rp) - 1, 2)).to(dev))
