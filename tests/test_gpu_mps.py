"""Particle (MPS) matrix assembled on the device (mgcg_mps_*) against its CPU statement oracle/mps_ref.c: bit-exact
ordering, structure and values; then solved with MGCG (strength filter + MIS-2 aggregation) against the CPU twin."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cloud(n_side, jitter, seed):
    rng = np.random.default_rng(seed)
    g = np.arange(n_side)
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    pos = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1) + rng.uniform(-jitter, jitter, size=(n_side ** 3, 3))
    return pos[rng.permutation(len(pos))]


@pytest.mark.parametrize("n_side,re", [(8, 2.1), (13, 2.1), (9, 3.1)])
def test_device_assembly_bit_exact(oracle, n_side, re):
    import torch
    from conjugategradient_b200 import solver as S, device_generators as dg
    pos = _cloud(n_side, 0.2, n_side)
    m = S.Mgcg(0)
    val, rp, col, order, ps = dg.mps_particles(m.ctx, torch.as_tensor(pos).cuda(), re, shift_abs=0.125, max_per_row=160)
    torch.cuda.synchronize()
    ov, orp, oc, oorder = oracle.mps_assemble(pos, re, shift_abs=0.125, max_per_row=160)
    assert np.array_equal(order.cpu().numpy(), oorder)
    assert np.array_equal(ps.cpu().numpy(), pos[oorder])
    assert np.array_equal(rp.cpu().numpy(), orp) and np.array_equal(col.cpu().numpy(), oc)
    assert np.array_equal(val.cpu().numpy(), ov)
    m.close()


def test_device_assembly_row_limit(oracle):
    import torch
    from conjugategradient_b200 import _lib, solver as S, device_generators as dg
    pos = _cloud(8, 0.2, 3)
    m = S.Mgcg(0)
    with pytest.raises(_lib.MgcgError):
        dg.mps_particles(m.ctx, torch.as_tensor(pos).cuda(), 2.1, max_per_row=8)
    m.close()


def test_device_assembled_system_solves(oracle):
    import torch
    from conjugategradient_b200 import solver as S, device_generators as dg
    pos = _cloud(16, 0.2, 5)
    m = S.Mgcg(0)
    val, rp, col, order, ps = dg.mps_particles(m.ctx, torch.as_tensor(pos).cuda(), 2.1, shift_abs=0.05)
    n = len(pos)
    b = torch.as_tensor(np.random.default_rng(0).uniform(-1, 1, n)).cuda()
    x = torch.zeros(n, dtype=torch.float64, device="cuda")
    m.setup(val, rp, col, S.default_options(precond=1, norm=2, tolerance=1e-8, max_iteration=500))
    res = m.solve(b, x)
    xo, ito, _, st = oracle.cg_solve(val.cpu().numpy(), rp.cpu().numpy(), col.cpu().numpy(), b.cpu().numpy(), tol=1e-10,
                                     norm=oracle.NORM_REL_L2)
    assert st == oracle.OK and res.iterations_done < ito
    assert np.abs(x.cpu().numpy() - xo).max() <= 1e-6 * np.abs(xo).max()
    m.close()


@pytest.mark.parametrize("re", [2.1, 3.1])
def test_c3_parity_at_64_cubed_particles(oracle, re):
    """BASELINE config C3 at 64^3 = 262 144 particles (the bench runs 160^3): the device-assembled matrix is bit-identical to the
    CPU statement; the hierarchy built from it (strength filter, MIS-2 aggregates, filtered smoothed prolongator, Galerkin products
    through the warp-merge SpGEMM with 4 / 8 lists per lane) is bit-identical to the CPU twin's, level by level; MGCG converges at
    the twin's iteration index with the solution within 1e-10."""
    import torch
    from conjugategradient_b200 import solver as S, device_generators as dg
    n_side = 64
    rng = np.random.default_rng(1)
    g = np.arange(n_side, dtype=np.float64)
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    pos = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1) + rng.uniform(-0.2, 0.2, size=(n_side ** 3, 3))
    m = S.Mgcg(0)
    val, rp, col, order, ps = dg.mps_particles(m.ctx, torch.as_tensor(pos).cuda(), re, shift_abs=0.02, max_per_row=160)
    torch.cuda.synchronize()
    ov, orp, oc, oorder = oracle.mps_assemble(pos, re, shift_abs=0.02, max_per_row=160)
    assert np.array_equal(order.cpu().numpy(), oorder)
    assert np.array_equal(rp.cpu().numpy(), orp) and np.array_equal(col.cpu().numpy(), oc) and np.array_equal(val.cpu().numpy(), ov)
    n = len(pos)
    b = np.random.default_rng(2).uniform(-1, 1, n)
    m.setup(val, rp, col, S.default_options(precond=1, norm=2, tolerance=1e-8, max_iteration=500))
    H = oracle.MgRef(ov, orp, oc)
    assert m.levels() == H.levels() >= 3
    for l in range(m.levels()):
        if l + 1 < m.levels():
            assert np.array_equal(m.aggregates(l), H.aggregates(l)), f"level {l} aggregates"
        gv, gr, gc = m.level(l)
        hv, hr, hc = H.level(l)
        assert np.array_equal(gr, hr) and np.array_equal(gc, hc), f"level {l} structure"
        assert np.array_equal(gv, hv), f"level {l} values: max diff {np.abs(gv - hv).max()}"
    tb, tx = torch.as_tensor(b).cuda(), torch.zeros(n, dtype=torch.float64, device="cuda")
    res = m.solve(tb, tx)
    xt, itt, _, st = H.pcg(b, tol=1e-8, norm=oracle.NORM_REL_L2, max_iteration=500)
    x = tx.cpu().numpy()
    err = np.abs(x - xt).max() / np.abs(xt).max()
    print(f"C3 64^3 re={re}: device iteration {res.iteration}, twin {itt}, rel err {err:.3e}, levels {res.levels}, opc {res.operator_complexity:.3f}")
    assert st == 0 and abs(res.iteration - itt) <= 1
    assert err <= (1e-10 if res.iteration == itt else 1e-6)
    m.close(); H.close()
