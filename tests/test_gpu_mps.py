"""Particle (MPS) matrix assembled on the device (mgcg_mps_*) against its CPU statement oracle/mps_ref.c: bit-exact
ordering, structure and values; then solved with MGCG.  NOT YET RUN ON A GPU when written (round 1 budget spent)."""
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
