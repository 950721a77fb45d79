"""Device-side stencil generators (mgcg_generate_*) against the host generators: bit-exact CSR, whole grids and z-slabs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import torch
    from conjugategradient_b200 import solver as S
    assert torch.cuda.is_available()
    m = S.Mgcg(0)
    yield m
    m.close()


@pytest.mark.parametrize("dims,zr", [((16, 16, 16), None), ((9, 5, 7), None), ((9, 5, 7), (2, 5)), ((9, 5, 7), (0, 1)), ((9, 5, 7), (6, 7)),
                                     ((1, 1, 5), None), ((4, 1, 1), None), ((1, 1, 1), None), ((64, 48, 3), (1, 3)), ((33, 17, 40), (8, 24))])
def test_poisson3d7_device_matches_host(ctx, gen, dims, zr):
    import torch
    from conjugategradient_b200 import device_generators as dg
    v, rp, c = gen.poisson3d7(*dims, zrange=zr)
    tv, tr, tc = dg.poisson3d7(ctx.ctx, *dims, zrange=zr, device="cuda")
    torch.cuda.synchronize()
    assert np.array_equal(tr.cpu().numpy(), rp) and np.array_equal(tc.cpu().numpy(), c) and np.array_equal(tv.cpu().numpy(), v)


@pytest.mark.parametrize("dims,zr", [((12, 12, 12), None), ((9, 5, 7), None), ((9, 5, 7), (2, 5)), ((9, 5, 7), (0, 1)), ((9, 5, 7), (6, 7)),
                                     ((1, 1, 5), None), ((3, 1, 1), None), ((20, 17, 24), (8, 16))])
def test_poisson3d27_device_matches_host(ctx, gen, dims, zr):
    import torch
    from conjugategradient_b200 import device_generators as dg
    v, rp, c = gen.poisson3d27_varcoef(*dims)
    if zr is not None:
        r0, r1 = zr[0] * dims[0] * dims[1], zr[1] * dims[0] * dims[1]
        e0, e1 = int(rp[r0]), int(rp[r1])
        v, c, rp = v[e0:e1], c[e0:e1], (rp[r0:r1 + 1] - rp[r0]).astype(np.int32)
    tv, tr, tc = dg.poisson3d27_varcoef(ctx.ctx, *dims, zrange=zr, device="cuda")
    torch.cuda.synchronize()
    assert np.array_equal(tr.cpu().numpy(), rp) and np.array_equal(tc.cpu().numpy(), c)
    assert np.array_equal(tv.cpu().numpy(), v)          # bit-exact: same products, same order, no FMA contraction
