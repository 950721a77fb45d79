"""The block cache under the hierarchy build (csrc/alloc.cu) through the C ABI: slab sizing and reuse of dropped blocks.

The reference allocates every vector and matrix with its own cudaMalloc (Vector_Double.cu / Mgcg.cu:57-85) and has nothing like it;
the cache exists because a cold build made hundreds of driver allocations.  What is pinned here:
  * a small system does not reserve a 512 MiB slab any more (the slab follows the set-up's own size hint);
  * blocks that contexts leave behind when they are destroyed are handed out again even while ANOTHER context keeps the same slab
    alive — the slab total stays flat over repeated create / set-up / solve / destroy cycles, and the solves stay bit-identical;
  * trimming gives everything back once nothing is live.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _stats(L, trim=0):
    out = (C.c_longlong * 5)()
    assert L.mgcg_test_pool_stats(0, trim, out) == 0
    return dict(zip(("cached", "slabs", "slab_bytes", "used", "holes"), [int(v) for v in out]))


def _system(torch, gen, n):
    v, rp, c = gen.poisson3d7(n)
    b = gen.rhs_uniform(n ** 3, 0)
    return [torch.as_tensor(a).to("cuda:0") for a in (v, rp, c, b)]


def _solve(torch, S, sysm, n):
    tv, tr, tc, tb = sysm
    m = S.Mgcg(0)
    m.setup(tv, tr, tc, S.default_options(precond=1, nx=n, ny=n, nz=n, norm=2, tolerance=1e-8, max_iteration=200))
    tx = torch.zeros(n ** 3, dtype=torch.float64, device="cuda:0")
    res = m.solve(tb, tx)
    return m, res, tx


def test_small_system_opens_a_small_slab(gen):
    import torch
    from conjugategradient_b200 import _lib, solver as S
    L = _lib.lib()
    base = _stats(L, trim=1)
    m, res, _ = _solve(torch, S, _system(torch, gen, 16), 16)
    assert res.status == 0
    grown = _stats(L)["slab_bytes"] - base["slab_bytes"]
    m.close()
    assert grown <= (128 << 20), grown          # was >= 512 MiB for any first allocation


def test_dropped_blocks_are_reused_beside_a_live_context(gen):
    import torch
    from conjugategradient_b200 import _lib, solver as S
    L = _lib.lib()
    n = 48
    sysm = _system(torch, gen, n)
    keeper, res0, x0 = _solve(torch, S, sysm, n)   # stays alive: pins the slab(s) its blocks were carved from
    assert res0.status == 0
    totals = []
    for _ in range(8):
        m, res, x = _solve(torch, S, sysm, n)
        assert res.status == 0 and res.iteration == res0.iteration
        assert torch.equal(x, x0)                  # recycled space, same bits
        m.close()                                  # its cached blocks are dropped -> holes of the slab the keeper pins
        totals.append(_stats(L)["slab_bytes"])
    assert totals[-1] == totals[1], totals         # flat after the first cycles (used to grow by a slab every few cycles)
    r2 = keeper.solve(sysm[3], x0.zero_())         # the keeper's own blocks were never handed to anyone else
    assert r2.status == 0 and r2.iteration == res0.iteration
    keeper.close()
    after = _stats(L, trim=1)
    assert after["cached"] == 0
    assert after["used"] - after["holes"] <= after["slab_bytes"]
