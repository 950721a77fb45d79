"""Host-side logic of the sharded (N > 1) path on the CPU: world-size-2 gloo group, shard layout, halo exchange and
rank-ordered reductions, checked against the oracle on the whole system."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from conjugategradient_b200 import generators as g, sharding as sh
    from oracle import oracle as o
    n = 12
    N = n ** 3
    v, rp, c = g.poisson3d7(n)
    zs, offs = sh.plane_partition(n, world)
    assert offs == sh.partition_offsets(N, world)                      # world | n: the reference's rule
    r0, r1 = offs[rank], offs[rank + 1]
    lv, lrp, lc = sh.shard_csr(v, rp, c, r0, r1)
    sv, srp, sc = g.poisson3d7(n, zrange=(zs[rank], zs[rank + 1]))     # the slab generator builds the same shard
    assert np.array_equal(lv, sv) and np.array_equal(lrp, srp) and np.array_equal(lc, sc)
    # halo width = global max of the local half bandwidths
    hb = torch.tensor([sh.half_bandwidth(lrp, lc, r0)])
    dist.all_reduce(hb, op=dist.ReduceOp.MAX)
    halo = int(hb.item())
    assert halo == n * n
    lo, hi = sh.ghost_widths(rank, world, halo)
    # a sharded vector p = [lo | own | hi]; exchange ghosts with the neighbours
    pfull = np.random.default_rng(7).standard_normal(N)
    ext = np.zeros(lo + (r1 - r0) + hi)
    ext[lo:lo + r1 - r0] = pfull[r0:r1]
    own = torch.from_numpy(ext[lo:lo + r1 - r0])
    reqs = []
    if rank > 0:
        reqs.append(dist.isend(own[:halo].clone(), rank - 1))
        glo = torch.empty(lo, dtype=torch.float64); reqs.append(dist.irecv(glo, rank - 1))
    if rank < world - 1:
        reqs.append(dist.isend(own[-halo:].clone(), rank + 1))
        ghi = torch.empty(hi, dtype=torch.float64); reqs.append(dist.irecv(ghi, rank + 1))
    for r in reqs:
        r.wait()
    if rank > 0:
        ext[:lo] = glo.numpy()
    if rank < world - 1:
        ext[lo + r1 - r0:] = ghi.numpy()
    # local SpMV through the shifted (globally indexed) view == the rows of the global product
    y = o.spmv(lv, lrp, lc, np.nan_to_num(sh.shifted_view(ext, r0, lo, N), nan=1e300))
    yfull = o.spmv(v, rp, c, pfull)
    assert np.array_equal(y, yfull[r0:r1])
    # p.Ap: per-rank partials, all-gathered, added in rank order on every rank
    part = torch.tensor([o.dot(pfull[r0:r1].copy(), y)], dtype=torch.float64)
    allp = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(allp, part)
    total = sh.rank_ordered_sum([t.item() for t in allp])
    ref = o.dot(pfull, yfull)
    assert abs(total - ref) <= 1e-13 * abs(ref)
    q.put((rank, total))
    dist.destroy_process_group()


def test_two_rank_shard_layout_and_reductions():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = sorted(q.get(timeout=5) for _ in range(world))
    assert got[0][1] == got[1][1]          # identical on every rank (same data, same order)


def test_bench_shard_planes_rules():
    """bench.py's z-plane split: the default is the reference's floor rule (ConjugateGradientParallelGpu.cs:271-277) in planes;
    --align-planes K puts every cut on a multiple of K planes without losing or duplicating a plane."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    from conjugategradient_b200 import sharding
    for n, w in ((256, 8), (512, 8), (48, 4), (50, 4), (24, 8)):
        p = b.shard_planes(n, w, 0)
        assert sum(p) == n and len(p) == w
        offs = sharding.partition_offsets(n * n * n, w)          # the reference's rule on rows
        if n % w == 0:
            assert [int(o) for o in offs] == [sum(p[:q]) * n * n for q in range(w + 1)]
        for k in (6, 12):
            pk = b.shard_planes(n, w, k)
            assert sum(pk) == n and len(pk) == w and min(pk) > 0
            if n // k >= w:
                assert all(sum(pk[:q]) % k == 0 for q in range(1, w))     # every interior cut on a multiple of k
            else:
                assert pk == p                                            # too few units: falls back to the floor rule
    assert b.shard_planes(512, 8, 6) == [66, 66, 66, 66, 66, 60, 60, 62]
