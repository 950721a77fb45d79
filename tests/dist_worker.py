"""Worker of the sharded-solve tests: run under torchrun, one rank per GPU (tests/test_gpu_dist.py launches it).
Solves the same system (a) sharded over WORLD_SIZE GPUs and (b) on this rank's GPU alone, and compares."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conjugategradient_b200 import generators as g  # noqa: E402
from conjugategradient_b200 import solver as S  # noqa: E402


def main():
    cfg = json.loads(sys.argv[1])
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("gloo")
    n = cfg["n"]
    N = n ** 3
    if cfg.get("matrix", "p3d7") == "p3d7":
        v, rp, c = g.poisson3d7(n)
    else:
        v, rp, c = g.poisson3d27_varcoef(n)
    b = g.rhs_uniform(N, 0)
    # plane-aligned row blocks (== the reference's floor rule, ConjugateGradientParallelGpu.cs:271-277, when world | n)
    planes = [n // world] * world
    planes[-1] += n - sum(planes)
    offs = np.concatenate([[0], np.cumsum(planes)]) * n * n
    r0, r1 = int(offs[rank]), int(offs[rank + 1])
    e0, e1 = int(rp[r0]), int(rp[r1])
    tv = torch.as_tensor(v[e0:e1]).to(dev)
    tc = torch.as_tensor(c[e0:e1]).to(dev)
    tr = torch.as_tensor((rp[r0:r1 + 1] - rp[r0]).astype(np.int32)).to(dev)
    tb = torch.as_tensor(b[r0:r1]).to(dev)
    tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
    ids = [S.MgcgDist.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    m = S.MgcgDist(lr, rank, world, ids[0])
    kw = dict(cfg.get("options", {}))
    if kw.get("precond", 0) == 1 and cfg.get("hint", True):
        kw.update(nx=n, ny=n, nz=n)
    elif kw.get("precond", 0) == 1:
        kw.update(mg_aggregation=1)     # without a grid the sharded hierarchy aggregates by blocks of 8 rows (csrc/dist.cu agg_of)
    o = S.default_options(norm=2, tolerance=cfg.get("tolerance", 1e-8), max_iteration=2000, **kw)
    m.setup(N, r0, tv, tr, tc, o)
    res = m.solve(tb, tx)
    for _ in range(cfg.get("setups", 1) - 1):     # a re-setup reuses the peer arena, the private coarse context and cached blocks
        x_first = tx.clone()
        m.setup(N, r0, tv, tr, tc, o)
        res2 = m.solve(tb, tx.zero_())
        assert res2.iterations_done == res.iterations_done and torch.equal(tx, x_first), "solve after a re-setup differs"
        res = res2
    for _ in range(cfg.get("solves", 1) - 1):     # repeated solves replay the captured iteration; results must not move
        x_first = tx.clone()
        res2 = m.solve(tb, tx.zero_())
        assert res2.iterations_done == res.iterations_done and torch.equal(tx, x_first), "repeated sharded solve differs"
    torch.cuda.synchronize()
    import ctypes as C
    nd = C.c_int(0)
    m.L.mgcg_matrix_codes(m.ctx, C.byref(nd), None, None)
    if cfg.get("coded"):
        # coded columns restate the shard's columns exactly: the solve on int32 columns gives the same bits
        x_coded = tx.clone()
        kw0 = dict(kw); kw0["spmv_coding"] = 1      # MGCG_CODING_OFF
        m.setup(N, r0, tv, tr, tc, S.default_options(norm=2, tolerance=cfg.get("tolerance", 1e-8), max_iteration=2000, **kw0))
        res0 = m.solve(tb, tx.zero_())
        if cfg.get("matrix", "p3d7") == "p3d7":     # same kernel family, storage-order sums on both sides: the same bits
            assert res0.iterations_done == res.iterations_done and torch.equal(tx, x_coded), "sharded solve on coded columns differs from int32 columns"
        else:                                        # 27-point: int32 columns run on the lanes-per-row kernel (other summation order)
            assert abs(res0.iterations_done - res.iterations_done) <= 1
            # (two solves of a contrast-1e6 system stopped at 1e-8 differ by rounding order times conditioning; the comparison against the
            # 1e-12 solve below is the parity check proper)
            assert float((tx - x_coded).abs().max() / x_coded.abs().max()) < 1e-5, "sharded solve on coded columns differs from int32 columns"
        tx.copy_(x_coded)
        torch.cuda.synchronize()
    # the same system on this GPU alone
    ms = S.Mgcg(lr)
    fv, fr, fc, fb = (torch.as_tensor(a).to(dev) for a in (v, rp, c, b))
    fx = torch.zeros(N, dtype=torch.float64, device=dev)
    ms.setup(fv, fr, fc, o)
    rs = ms.solve(fb, fx)
    torch.cuda.synchronize()
    err = float((tx - fx[r0:r1]).abs().max() / fx.abs().max())
    tight = {}
    if cfg.get("tight"):
        # ill-conditioned case: both solutions against a solve of the same system taken four orders further down.  The sharded x
        # has to be as close to it as the single-GPU x is (+1e-10) — the direct difference of two 1e-8 solves carries the
        # rounding-order noise of both, amplified by the conditioning.
        ot = S.default_options(norm=2, tolerance=cfg["tight"], max_iteration=4000, **kw)
        ms.setup(fv, fr, fc, ot)
        xt = torch.zeros(N, dtype=torch.float64, device=dev)
        rt = ms.solve(fb, xt, raise_on_fail=False)
        torch.cuda.synchronize()
        sc = float(xt.abs().max())
        tight = {"tight_residual": rt.residual, "tight_it": rt.iterations_done,
                 "err_tight_dist": float((tx - xt[r0:r1]).abs().max() / sc), "err_tight_single": float((fx - xt).abs().max() / sc),
                 "err_tight_single_local": float((fx[r0:r1] - xt[r0:r1]).abs().max() / sc)}
    # oracle leg: the same system on rank 0's host cores by the CPU restatements (oracle/cg_ref.c for plain CG, the twin oracle/mg_ref.c
    # for MGCG with a grid hint) — the reference's own check of its multi-device solver is against its CPU solver too
    # (Mgcg/cuBlas/Mgcg/MgcgMain.cs:151-162).  Without a grid hint the sharded levels aggregate per shard; that variant has no twin.
    twin = {}
    pre = kw.get("precond", 0)
    if pre == 0 or cfg.get("hint", True):
        xt = np.empty(N)
        if rank == 0:
            from oracle import oracle as orc
            tol = cfg.get("tolerance", 1e-8)
            if pre == 0:
                xt, itt, _, stt = orc.cg_solve(v, rp, c, b, tol=tol, norm=orc.NORM_REL_L2, max_iteration=2000)
            else:
                H = orc.MgRef(v, rp, c, n, n, n)
                if kw.get("mg_smoother", 0) == 2:
                    H.set_chebyshev(2, 6.0)                      # the library defaults (mgcg_options_default)
                xt, itt, _, stt = H.pcg(b, tol=tol, norm=orc.NORM_REL_L2, max_iteration=2000)
                H.close()
            twin = {"it_twin": itt + 1, "twin_status": stt}      # iterations done = 0-based index + 1
        bx = torch.from_numpy(np.ascontiguousarray(xt))
        dist.broadcast(bx, src=0)
        xt = bx.numpy()
        tw = [twin]
        dist.broadcast_object_list(tw, src=0)
        twin = dict(tw[0])
        twin["err_twin"] = float(np.abs(tx.cpu().numpy() - xt[r0:r1]).max() / np.abs(xt).max())
    out = {"rank": rank, **twin, "it_dist": res.iterations_done, "it_single": rs.iterations_done, "err": err, "residual": res.residual,
           "levels": res.levels, "first_replicated": res.reserved[2], "halo_exchanges": res.reserved[0], "allgathers": res.reserved[1],
           "solve_ms": res.solve_ms, "setup_ms": res.setup_ms, "single_ms": rs.solve_ms, "launches": res.kernel_launches,
           "transport": res.reserved[3], "graph": res.reserved[4], "coded": nd.value, **tight}
    allo = [None] * world
    dist.all_gather_object(allo, out)
    if rank == 0:
        print("DIST_RESULT " + json.dumps(allo), flush=True)
    m.close(); ms.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
