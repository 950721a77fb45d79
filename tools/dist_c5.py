"""BASELINE.json config 5: 27-point variable-coefficient (contrast 1e6, anisotropy 1e-2) Poisson, sharded over the ranks of one
node, assembled per shard in device memory (the global 512^3 matrix has 3.6e9 non-zeros: more than int32, 44 GB).
Run under torchrun; prints one JSON line on rank 0.  One-off measurement, not a test."""
import json, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conjugategradient_b200 import generators as g, solver as S, device_generators as dg

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
planes = [n // world] * world
planes[-1] += n - sum(planes)
zo = np.concatenate([[0], np.cumsum(planes)])
z0, z1 = int(zo[rank]), int(zo[rank + 1])
r0, r1 = z0 * n * n, z1 * n * n
ids = [S.MgcgDist.unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
m = S.MgcgDist(lr, rank, world, ids[0])
tv, tr, tc = dg.poisson3d27_varcoef(m.ctx, n, zrange=(z0, z1), device=dev)
tb = torch.from_numpy(g.rhs_uniform(n ** 3, 0)[r0:r1].copy()).to(dev)
tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
o = S.default_options(precond=1, nx=n, ny=n, nz=n, norm=2, tolerance=1e-8, max_iteration=2000)
m.setup(n ** 3, r0, tv, tr, tc, o)
m.solve(tb, tx); tx.zero_()
torch.cuda.synchronize(); dist.barrier()
res = m.solve(tb, tx)
t = torch.tensor([res.solve_ms], dtype=torch.float64, device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
nnz = torch.tensor([tv.numel()], dtype=torch.int64, device=dev)
dist.all_reduce(nnz)
m.setup(n ** 3, r0, tv, tr, tc, o)        # steady-state setup time
res2 = m.solve(tb, tx.zero_())
if rank == 0:
    print(json.dumps({"config": f"C5 27-point variable coefficient {n}^3 on {world} GPUs", "rows": n ** 3, "nnz_global": int(nnz.item()),
                      "nnz_rank0": tv.numel(), "iterations": res.iterations_done, "solve_ms_max_over_ranks": round(float(t.item()), 2),
                      "relative_residual": res.residual, "setup_ms_steady": round(res2.setup_ms, 1), "levels": res.levels,
                      "distributed_levels": res.reserved[2], "transport": "peer" if res.reserved[3] == 2 else "nccl"}), flush=True)
m.close()
dist.barrier(); dist.destroy_process_group()
