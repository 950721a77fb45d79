#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on its own workload: MGCG time-to-1e-8 (+ SpMV HBM GB/s) on the 3-D 7-point Poisson
256^3 system, fp64, at 1/2/4/8 B200 (row-block sharded for N > 1).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size 256]

A "step" is ONE complete solve of the system to 1e-8 relative residual (x0 = 0).  One JSON line is printed by rank 0:
  value / ms_per_step : solve time with the matrix, the hierarchy and b resident in HBM (CUDA events, max over ranks)
  e2e                 : the same solve through the host-pointer C-ABI call (mgcg_solve_host): pinned host CSR + b in,
                        x out, upload + hierarchy build + solve + download all inside the timed region
  roofline            : the fine-level SpMV(+p.Ap) kernel, algorithmic bytes 12*nnz + 20*N per launch, timed live with CUDA
                        events inside a solve, against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline        : the reference's CPU path (oracle port of ConjugateGradientCpu.Solve: plain, unpreconditioned,
                        single-threaded like the reference runs it) on a bounded sample of the same system
`--impl reference` times that CPU path alone (the reference has no multigrid and cannot be compiled here).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "mgcg_time_to_1e-8_poisson3d7"
TOL = 1e-8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=256, help="grid points per side (BASELINE: 256)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--device-gen", action="store_true", help="assemble the matrix on the device (mgcg_generate_poisson3d7); implied for --size > 256, "
                    "where the e2e and CPU legs (which need the host copy) are skipped")
    ap.add_argument("--graph", type=int, default=-1, help="use_graph option: 0 eager, 1 one iteration per graph launch, 2 device-side while loop "
                    "(library default); profilers want 0")
    ap.add_argument("--transport", type=int, default=0, help="sharded arm: 0 auto (peer memory on one node), 1 NCCL, 2 peer memory")
    ap.add_argument("--replicate-below", type=int, default=0, help="sharded arm: replicate multigrid levels smaller than this (0 = library default)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel_substr, n_side):
    """DRAM bytes (read + write) of one launch of the dominant kernel from the committed `ncu --set full` capture of this
    workload (profiles/r01_ncu_full_spmv_tma_256.csv, taken at 256^3); None for other sizes or if the file is absent."""
    import csv
    p = os.path.join(ROOT, "profiles", "r01_ncu_full_spmv_tma_256.csv")
    if n_side != 256 or not os.path.exists(p):
        return None
    try:
        rows = list(csv.reader(open(p)))
        hdr, units = rows[0], rows[1]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        for r in rows[2:]:
            if kernel_substr in r[0]:
                return float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
    except Exception:
        return None
    return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# workload (synthetic): 3-D 7-point Poisson, diag 6 / off -1, diagonal stored first, b ~ U(-1,1) seed 0
def make_problem(n_side):
    from conjugategradient_b200 import generators as g
    val, rp, col = g.poisson3d7(n_side)
    b = g.rhs_uniform(n_side ** 3, 0)
    return val, rp, col, b


# Iterations the reference's (unpreconditioned) CG needs for ||r||/||r0|| < 1e-8 on this workload.  Measured with this
# repo's GPU plain-CG path, whose iteration count is parity-tested (+-1) against the oracle (tests/test_gpu_cg.py); the
# CPU legs time a bounded number of iterations and scale to this count ("time a fixed number of
# iterations and extrapolate, stated as such").
PLAIN_CG_ITERATIONS = {32: 111, 48: 177, 64: 230, 128: 445}


def plain_cg_iterations(n_side):
    if n_side in PLAIN_CG_ITERATIONS:
        return PLAIN_CG_ITERATIONS[n_side], "table"
    return int(round(3.48 * n_side)), "estimate 3.48*n (fit to 64/128)"


def cpu_sample(problem, n_side, iters, threads):
    """Run `iters` iterations of the reference CPU CG (oracle port) and return seconds per iteration."""
    from oracle import oracle as o
    val, rp, col, b = problem
    t0 = time.perf_counter()
    if threads == 1:
        _, it, _, st = o.cg_solve(val, rp, col, b, tol=0.0, max_iteration=iters - 2, norm=o.NORM_REL_L2)
    else:
        _, it, _, st = o.cg_solve_mt(val, rp, col, b, tol=0.0, max_iteration=iters - 2, norm=o.NORM_REL_L2, threads=threads)
    dt = time.perf_counter() - t0
    done = it + 1
    return dt / done, done


def cpu_baseline_leg(problem, n_side, budget_s):
    """cpu_baseline object: the reference's CPU path, all host cores, bounded sample, scaled to time-to-1e-8."""
    from oracle import oracle as o
    cores = o.max_threads()
    per_it, _ = cpu_sample(problem, n_side, 3, cores)                  # calibration
    iters = max(3, min(200, int(budget_s / max(per_it, 1e-6))))
    per_it, done = cpu_sample(problem, n_side, iters, cores)
    need, how = plain_cg_iterations(n_side)
    per_it1, done1 = cpu_sample(problem, n_side, max(3, min(20, int(0.3 * budget_s / max(per_it * cores, 1e-6)) or 3)), 1)
    return {"value": per_it * (need + 1) * 1e3, "unit": "ms", "cores": cores, "kind": "port",
            "sample": f"{done} iterations of the reference's unpreconditioned CPU CG (oracle/cg_ref_mt.c, OpenMP over its "
                      f"commented-out Parallel.For loops) on the same {n_side}^3 system, {per_it * 1e3:.1f} ms/iteration, scaled to "
                      f"the {need + 1} iterations plain CG needs for 1e-8 ({how}); the reference has no multigrid (SURVEY F1)",
            "ms_per_iteration": per_it * 1e3,
            "single_thread_ms_per_iteration": per_it1 * 1e3,
            "single_thread_value": per_it1 * (need + 1) * 1e3}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle port; the C#/boost/CUDA-5 sources cannot
    be compiled here) on the box's host cores.  Each step = a bounded number of CG iterations on the same
    system, scaled to the iteration count plain CG needs for 1e-8."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as o
    n = args.size
    problem = make_problem(n)
    cores = o.max_threads()
    need, how = plain_cg_iterations(n)
    per_it, _ = cpu_sample(problem, n, 3, cores)
    total_steps = args.steps + args.warmup
    iters = max(3, min(100, int(120.0 / total_steps / max(per_it, 1e-6))))
    for _ in range(args.warmup):
        cpu_sample(problem, n, iters, cores)
    t = []
    for _ in range(args.steps):
        p, done = cpu_sample(problem, n, iters, cores)
        t.append(p)
    per_it = float(np.mean(t))
    value = per_it * (need + 1) * 1e3
    sample = (f"{iters} iterations per step of the reference's unpreconditioned CPU CG (oracle/cg_ref_mt.c) on {n}^3, "
              f"{per_it * 1e3:.1f} ms/iteration x {need + 1} iterations to 1e-8 ({how})")
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": value, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"3-D 7-point Poisson {n}^3, fp64, b~U(-1,1) seed 0, x0=0, relative residual 1e-8",
                      "rows": n ** 3, "nnz": int(problem[1][-1]), "preconditioner": "none (the reference has none)"},
           "cpu_baseline": {"value": value, "unit": "ms", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------------------------
def run_ours_single(args):
    """N = 1: the whole system on cuda:0 through the C ABI (mgcg_setup / mgcg_solve / mgcg_solve_host)."""
    import torch
    from conjugategradient_b200 import solver as S
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    torch.cuda.set_stream(torch.cuda.Stream(dev))   # a real (capturable) stream shared by torch's events and the library
    n_side = args.size
    n = n_side ** 3
    m = S.Mgcg(0, use_torch_stream=True)     # run on torch's current stream so torch.cuda.Event brackets the library's work
    device_gen = args.device_gen or n_side > 256
    if device_gen:
        from conjugategradient_b200 import device_generators as dg
        from conjugategradient_b200 import generators as g
        tv, tr, tc = dg.poisson3d7(m.ctx, n_side, device=dev)
        tb = torch.from_numpy(g.rhs_uniform(n, 0)).to(dev)
        nnz = tv.numel()
        problem = None
        args.no_e2e = True; args.no_cpu = True
    else:
        problem = make_problem(n_side)
        val, rp, col, b = problem
        nnz = int(rp[-1])
        # pinned host copies (the e2e leg uploads from these)
        hv, hr, hc, hb = (torch.from_numpy(a).pin_memory() for a in (val, rp, col, b))
        hx = torch.zeros(n, dtype=torch.float64).pin_memory()
        tv, tr, tc, tb = (t.to(dev, non_blocking=True) for t in (hv, hr, hc, hb))
    tx = torch.zeros(n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    gopt = {} if args.graph < 0 else {"use_graph": args.graph}
    opts = S.default_options(precond=1, nx=n_side, ny=n_side, nz=n_side, norm=2, tolerance=TOL, max_iteration=1000, **gopt)
    m.setup(tv, tr, tc, opts)

    def step():
        tx.zero_()
        return m.solve(tb, tx)

    for _ in range(args.warmup):
        res = step()
    torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    e0.record()
    for _ in range(args.steps):
        res = step()
        launches += res.kernel_launches + 1
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    clk = clocks.stop()
    iters = res.iterations_done
    setup_ms = res.setup_ms
    # true residual check of the last solve (not timed): ||b - A x|| / ||b||
    ty = torch.empty_like(tx)
    S.check(m.L.mgcg_spmv(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(tx), S._p(ty), 1.0, 0.0, 0), "mgcg_spmv")
    torch.cuda.synchronize()
    true_rel = float(torch.linalg.vector_norm(tb - ty) / torch.linalg.vector_norm(tb))

    # roofline leg: the fine-level SpMV(+p.Ap) launches timed with CUDA events inside one more solve (profile=1)
    popts = S.default_options(precond=1, nx=n_side, ny=n_side, nz=n_side, norm=2, tolerance=TOL, max_iteration=1000, profile=1)
    m.setup(tv, tr, tc, popts)
    tx.zero_()
    pres = m.solve(tb, tx)
    spmv_ms = pres.spmv_ms_total / max(pres.spmv_launches, 1)
    # ... and back-to-back in isolation (inputs 1.6 GB >> 126 MB L2, so no flush is needed)
    td = torch.zeros(1, dtype=torch.float64, device=dev)
    for _ in range(3):
        m.L.mgcg_spmv_dot(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(tb), S._p(ty), S._p(tb), S._p(td), 0)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    s0.record()
    for _ in range(reps):
        m.L.mgcg_spmv_dot(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(tb), S._p(ty), S._p(tb), S._p(td), 0)
    s1.record()
    torch.cuda.synchronize()
    spmv_iso_ms = s0.elapsed_time(s1) / reps
    spmv_bytes = 12.0 * nnz + 20.0 * n
    peak, peak_how = measured_peak()
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "k_spmv_tma<EPI_DOT> (fine-level A*p with fused p.Ap, TMA-staged stream kernel)", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_8TBs_nominal": achieved / 8000.0,
                "traffic": ncu_traffic("k_spmv_tma<2>", n_side),
                "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full (profiles/r01_ncu_full_spmv_tma_256.csv)",
                "peak_source": peak_how, "bytes_per_launch": spmv_bytes, "ms_per_launch": spmv_ms,
                "launches_timed": pres.spmv_launches, "isolated_ms_per_launch": spmv_iso_ms,
                "isolated_gbs": spmv_bytes / (spmv_iso_ms * 1e-3) / 1e9}

    # e2e leg: host CSR + b in, x out, everything (upload, hierarchy, solve, download) inside the timed region
    e2e = None
    if not args.no_e2e:
        m2 = S.Mgcg(0, use_torch_stream=True)
        hx.zero_()
        m2.solve_host(hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx.numpy(), opts)     # warm-up (allocations)
        k = max(1, min(args.steps, 3))
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        g0.record()
        for _ in range(k):
            hx.zero_()
            r2 = m2.solve_host(hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx.numpy(), opts)
        g1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 / k
        e2e = {"value": max(g0.elapsed_time(g1) / k, wall), "unit": "ms",
               "h2d_bytes_per_step": int(12 * nnz + 4 * (n + 1) + 16 * n), "d2h_bytes_per_step": int(8 * n),
               "setup_ms": r2.setup_ms, "solve_ms": r2.solve_ms, "iterations": r2.iterations_done,
               "api": "mgcg_solve_host (include/mgcg.h): pinned host CSR + b + x0 -> x"}
        assert abs(r2.iterations_done - iters) <= 1
        m2.close()

    cpu = None if args.no_cpu else cpu_baseline_leg(problem, n_side, args.cpu_seconds)
    out = {"metric": METRIC, "value": ms, "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": f"3-D 7-point Poisson {n_side}^3 MGCG, fp64, b~U(-1,1) seed 0, x0=0, relative residual 1e-8",
                      "rows": n, "nnz": nnz, "preconditioner": "smoothed-aggregation V(1,1), weighted Jacobi 0.85, built on device",
                      "l2": "inputs (1.6 GB matrix + 128 MiB vectors at 256^3) larger than the 126 MB L2; no flush",
                      "matrix_assembly": "device (mgcg_generate_poisson3d7)" if device_gen else "host numpy, uploaded",
                      "parallelism": "1 GPU"},
           "iterations": iters, "iterations_per_s": iters / (ms * 1e-3), "final_relative_residual": res.residual,
           "true_relative_residual": true_rel, "setup_ms": setup_ms, "levels": res.levels,
           "operator_complexity": res.operator_complexity, "spmv_gbs": achieved,
           "gpu_launches": launches, "clocks": clk, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e}
    print(json.dumps(out), flush=True)
    m.close()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 or args.gpus > 1:
        return run_ours_dist(args)
    return run_ours_single(args)


def run_ours_dist(args):
    """N > 1 (torchrun, one rank per GPU): the SAME system row-block sharded over the ranks (strong scaling): z-slab shards,
    halo exchange of p and of the multigrid level vectors plus rank-ordered scalar reductions over NCCL."""
    import torch
    import torch.distributed as dist
    from conjugategradient_b200 import generators as g
    from conjugategradient_b200 import solver as S
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs CUDA devices (there is no CPU fallback)"
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    torch.cuda.set_stream(torch.cuda.Stream(dev))
    n_side = args.size
    n = n_side ** 3
    planes = [n_side // world] * world          # == offsets[i] = offsets[i-1] + floor(Count/D) (cs:271-277) when world | n
    planes[-1] += n_side - sum(planes)
    zo = np.concatenate([[0], np.cumsum(planes)])
    z0, z1 = int(zo[rank]), int(zo[rank + 1])
    r0, r1 = z0 * n_side * n_side, z1 * n_side * n_side
    ids = [S.MgcgDist.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    m = S.MgcgDist(lr, rank, world, ids[0], use_torch_stream=True)
    b = g.rhs_uniform(n, 0)[r0:r1].copy()
    device_gen = args.device_gen or n_side > 256
    if device_gen:
        from conjugategradient_b200 import device_generators as dg
        tv, tr, tc = dg.poisson3d7(m.ctx, n_side, zrange=(z0, z1), device=dev)
        tb = torch.from_numpy(b).to(dev)
        nnz_local = tv.numel()
        args.no_e2e = True
    else:
        val, rp, col = g.poisson3d7(n_side, zrange=(z0, z1))
        nnz_local = int(rp[-1])
        hv, hr, hc, hb = (torch.from_numpy(a).pin_memory() for a in (val, rp, col, b))
        tv, tr, tc, tb = (t.to(dev, non_blocking=True) for t in (hv, hr, hc, hb))
    tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
    extra = {"dist_transport": args.transport}
    if args.graph >= 0:
        extra["use_graph"] = args.graph
    if args.replicate_below > 0:
        extra["dist_replicate_below"] = args.replicate_below
    opts = S.default_options(precond=1, nx=n_side, ny=n_side, nz=n_side, norm=2, tolerance=TOL, max_iteration=1000, **extra)
    m.setup(n, r0, tv, tr, tc, opts)

    def step():
        tx.zero_()
        return m.solve(tb, tx)

    for _ in range(args.warmup):
        res = step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(lr)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    e0.record()
    for _ in range(args.steps):
        res = step()
        launches += res.kernel_launches + 1
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clk = clocks.stop() if rank == 0 else None
    micro = (C.c_double * 3)()
    S.check(m.L.mgcg_dist_microbench(m.ctx, 200, micro), "mgcg_dist_microbench")
    # local SpMV(+p.Ap) in isolation on this rank's shard (x = a [lo|own|hi]-sized vector addressed globally, like p)
    L = m.L
    halo = n_side * n_side
    tp = torch.zeros(r1 - r0 + 2 * halo, dtype=torch.float64, device=dev)
    ty = torch.empty(r1 - r0, dtype=torch.float64, device=dev)
    td = torch.zeros(1, dtype=torch.float64, device=dev)
    xsh = C.c_void_p(tp.data_ptr() - 8 * (r0 - halo))
    wown = C.c_void_p(tp.data_ptr() + 8 * halo)
    for _ in range(3):
        L.mgcg_spmv_dot(m.ctx, r1 - r0, n, nnz_local, S._p(tv), S._p(tr), S._p(tc), xsh, S._p(ty), wown, S._p(td), 0)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    s0.record()
    for _ in range(reps):
        L.mgcg_spmv_dot(m.ctx, r1 - r0, n, nnz_local, S._p(tv), S._p(tr), S._p(tc), xsh, S._p(ty), wown, S._p(td), 0)
    s1.record()
    torch.cuda.synchronize()
    spmv_ms = s0.elapsed_time(s1) / reps
    spmv_bytes = 12.0 * nnz_local + 20.0 * (r1 - r0)
    peak, peak_how = measured_peak()
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9

    # e2e: pinned host shard -> device, hierarchy build, solve, x back to the host, all inside the timed region
    e2e = None
    if not args.no_e2e:
        hx = torch.zeros(r1 - r0, dtype=torch.float64).pin_memory()
        k = max(1, min(args.steps, 3))
        dist.barrier()
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        g0.record()
        for _ in range(k):
            dv, dr, dc, db = (h.to(dev, non_blocking=True) for h in (hv, hr, hc, hb))
            dx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
            m.setup(n, r0, dv, dr, dc, opts)
            r2 = m.solve(db, dx)
            hx.copy_(dx, non_blocking=True)
        g1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 / k
        t2 = torch.tensor([max(g0.elapsed_time(g1) / k, wall)], dtype=torch.float64, device=dev)
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": float(t2.item()), "unit": "ms", "h2d_bytes_per_step": int(12 * nnz_local + 4 * (r1 - r0 + 1) + 8 * (r1 - r0)),
               "d2h_bytes_per_step": int(8 * (r1 - r0)), "setup_ms": r2.setup_ms, "solve_ms": r2.solve_ms, "iterations": r2.iterations_done,
               "api": "MgcgDist.setup + solve (mgcg_setup_dist / mgcg_solve_dist) from pinned host shards; bytes are per rank"}
    if rank == 0:
        out = {"metric": METRIC, "value": ms, "unit": "ms", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": {"workload": f"3-D 7-point Poisson {n_side}^3 MGCG, fp64, b~U(-1,1) seed 0, x0=0, relative residual 1e-8",
                          "rows": n, "nnz": 7 * n - 6 * n_side * n_side,
                          "preconditioner": "smoothed-aggregation V(1,1), weighted Jacobi 0.85, built on device; coarse levels replicated",
                          "l2": "per-rank shard (matrix + vectors) larger than the 126 MB L2 at 256^3 up to 8 ranks; no flush",
                          "parallelism": f"row-block (z-slab) shards over {world} GPUs; halo exchange, coarse all-gather and rank-ordered "
                                         "scalar reductions by " + ("the library's own kernels over NVLink peer memory, iteration replayed as a CUDA graph"
                                                                    if res.reserved[3] == 2 else "NCCL send/recv + all-gather")},
               "transport": "peer" if res.reserved[3] == 2 else "nccl", "graph": bool(res.reserved[4]),
               "exchange_us": {"halo_level0": micro[0], "coarse_allgather": micro[1], "scalar_reduction": micro[2]},
               "iterations": res.iterations_done, "iterations_per_s": res.iterations_done / (ms * 1e-3),
               "final_relative_residual": res.residual, "setup_ms": res.setup_ms, "levels": res.levels,
               "distributed_levels": res.reserved[2], "halo_exchanges_per_solve": res.reserved[0], "allgathers_per_solve": res.reserved[1],
               "spmv_gbs": achieved * world, "gpu_launches": launches, "clocks": clk,
               "roofline": {"bound": "hbm", "kernel": "k_spmv_tma<EPI_DOT> on rank 0's shard (isolated, back to back)", "achieved": achieved,
                            "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_8TBs_nominal": achieved / 8000.0, "traffic": None,
                            "peak_source": peak_how, "bytes_per_launch": spmv_bytes, "ms_per_launch": spmv_ms},
               "cpu_baseline": None, "e2e": e2e}
        print(json.dumps(out), flush=True)
    m.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
