#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on its own workloads: MGCG time-to-1e-8 (+ SpMV HBM GB/s), fp64.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c4|c2|c3|c5] [--size n] [--re 2.1]

  --config c4 (default): 3-D 7-point Poisson 256^3 at 1/2/4/8 B200 (row-block sharded for N > 1) — the configuration the metric
                         is quoted on.  c2: the same at 128^3, one GPU.  c3: particle (MPS) pressure matrix, 160^3 = 4 096 000
                         particles, neighbour radius --re (2.1 or 3.1 lattice spacings), assembled on the device, no grid given to
                         the solver (strength filter + MIS-2 aggregation), one GPU.  c5: 27-point variable-coefficient 512^3,
                         sharded (needs N >= 2: the global matrix has 3.6e9 non-zeros), assembled per shard on the device
                         (--size 256 or less runs the same operator on one GPU).
A "step" is ONE complete solve of the system to 1e-8 relative residual (x0 = 0).  One JSON line is printed by rank 0:
  value / ms_per_step : solve time with the matrix, the hierarchy and b resident in HBM (CUDA events, max over ranks)
  e2e                 : the same solve through ONE host-pointer C-ABI call (mgcg_solve_host / mgcg_solve_host_dist): pinned host
                        CSR + b in, x out; upload + hierarchy build + solve + download inside the timed region.  e2e_cold is the
                        first such call of the process (allocations, peer mappings, graph instantiation included).
  roofline            : the fine-level SpMV(+p.Ap) kernel timed live with CUDA events inside a solve, against MEASURED_PEAKS.json's HBM
                        copy bandwidth.  Bytes per launch = what the kernel's layout must move: 9*nnz + 20*N on the library's coded
                        columns (fp64 value + 1-byte code per entry; the default for stencil operators), 12*nnz + 20*N on the
                        caller's int32 columns (--opt spmv_coding=1); the CSR figure is reported beside it (csr_equivalent_gbs)
  parity              : x of the timed solve against a 1e-12 solve of the same system on the same device (the CPU twin at 256^3
                        takes minutes: it is compared at 128^3 in tests/test_gpu_parity_c2.py and once at 256^3 by
                        tools/parity_c4.py -> profiles/); the plain-CG iteration count measured on the GPU; N > 1: the sharded
                        x against the single-GPU x of the same system (dist_x_rel_err)
  cpu_baseline        : the reference's CPU path (oracle port of ConjugateGradientCpu.Solve: plain, unpreconditioned) on the host
                        cores: one COMPLETE solve when that fits the time budget, else a bounded number of iterations scaled to
                        the iteration count MEASURED by the GPU plain-CG run of the same system in this run
`--impl reference` times that CPU path alone (the reference has no multigrid and cannot be compiled here): the
first warm-up step is one complete CPU solve (which measures the iteration count), every other step a bounded sample.
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

TOL = 1e-8
NCU_FULL = os.path.join("profiles", "r02_ncu_full_kernels_256.csv")       # refreshed capture of the shipped kernels
NCU_FULL_STENCIL = os.path.join("profiles", "r02_ncu_full_stencil_256.csv")  # capture of the coded-column (STENCIL) kernels
NCU_FULL_OLD = os.path.join("profiles", "r01_ncu_full_spmv_tma_256.csv")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c4", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--size", type=int, default=0, help="grid points / particles per side (default: the config's own: 128 / 160 / 256 / 512)")
    ap.add_argument("--re", type=float, default=2.1, help="c3: neighbour radius in lattice spacings (2.1: ~38 per row, 3.1: ~123 per row)")
    ap.add_argument("--theta", type=float, default=0.0, help="c3: strength threshold of the MIS-2 aggregation (0 = library default 0.5)")
    ap.add_argument("--cpu-seconds", type=float, default=30.0, help="budget of the CPU baseline leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra-512", action="store_true", help="default config only: skip the extra_512 object (the 7-point 512^3 solve on the same GPUs)")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity block (tight solve, plain-CG count, sharded-vs-single)")
    ap.add_argument("--device-gen", action="store_true", help="assemble the matrix on the device; implied for c3, c5 and --size > 256")
    ap.add_argument("--graph", type=int, default=-1, help="use_graph option: 0 eager, 1 one iteration per graph launch, 2 device-side while loop "
                    "(library default); profilers want 0")
    ap.add_argument("--transport", type=int, default=0, help="sharded arm: 0 auto (peer memory on one node), 1 NCCL, 2 peer memory")
    ap.add_argument("--align-planes", type=int, default=0, help="sharded arm: shard cuts on multiples of this many z-planes (0 = the reference's floor "
                    "rule); 6 keeps the 3- and 2-plane aggregates of two multigrid levels whole")
    ap.add_argument("--replicate-below", type=int, default=0, help="sharded arm: replicate multigrid levels smaller than this (0 = library default)")
    ap.add_argument("--smoother", type=int, default=-1, help="multigrid smoother (-1 = library default)")
    ap.add_argument("--opt", action="append", default=[], help="extra mgcg_options field, name=value (repeatable)")
    a = ap.parse_args()
    if a.size <= 0:
        a.size = {"c2": 128, "c3": 160, "c4": 256, "c5": 512}[a.config]
    return a


def metric_name(cfg):
    return {"c2": "mgcg_time_to_1e-8_poisson3d7", "c4": "mgcg_time_to_1e-8_poisson3d7", "c3": "mgcg_time_to_1e-8_mps_particles",
            "c5": "mgcg_time_to_1e-8_poisson3d27_varcoef"}[cfg]


def workload_text(args):
    n = args.size
    if args.config in ("c2", "c4"):
        return f"3-D 7-point Poisson {n}^3 MGCG, fp64, b~U(-1,1) seed 0, x0=0, relative residual 1e-8"
    if args.config == "c3":
        return (f"MPS particle pressure Poisson, {n}^3 = {n ** 3} particles on a unit lattice jittered by U(-0.2,0.2) (seed 1), neighbours r < "
                f"{args.re} l0, a_ij = -(re/r - 1), diagonal = sum + 1e-3 * mean diagonal; cell-ordered; MGCG fp64, b~U(-1,1) seed 2, x0=0, "
                "relative residual 1e-8")
    return (f"3-D 27-point variable-coefficient Poisson {n}^3 (k = 10^U(-3,3) on 8^3 blocks, seed 3, anisotropy 1:1:0.01) MGCG, fp64, "
            "b~U(-1,1) seed 0, x0=0, relative residual 1e-8")


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel_substr, args, files=None):
    """DRAM bytes (read + write) of one launch of the dominant kernel from the committed `ncu --set full` capture of the c4
    workload; None for other workloads or if no capture is committed."""
    import csv
    if args.config != "c4" or args.size != 256:
        return None, None
    for rel in (files or (NCU_FULL, NCU_FULL_OLD)):
        p = os.path.join(ROOT, rel)
        if not os.path.exists(p):
            continue
        try:
            rows = list(csv.reader(open(p)))
            hdr, units = rows[0], rows[1]
            ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            ik = hdr.index("Kernel Name") if "Kernel Name" in hdr else 0
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
            for r in rows[2:]:
                if kernel_substr in r[ik]:
                    return float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]], rel
        except Exception:
            continue
    return None, None


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
# workloads (synthetic)
def mps_positions(n_side, jitter=0.2, seed=1):
    rng = np.random.default_rng(seed)
    g = np.arange(n_side, dtype=np.float64)
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    pos = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1)
    pos += rng.uniform(-jitter, jitter, size=pos.shape)
    return pos


def host_problem(args):
    """(val, rowptr, col, b) as numpy arrays — only for workloads whose host copy is cheap to build (7-point grids <= 256^3)."""
    from conjugategradient_b200 import generators as g
    n = args.size
    assert args.config in ("c2", "c4")
    val, rp, col = g.poisson3d7(n)
    return val, rp, col, g.rhs_uniform(n ** 3, 0)


def device_problem(args, m, dev, zrange=None):
    """Assemble the workload (or the z-slab `zrange` of it) in device memory.  -> (tv, tr, tc, tb_full_or_local, info)."""
    import torch
    from conjugategradient_b200 import device_generators as dg
    from conjugategradient_b200 import generators as g
    n = args.size
    if args.config in ("c2", "c4"):
        tv, tr, tc = dg.poisson3d7(m.ctx, n, zrange=zrange, device=dev)
        b = g.rhs_uniform(n ** 3, 0)
    elif args.config == "c5":
        tv, tr, tc = dg.poisson3d27_varcoef(m.ctx, n, zrange=zrange, device=dev)
        b = g.rhs_uniform(n ** 3, 0)
    else:
        assert zrange is None
        pos = torch.from_numpy(mps_positions(n)).to(dev)
        t0 = time.perf_counter()
        try:       # the reference's container holds at most 160 entries per row (MgcgMain.cs:20)
            tv, tr, tc, order, _ = dg.mps_particles(m.ctx, pos, args.re, shift_abs=0.0, max_per_row=160)
            device_problem.row_cap = 160
        except Exception:
            tv, tr, tc, order, _ = dg.mps_particles(m.ctx, pos, args.re, shift_abs=0.0, max_per_row=1 << 20)
            device_problem.row_cap = None
        diag = tv[tr[:-1].long()]
        tv[tr[:-1].long()] = diag + 1e-3 * float(diag.mean())          # the stand-in for the free-surface Dirichlet rows
        torch.cuda.synchronize()
        device_problem.assembly_ms = (time.perf_counter() - t0) * 1e3
        del pos, order, diag
        b = g.rhs_uniform(n ** 3, 2)
    if zrange is not None:
        r0, r1 = zrange[0] * n * n, zrange[1] * n * n
        b = b[r0:r1].copy()
    return tv, tr, tc, torch.from_numpy(b).to(dev)


device_problem.assembly_ms = None
device_problem.row_cap = None


def solver_options(args, S, **extra):
    n = args.size
    kw = dict(precond=1, norm=2, tolerance=TOL, max_iteration=5000)
    if args.config != "c3":
        kw.update(nx=n, ny=n, nz=n)
    elif args.theta > 0:
        kw.update(mg_strength_theta=args.theta)
    if args.graph >= 0:
        kw["use_graph"] = args.graph
    if args.smoother >= 0:
        kw["mg_smoother"] = args.smoother
    for kv in args.opt:
        k, v = kv.split("=")
        kw[k] = float(v) if ("." in v or "e" in v.lower()) else int(v)
    kw.update(extra)
    return S.default_options(**kw)


def precond_text(args):
    if args.config == "c3":
        return "smoothed aggregation on strength-filtered MIS-2 aggregates (no grid given), V(1,1), weighted Jacobi 0.85, built on device"
    return "smoothed-aggregation V(1,1) on geometric aggregates (grid hint), weighted Jacobi 0.85, built on device"


# ---------------------------------------------------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
def cpu_sample(problem, iters, threads):
    """Run `iters` iterations of the reference CPU CG (oracle port); -> (seconds per iteration, iterations done)."""
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


def cpu_full_solve(problem, threads):
    """One COMPLETE solve to 1e-8 relative residual with the reference's CPU CG; -> (seconds, iterations performed, status)."""
    from oracle import oracle as o
    val, rp, col, b = problem
    t0 = time.perf_counter()
    _, it, _, st = o.cg_solve_mt(val, rp, col, b, tol=TOL, norm=o.NORM_REL_L2, threads=threads)
    return time.perf_counter() - t0, it + 1, st


def cpu_baseline_leg(problem, label, budget_s, gpu_plain_iterations):
    """cpu_baseline object.  A complete CPU solve when it fits the budget (nothing extrapolated); else a bounded sample scaled
    to the plain-CG iteration count measured on the GPU in this run."""
    from oracle import oracle as o
    cores = o.max_threads()
    per_it, _ = cpu_sample(problem, 3, cores)                  # calibration
    need = gpu_plain_iterations
    base = {"unit": "ms", "cores": cores, "kind": "port"}
    if need is not None and per_it * need <= budget_s:
        secs, done, st = cpu_full_solve(problem, cores)
        base.update({"value": secs * 1e3, "iterations": done, "ms_per_iteration": secs * 1e3 / done, "extrapolated": False,
                     "sample": f"one COMPLETE solve of the same {label} system by the reference's unpreconditioned CPU CG (oracle/cg_ref_mt.c, "
                               f"OpenMP over its commented-out Parallel.For loops, {cores} threads): {done} iterations to 1e-8 "
                               f"(GPU plain CG of this run: {need}); the reference has no multigrid (SURVEY F1)"})
        return base
    iters = max(3, min(200, int(budget_s / max(per_it, 1e-6))))
    per_it, done = cpu_sample(problem, iters, cores)
    how = "measured by this run's GPU plain-CG solve" if need is not None else "unknown: per-iteration time only"
    base.update({"value": per_it * need * 1e3 if need is not None else None, "ms_per_iteration": per_it * 1e3, "extrapolated": True,
                 "iterations": need,
                 "sample": f"{done} iterations of the reference's unpreconditioned CPU CG (oracle/cg_ref_mt.c, {cores} threads) on the same "
                           f"{label} system, {per_it * 1e3:.1f} ms/iteration, scaled to the {need} iterations plain CG needs for 1e-8 ({how})"})
    return base


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle port; the C#/boost/CUDA-5 sources cannot
    be compiled here) on the box's host cores, all threads.  The first warm-up step is one COMPLETE solve — it
    measures the iteration count plain CG needs; every other step times a bounded number of iterations and scales to it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as o
    if args.config not in ("c2", "c4") or args.size > 256:
        print(json.dumps({"impl": "reference", "unavailable": f"config {args.config} at size {args.size} has no host-assembled matrix for the CPU arm "
                          "(assembled on the device only); run --config c4 / c2"}), flush=True)
        return 0
    n = args.size
    problem = host_problem(args)
    cores = o.max_threads()
    t_full, need, st = cpu_full_solve(problem, cores)          # warm-up step 1: complete solve, measures the iteration count
    per_it_full = t_full / need
    total_steps = args.steps + args.warmup
    iters = max(3, min(100, int(90.0 / max(total_steps, 1) / max(per_it_full, 1e-6))))
    if t_full * (total_steps - 1) <= 100.0:        # every step a COMPLETE solve: nothing is scaled
        for _ in range(max(args.warmup - 1, 0)):
            cpu_full_solve(problem, cores)
        t = [cpu_full_solve(problem, cores)[0] for _ in range(args.steps)]
        value = float(np.mean(t)) * 1e3
        sample = (f"every step is one COMPLETE solve ({need} iterations to 1e-8) by the reference's unpreconditioned CPU CG "
                  f"(oracle/cg_ref_mt.c, {cores} threads) on {n}^3; nothing extrapolated")
    else:
        for _ in range(max(args.warmup - 1, 0)):
            cpu_sample(problem, iters, cores)
        t = []
        for _ in range(args.steps):
            p, done = cpu_sample(problem, iters, cores)
            t.append(p)
        per_it = float(np.mean(t))
        value = per_it * need * 1e3
        sample = (f"{iters} iterations per step of the reference's unpreconditioned CPU CG (oracle/cg_ref_mt.c, {cores} threads) on {n}^3, "
                  f"{per_it * 1e3:.1f} ms/iteration x {need} iterations to 1e-8; the iteration count is MEASURED by one complete CPU solve in "
                  f"this run (first warm-up step: {t_full:.1f} s = {t_full * 1e3 / need:.1f} ms/iteration, status {st})")
    out = {"impl": "reference", "metric": metric_name(args.config), "value": value, "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": value, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": {"workload": workload_text(args).replace(" MGCG", ""), "rows": n ** 3, "nnz": int(problem[1][-1]),
                      "preconditioner": "none (the reference has none)"},
           "plain_cg_iterations": need, "complete_solve_ms": t_full * 1e3,
           "cpu_baseline": {"value": value, "unit": "ms", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": value, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------------------------
def gpu_plain_cg_iterations(S, m, tv, tr, tc, tb, tx, opts_mg):
    """Iterations the reference's algorithm (unpreconditioned CG) needs for 1e-8 on this system, measured on the GPU outside the
    timed region (its +-1 parity with the CPU oracle is tested in tests/test_gpu_cg.py and tests/test_gpu_parity_c2.py)."""
    o = S.default_options(precond=0, norm=2, tolerance=TOL, max_iteration=200000)
    m.setup(tv, tr, tc, o)
    tx.zero_()
    res = m.solve(tb, tx, raise_on_fail=False)
    out = (res.iterations_done, res.solve_ms, tx.clone()) if res.status == 0 else (None, res.solve_ms, None)
    m.setup(tv, tr, tc, opts_mg)
    return out


def run_ours_single(args):
    """N = 1: the whole system on cuda:0 through the C ABI (mgcg_setup / mgcg_solve / mgcg_solve_host)."""
    import torch
    from conjugategradient_b200 import solver as S
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    if args.config == "c5" and args.size > 256:
        raise SystemExit("config c5 (27-point 512^3: 3.6e9 non-zeros) needs --gpus >= 2; per-rank shards must stay below 2^31 entries "
                         "(--size 256 runs the same operator on one GPU)")
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    torch.cuda.set_stream(torch.cuda.Stream(dev))   # a real (capturable) stream shared by torch's events and the library
    n_side = args.size
    n = n_side ** 3
    m = S.Mgcg(0, use_torch_stream=True)     # run on torch's current stream so torch.cuda.Event brackets the library's work
    device_gen = args.device_gen or args.config in ("c3", "c5") or n_side > 256
    problem = None
    if device_gen:
        tv, tr, tc, tb = device_problem(args, m, dev)
        nnz = tv.numel()
        if not args.no_e2e or not args.no_cpu:     # the host legs need a host copy: bring the device-assembled system back once
            hv, hr, hc, hb = (torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in (tv, tr, tc, tb))
            for h, t in ((hv, tv), (hr, tr), (hc, tc), (hb, tb)):
                h.copy_(t)
            problem = (hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy())
    else:
        problem = host_problem(args)
        val, rp, col, b = problem
        nnz = int(rp[-1])
        hv, hr, hc, hb = (torch.from_numpy(a).pin_memory() for a in (val, rp, col, b))   # pinned host copies (the e2e leg uploads from these)
        tv, tr, tc, tb = (t.to(dev, non_blocking=True) for t in (hv, hr, hc, hb))
    hx = torch.zeros(n, dtype=torch.float64).pin_memory()
    tx = torch.zeros(n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    opts = solver_options(args, S)
    t0 = time.perf_counter()
    m.setup(tv, tr, tc, opts)
    torch.cuda.synchronize()
    setup_cold_ms = (time.perf_counter() - t0) * 1e3            # first hierarchy build of the process: allocations included

    def step():
        tx.zero_()
        return m.solve(tb, tx)

    t0 = time.perf_counter()
    res = step()
    torch.cuda.synchronize()
    solve_cold_ms = (time.perf_counter() - t0) * 1e3            # first solve: graph capture + instantiation included
    for _ in range(max(args.warmup - 1, 0)):
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
    x_timed = tx.clone()
    # steady-state setup (a second build reuses the cached blocks)
    m.setup(tv, tr, tc, opts)
    setup_ms = step().setup_ms
    # true residual check of the last solve (not timed): ||b - A x|| / ||b||.  kernel = MGCG_SPMV_VECTOR (2) on purpose: it differs from
    # the set-up's kernel option, so abi.cu's same_as_setup rule does not reroute the call to the coded-column / STENCIL kernel the
    # solve ran on — the check reads the caller's int32 columns with a different kernel family than the one under test
    ty = torch.empty_like(tx)
    S.check(m.L.mgcg_spmv(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(x_timed), S._p(ty), 1.0, 0.0, 2), "mgcg_spmv")
    torch.cuda.synchronize()
    true_rel = float(torch.linalg.vector_norm(tb - ty) / torch.linalg.vector_norm(tb))

    # parity block (not timed)
    parity, plain_it = None, None
    if not args.no_parity:
        tight = solver_options(args, S, tolerance=1e-12)
        m.setup(tv, tr, tc, tight)
        tx.zero_()
        rt = m.solve(tb, tx)
        x_tight = tx.clone()
        scale = float(x_tight.abs().max())
        m.setup(tv, tr, tc, opts)
        plain_it, plain_ms, x_plain = gpu_plain_cg_iterations(S, m, tv, tr, tc, tb, tx, opts)
        parity = {"x_rel_err_vs_tight_solve": float((x_timed - x_tight).abs().max()) / scale,
                  "reference_solution": "1e-12 relative-residual MGCG solve of the same system on the same device (not the CPU twin: it takes "
                                        "minutes at this size; twin parity is tested at 128^3 in tests/test_gpu_parity_c2.py and recorded at 256^3 by "
                                        "tools/parity_c4.py in profiles/r03_parity_c4.json: hierarchy bit-exact, iteration 15 == 15, x 2.3e-15)",
                  "tight_iterations": rt.iterations_done, "tight_relative_residual": rt.residual,
                  "plain_cg_iterations_gpu": plain_it, "plain_cg_solve_ms_gpu": plain_ms,
                  "x_rel_err_plain_cg_vs_tight_solve": (float((x_plain - x_tight).abs().max()) / scale) if x_plain is not None else None,
                  "norm": "max|dx| / max|x|"}
        del x_tight, x_plain

    # roofline leg: the fine-level SpMV(+p.Ap) launches timed with CUDA events inside one more solve (profile=1)
    popts = solver_options(args, S, profile=1, cg_variant=0)    # (the product kernel is the same in both CG recurrences)
    m.setup(tv, tr, tc, popts)
    tx.zero_()
    pres = m.solve(tb, tx)
    spmv_ms = pres.spmv_ms_total / max(pres.spmv_launches, 1)
    # ... and back-to-back in isolation (inputs >> 126 MB L2, so no flush is needed)
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
    kname = pres.reserved[5]
    coded = (kname == 4)        # coded columns (csrc/spmv_code.cu; --opt spmv_coding=1 switches them off): 8 + 1 instead of 8 + 4 bytes per entry
    csr_bytes = 12.0 * nnz + 20.0 * n
    spmv_bytes = (9.0 if coded else 12.0) * nnz + 20.0 * n
    peak, peak_how = measured_peak()
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9
    kernel_text = {1: "k_spmv_tma<EPI_DOT> (fine-level A*p with fused p.Ap, TMA-staged stream kernel)",
                   4: "k_spmv_stencil<EPI_DOT> (fine-level A*p with fused p.Ap on coded columns: values, 1-byte codes and the x runs of "
                      "every distance staged by TMA, 9 B per entry instead of CSR's 12)",
                   2: "k_spmv_vector<EPI_DOT> (fine-level A*p with fused p.Ap, lanes-per-row kernel)",
                   3: "k_spmv_rowblock<EPI_DOT> (fine-level A*p with fused p.Ap, TMA-staged long-row kernel)",
                   0: "k_spmv_stream<EPI_DOT>"}.get(kname, "fine-level SpMV + p.Ap")
    traffic, traffic_src = ncu_traffic("k_spmv_stencil<2", args, (NCU_FULL_STENCIL,)) if coded else ncu_traffic("k_spmv_tma<2", args)
    roofline = {"bound": "hbm", "kernel": kernel_text, "achieved": achieved,
                "layout": "coded columns (1 B) + fp64 values" if coded else "CSR (int32 columns + fp64 values)",
                "bytes_rule": ("9 nnz + 20 N (the layout the kernel reads)" if coded else "12 nnz + 20 N (SURVEY 8d)"),
                "csr_bytes_per_launch": csr_bytes, "csr_equivalent_gbs": csr_bytes / (spmv_ms * 1e-3) / 1e9,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_8TBs_nominal": achieved / 8000.0,
                "traffic": traffic,
                "traffic_source": f"dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full ({traffic_src})" if traffic else None,
                "peak_source": peak_how, "bytes_per_launch": spmv_bytes, "ms_per_launch": spmv_ms,
                "launches_timed": pres.spmv_launches, "isolated_ms_per_launch": spmv_iso_ms,
                "isolated_gbs": spmv_bytes / (spmv_iso_ms * 1e-3) / 1e9}
    m.setup(tv, tr, tc, opts)

    # e2e leg: host CSR + b in, x out, everything (upload, hierarchy, solve, download) inside the timed region
    e2e = None
    if not args.no_e2e:
        m2 = S.Mgcg(0, use_torch_stream=True)
        hx.zero_()
        t0 = time.perf_counter()
        rc = m2.solve_host(hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx.numpy(), opts)     # the first call of a fresh context
        torch.cuda.synchronize()
        e2e_cold = (time.perf_counter() - t0) * 1e3
        k = max(1, min(args.steps, 3))
        # x is in/out (x0 = 0 in, solution out): one zeroed pinned buffer per timed call, prepared outside the timed region like the
        # caller's other inputs (clearing 8 N bytes of host memory is not part of the solve)
        hxs = [torch.zeros_like(hx).pin_memory() for _ in range(k)]
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        g0.record()
        for i in range(k):
            r2 = m2.solve_host(hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hxs[i].numpy(), opts)
        g1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 / k
        e2e = {"value": max(g0.elapsed_time(g1) / k, wall), "unit": "ms",
               "h2d_bytes_per_step": int(12 * nnz + 4 * (n + 1) + 16 * n), "d2h_bytes_per_step": int(8 * n),
               "setup_ms": r2.setup_ms, "solve_ms": r2.solve_ms, "iterations": r2.iterations_done,
               "e2e_cold_ms": e2e_cold, "cold_iterations": rc.iterations_done,
               "x_rel_err_vs_device_path": float((hxs[-1].to(dev) - x_timed).abs().max() / x_timed.abs().max()),
               "api": "mgcg_solve_host (include/mgcg.h): pinned host CSR + b + x0 -> x"}
        assert abs(r2.iterations_done - iters) <= 1
        m2.close()

    label = f"{args.config} {n_side}^3"
    cpu = None if (args.no_cpu or problem is None) else cpu_baseline_leg(problem, label, args.cpu_seconds, plain_it)
    x512 = None
    if args.config == "c4" and n_side == 256 and not args.no_extra_512:
        try:
            x512 = extra_512_single(args, S, dev)
        except Exception as e:      # the extra leg never takes the headline line down
            x512 = {"error": repr(e)[:300]}
    out = {"metric": metric_name(args.config), "value": ms, "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic",
           "config": {"workload": workload_text(args), "name": args.config,
                      "rows": n, "nnz": nnz, "preconditioner": precond_text(args),
                      "l2": f"inputs ({12 * nnz / 1e9:.2f} GB matrix + {8 * n / 2 ** 20:.0f} MiB vectors) larger than the 126 MB L2; no flush",
                      "matrix_assembly": "device" if device_gen else "host numpy, uploaded",
                      "parallelism": "1 GPU"},
           "iterations": iters, "iterations_per_s": iters / (ms * 1e-3), "final_relative_residual": res.residual,
           "true_relative_residual": true_rel, "true_residual_kernel": "VECTOR on int32 columns (not the solve's coded kernel)", "setup_ms": setup_ms, "setup_cold_ms": setup_cold_ms, "first_solve_ms": solve_cold_ms,
           "levels": res.levels, "operator_complexity": res.operator_complexity, "spmv_gbs": achieved,
           "assembly_ms": device_problem.assembly_ms,
           "gpu_launches": launches, "clocks": clk, "roofline": roofline, "parity": parity, "cpu_baseline": cpu, "e2e": e2e, "extra_512": x512}
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


def dist_vs_twin(S, m, dist, dev, rank, world, extra):
    """Oracle leg of the sharded path (not timed; one of the places that execute oracle/): two small systems — 48^3 7-point and
    24^3 27-point variable-coefficient — solved sharded over ALL the ranks of this run and, on rank 0's host cores, by the CPU twin
    oracle/mg_ref.c (MgRef.pcg, the restatement the single-GPU path is pinned to).  Iteration index at 1e-8 on both sides, and the
    solutions of both taken to 1e-12 compared as max|dx| / max|x|.  The reference compares its multi-device solve with its CPU solve
    the same way (Mgcg/cuBlas/Mgcg/MgcgMain.cs:151-162)."""
    import torch
    from conjugategradient_b200 import generators as g
    out = {}
    for name, ns, make in (("poisson3d7_48", 48, g.poisson3d7), ("poisson3d27_varcoef_24", 24, g.poisson3d27_varcoef)):
        try:
            v, rp, c = make(ns)
            N = ns ** 3
            b = g.rhs_uniform(N, 0)
            planes = [ns // world] * world
            planes[-1] += ns - sum(planes)
            zo = np.concatenate([[0], np.cumsum(planes)]) * ns * ns
            r0, r1 = int(zo[rank]), int(zo[rank + 1])
            k0, k1 = int(rp[r0]), int(rp[r1])
            tv, tc = torch.as_tensor(v[k0:k1]).to(dev), torch.as_tensor(c[k0:k1]).to(dev)
            tr = torch.as_tensor((rp[r0:r1 + 1] - rp[r0]).astype(np.int32)).to(dev)
            tb = torch.as_tensor(b[r0:r1]).to(dev)
            tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
            rec = {}
            xs = {}
            for tol in (1e-8, 1e-12):
                m.setup(N, r0, tv, tr, tc, S.default_options(precond=1, nx=ns, ny=ns, nz=ns, norm=2, tolerance=tol, max_iteration=2000, **extra))
                res = m.solve(tb, tx.zero_())
                torch.cuda.synchronize()
                parts = [None] * world
                dist.all_gather_object(parts, tx.cpu().numpy())
                xs[tol] = np.concatenate(parts)
                rec[tol] = (res.iteration, res.residual)
            if rank == 0:
                from oracle import oracle as o
                H = o.MgRef(v, rp, c, ns, ns, ns)
                _, it8, res8, st8 = H.pcg(b, tol=1e-8, norm=o.NORM_REL_L2, max_iteration=2000)
                xt, it12, res12, st12 = H.pcg(b, tol=1e-12, norm=o.NORM_REL_L2, max_iteration=2000)
                H.close()
                err = float(np.abs(xs[1e-12] - xt).max() / np.abs(xt).max())
                out[name] = {"it_dist": rec[1e-8][0], "it_twin": it8, "residual_dist": rec[1e-8][1], "residual_twin": res8,
                             "it_dist_1e-12": rec[1e-12][0], "it_twin_1e-12": it12, "x_rel_err": err, "twin_status": [st8, st12],
                             "pass": bool(abs(rec[1e-8][0] - it8) <= 1 and err <= 1e-10)}
        except Exception as e:      # a failed check is reported in the line, it does not take the measurement down
            out[name] = {"error": repr(e)[:300], "pass": False}
    if rank == 0:
        out["what"] = (f"sharded over {world} ranks vs the CPU twin oracle/mg_ref.c on rank 0's host: 0-based iteration index at 1e-8 (bar +-1), "
                       "max|dx|/max|x| of the two 1e-12 solves (bar 1e-10)")
    return out


def _args_512(args):
    a2 = argparse.Namespace(**vars(args))
    a2.config, a2.size = "c4", 512
    return a2


def extra_512_single(args, S, dev):
    """north_star's scaling target (7-point 512^3, T1 / T8 >= 5) under the same clock as the headline: the 512^3 system assembled on the
    device, set up and solved to 1e-8 on ONE GPU; device-timed like `value` (CUDA events, 2 warm-up + 3 timed solves).  Not part of
    `value`; reported as `extra_512` in the same JSON line."""
    import torch
    a2 = _args_512(args)
    n = a2.size ** 3
    m = S.Mgcg(0, use_torch_stream=True)
    try:
        tv, tr, tc, tb = device_problem(a2, m, dev)
        nnz = tv.numel()
        tx = torch.zeros(n, dtype=torch.float64, device=dev)
        opts = solver_options(a2, S)
        m.setup(tv, tr, tc, opts)
        for _ in range(2):
            res = m.solve(tb, tx.zero_())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            tx.zero_()
            res = m.solve(tb, tx)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        m.setup(tv, tr, tc, opts)
        setup_ms = m.solve(tb, tx.zero_()).setup_ms
        ty = torch.empty_like(tx)
        S.check(m.L.mgcg_spmv(m.ctx, n, n, nnz, S._p(tv), S._p(tr), S._p(tc), S._p(tx), S._p(ty), 1.0, 0.0, 2), "mgcg_spmv")   # int32 columns, VECTOR kernel
        torch.cuda.synchronize()
        true_rel = float(torch.linalg.vector_norm(tb - ty) / torch.linalg.vector_norm(tb))
        return {"workload": workload_text(a2), "rows": n, "nnz": nnz, "n_gpus": 1, "value": ms, "unit": "ms", "steps": 3, "warmup": 2,
                "iterations": res.iterations_done, "final_relative_residual": res.residual, "true_relative_residual": true_rel,
                "setup_ms": setup_ms, "levels": res.levels}
    finally:
        m.close()


def extra_512_dist(args, S, m, dist, dev, rank, world, lr, extra):
    """The same 512^3 system row-block sharded over the ranks of this run (see extra_512_single): max over ranks of the device time of
    3 solves after 2 warm-ups; x against the single-GPU solve of the same system on rank 0's GPU, both taken to 1e-12."""
    import torch
    a2 = _args_512(args)
    ns = a2.size
    n = ns ** 3
    planes = [ns // world] * world
    planes[-1] += ns - sum(planes)
    zo = np.concatenate([[0], np.cumsum(planes)])
    z0, z1 = int(zo[rank]), int(zo[rank + 1])
    r0, r1 = z0 * ns * ns, z1 * ns * ns
    tv, tr, tc, tb = device_problem(a2, m, dev, zrange=(z0, z1))
    tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
    opts = solver_options(a2, S, **extra)
    m.setup(n, r0, tv, tr, tc, opts)
    for _ in range(2):
        res = m.solve(tb, tx.zero_())
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        tx.zero_()
        res = m.solve(tb, tx)
    e1.record()
    torch.cuda.synchronize(); dist.barrier()
    m.setup(n, r0, tv, tr, tc, opts)
    setup_ms = m.solve(tb, tx.zero_()).setup_ms
    t = torch.tensor([e0.elapsed_time(e1) / 3, setup_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, setup_ms = (float(v) for v in t.tolist())
    m.setup(n, r0, tv, tr, tc, solver_options(a2, S, tolerance=1e-12, **extra))
    rt = m.solve(tb, tx.zero_())
    torch.cuda.synchronize()
    if rank == 0:
        parts = [tx] + [torch.empty(int(zo[q + 1] - zo[q]) * ns * ns, dtype=torch.float64, device=dev) for q in range(1, world)]
        for r_ in [dist.irecv(parts[q], src=q) for q in range(1, world)]:
            r_.wait()
        xd = torch.cat(parts)
        del parts
    else:
        dist.send(tx, dst=0)
    out = None
    if rank == 0:
        err = None
        ms1 = S.Mgcg(lr, use_torch_stream=True)
        try:
            fv, fr, fc, fb = device_problem(a2, ms1, dev)
            ft = torch.zeros(n, dtype=torch.float64, device=dev)
            ms1.setup(fv, fr, fc, solver_options(a2, S, tolerance=1e-12))
            r1g = ms1.solve(fb, ft)
            err = float((xd - ft).abs().max() / ft.abs().max())
            it1 = r1g.iterations_done
        finally:
            ms1.close()
        out = {"workload": workload_text(a2), "rows": n, "n_gpus": world, "value": ms, "unit": "ms", "steps": 3, "warmup": 2,
               "iterations": res.iterations_done, "final_relative_residual": res.residual, "setup_ms": setup_ms, "levels": res.levels,
               "distributed_levels": res.reserved[2], "dist_x_rel_err": err,
               "dist_x_rel_err_note": f"sharded vs single-GPU solution on rank 0's GPU, both solved to 1e-12 relative residual ({rt.iterations_done} / {it1} iterations); max|dx| / max|x|"}
    dist.barrier()
    return out


def shard_planes(n_side, world, align=0):
    """z-planes per rank.  align <= 1: the reference's rule, offsets[i] = offsets[i-1] + floor(Count / D), remainder to the last device
    (ConjugateGradientParallelGpu.cs:271-277; the same thing in planes when world divides n_side).  align = K > 1: every cut falls on
    a multiple of K planes, so that geometric aggregates (3 planes on level 0, 2 on the levels below: K = 6 keeps two levels whole,
    K = 12 three) are not clipped by a shard boundary; the K-plane units are dealt out as evenly as they go, the remainder planes
    go to the last rank."""
    if align <= 1 or n_side // align < world:
        planes = [n_side // world] * world
    else:
        units = n_side // align
        planes = [(units // world + (1 if q < units % world else 0)) * align for q in range(world)]
    planes[-1] += n_side - sum(planes)
    return planes


def run_ours_dist(args):
    """N > 1 (torchrun, one rank per GPU): the SAME system row-block sharded over the ranks (strong scaling): z-slab shards,
    halo exchange of p and of the multigrid level vectors plus rank-ordered scalar reductions over NVLink."""
    import torch
    import torch.distributed as dist
    from conjugategradient_b200 import solver as S
    if args.config == "c3":
        raise SystemExit("config c3 (particle matrix, MIS-2 aggregation) is a single-GPU configuration (BASELINE.json); run it with --gpus 1")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs CUDA devices (there is no CPU fallback)"
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    torch.cuda.set_stream(torch.cuda.Stream(dev))
    n_side = args.size
    n = n_side ** 3
    planes = shard_planes(n_side, world, args.align_planes)
    zo = np.concatenate([[0], np.cumsum(planes)])
    z0, z1 = int(zo[rank]), int(zo[rank + 1])
    r0, r1 = z0 * n_side * n_side, z1 * n_side * n_side
    ids = [S.MgcgDist.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    t0 = time.perf_counter()
    m = S.MgcgDist(lr, rank, world, ids[0], use_torch_stream=True)
    comm_init_ms = (time.perf_counter() - t0) * 1e3
    device_gen = args.device_gen or args.config == "c5" or n_side > 256
    if device_gen:
        tv, tr, tc, tb = device_problem(args, m, dev, zrange=(z0, z1))
        nnz_local = tv.numel()
        args.no_e2e = True
    else:
        from conjugategradient_b200 import generators as g
        val, rp, col = g.poisson3d7(n_side, zrange=(z0, z1))
        b = g.rhs_uniform(n, 0)[r0:r1].copy()
        nnz_local = int(rp[-1])
        hv, hr, hc, hb = (torch.from_numpy(a).pin_memory() for a in (val, rp, col, b))
        tv, tr, tc, tb = (t.to(dev, non_blocking=True) for t in (hv, hr, hc, hb))
    tx = torch.zeros(r1 - r0, dtype=torch.float64, device=dev)
    extra = {"dist_transport": args.transport}
    if args.replicate_below > 0:
        extra["dist_replicate_below"] = args.replicate_below
    opts = solver_options(args, S, **extra)
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    m.setup(n, r0, tv, tr, tc, opts)
    torch.cuda.synchronize()
    setup_cold_ms = (time.perf_counter() - t0) * 1e3            # first sharded setup: peer arena + IPC mappings included

    def step():
        tx.zero_()
        return m.solve(tb, tx)

    t0 = time.perf_counter()
    res = step()
    torch.cuda.synchronize()
    solve_cold_ms = (time.perf_counter() - t0) * 1e3
    for _ in range(max(args.warmup - 1, 0)):
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
    t = torch.tensor([e0.elapsed_time(e1) / args.steps, setup_cold_ms, solve_cold_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, setup_cold_ms, solve_cold_ms = (float(v) for v in t.tolist())
    clk = clocks.stop() if rank == 0 else None
    x_timed = tx.clone()
    m.setup(n, r0, tv, tr, tc, opts)             # steady-state setup time (arena, mappings and cached blocks reused)
    setup_ms = step().setup_ms
    micro = (C.c_double * 3)()
    S.check(m.L.mgcg_dist_microbench(m.ctx, 200, micro), "mgcg_dist_microbench")
    # local SpMV(+p.Ap) in isolation on this rank's shard (x = a [lo|own|hi]-sized vector addressed globally, like p)
    L = m.L
    halo = n_side * n_side + (n_side + 1 if args.config == "c5" else 0)     # ghost rows either side: one plane (+ one row + one cell, 27-point)
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
    ndict = C.c_int(0)
    L.mgcg_matrix_codes(m.ctx, C.byref(ndict), None, None)
    coded_used = ndict.value > 0      # (the coded-column kernels take any fp64-aligned x, also a shard's odd-shifted view)
    spmv_bytes = (9.0 if coded_used else 12.0) * nnz_local + 20.0 * (r1 - r0)
    peak, peak_how = measured_peak()
    achieved = spmv_bytes / (spmv_ms * 1e-3) / 1e9
    del tp, ty

    # parity: the sharded x against the single-GPU solve of the SAME system, run on rank 0's GPU after the timed region
    parity = None
    nnz_global = torch.tensor([nnz_local], dtype=torch.int64, device=dev)
    dist.all_reduce(nnz_global)
    nnz_global = int(nnz_global.item())
    if not args.no_parity and nnz_global < 2 ** 31:
        def gather_to_rank0(x_local):
            parts = [torch.empty(int(zo[q + 1] - zo[q]) * n_side * n_side, dtype=torch.float64, device=dev) for q in range(world)] if rank == 0 else None
            if rank == 0:
                parts[0].copy_(x_local)
                reqs = [dist.irecv(parts[q], src=q) for q in range(1, world)]
                for r_ in reqs:
                    r_.wait()
                return torch.cat(parts)
            dist.send(x_local, dst=0)
            return None
        xd = gather_to_rank0(x_timed)
        # the sharded and the single-GPU hierarchies differ where aggregates meet a shard boundary, so two solves stopped at 1e-8
        # differ by O(1e-8); the sharded PATH is judged on solves of the same system taken to 1e-12 on both sides
        m.setup(n, r0, tv, tr, tc, solver_options(args, S, tolerance=1e-12, **extra))
        tx.zero_()
        rt_d = m.solve(tb, tx)
        xdt = gather_to_rank0(tx)
        m.setup(n, r0, tv, tr, tc, opts)
        if rank == 0:
            ms1 = S.Mgcg(lr, use_torch_stream=True)
            fv, fr, fc, fb = device_problem(args, ms1, dev)
            fx = torch.zeros(n, dtype=torch.float64, device=dev)
            o1 = solver_options(args, S)
            ms1.setup(fv, fr, fc, o1)
            r1g = ms1.solve(fb, fx)
            scale = float(fx.abs().max())
            err = float((xd - fx).abs().max()) / scale
            ms1.setup(fv, fr, fc, solver_options(args, S, tolerance=1e-12))
            ft = torch.zeros(n, dtype=torch.float64, device=dev)
            ms1.solve(fb, ft)
            parity = {"dist_x_rel_err": float((xdt - ft).abs().max()) / scale,
                      "dist_x_rel_err_note": "sharded vs single-GPU solution of the same system, both solved to 1e-12 relative residual "
                                             f"(sharded: {rt_d.iterations_done} iterations, residual {rt_d.residual:.2e})",
                      "dist_x_rel_err_at_1e-8": err, "single_gpu_iterations": r1g.iterations_done, "single_gpu_solve_ms": r1g.solve_ms,
                      "dist_x_rel_err_vs_tight_solve": float((xd - ft).abs().max()) / scale,
                      "single_x_rel_err_vs_tight_solve": float((fx - ft).abs().max()) / scale,
                      "norm": "max|dx| / max|x|; single-GPU solves of the same device-assembled system on rank 0's GPU, outside the timed region"}
            ms1.close()
            del xd, xdt, fx, ft, fv, fr, fc, fb
        dist.barrier()
    if not args.no_parity:
        twin = dist_vs_twin(S, m, dist, dev, rank, world, extra)
        if rank == 0:
            parity = dict(parity or {})
            parity["dist_vs_twin"] = twin
        m.setup(n, r0, tv, tr, tc, opts)
        dist.barrier()

    # e2e: ONE C-ABI call per rank on pinned host shards (mgcg_solve_host_dist): upload + hierarchy + solve + download timed
    e2e = None
    if not args.no_e2e:
        hx = torch.zeros(r1 - r0, dtype=torch.float64).pin_memory()
        k = max(1, min(args.steps, 3))
        m.solve_host(n, r0, hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx.numpy(), opts)      # allocates the context's device copies
        dist.barrier()
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        g0.record()
        for _ in range(k):
            hx.zero_()
            r2 = m.solve_host(n, r0, hv.numpy(), hr.numpy(), hc.numpy(), hb.numpy(), hx.numpy(), opts)
        g1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 / k
        t2 = torch.tensor([max(g0.elapsed_time(g1) / k, wall)], dtype=torch.float64, device=dev)
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": float(t2.item()), "unit": "ms", "h2d_bytes_per_step": int(12 * nnz_local + 4 * (r1 - r0 + 1) + 16 * (r1 - r0)),
               "d2h_bytes_per_step": int(8 * (r1 - r0)), "setup_ms": r2.setup_ms, "solve_ms": r2.solve_ms, "iterations": r2.iterations_done,
               "e2e_cold_ms": comm_init_ms + setup_cold_ms + solve_cold_ms,
               "e2e_cold_note": "communicator init + first sharded setup (peer arena, IPC mappings) + first solve (graph capture), wall clock, max over ranks",
               "api": "mgcg_solve_host_dist (include/mgcg.h): one call per rank on its pinned host row block; bytes are per rank"}
    x512 = None
    if args.config == "c4" and n_side == 256 and not args.no_extra_512:
        try:
            x512 = extra_512_dist(args, S, m, dist, dev, rank, world, lr, extra)
        except Exception as e:      # (an exception on one rank only would leave the others in a collective: the legs above ran the same calls)
            x512 = {"error": repr(e)[:300]}
    if rank == 0:
        out = {"metric": metric_name(args.config), "value": ms, "unit": "ms", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": {"workload": workload_text(args), "name": args.config,
                          "rows": n, "nnz": nnz_global,
                          "preconditioner": precond_text(args) + "; coarse levels replicated",
                          "l2": "per-rank shard (matrix + vectors) larger than the 126 MB L2 at 256^3 up to 8 ranks; no flush",
                          "planes_per_rank": planes,
                          "parallelism": f"row-block (z-slab) shards over {world} GPUs; halo exchange, coarse all-gather and rank-ordered "
                                         "scalar reductions by " + ("the library's own kernels over NVLink peer memory, iteration replayed as a CUDA graph"
                                                                    if res.reserved[3] == 2 else "NCCL send/recv + all-gather")},
               "transport": "peer" if res.reserved[3] == 2 else "nccl", "graph": int(res.reserved[4]),
               "exchange_us": {"halo_level0": micro[0], "coarse_allgather": micro[1], "scalar_reduction": micro[2]},
               "iterations": res.iterations_done, "iterations_per_s": res.iterations_done / (ms * 1e-3),
               "final_relative_residual": res.residual, "setup_ms": setup_ms, "setup_cold_ms": setup_cold_ms, "first_solve_ms": solve_cold_ms,
               "comm_init_ms": comm_init_ms, "levels": res.levels,
               "distributed_levels": res.reserved[2], "halo_exchanges_per_solve": res.reserved[0], "allgathers_per_solve": res.reserved[1],
               "spmv_gbs": achieved * world, "gpu_launches": launches, "clocks": clk,
               "roofline": {"bound": "hbm", "kernel": ("k_spmv_stencil<EPI_DOT> (coded columns, 9 B per entry)" if coded_used else "fine-level SpMV + p.Ap kernel (int32 columns)") + " on rank 0's shard (isolated, back to back)",
                            "bytes_rule": "9 nnz + 20 N" if coded_used else "12 nnz + 20 N", "achieved": achieved,
                            "peak": peak, "unit": "GB/s", "frac": achieved / peak, "frac_of_8TBs_nominal": achieved / 8000.0, "traffic": None,
                            "peak_source": peak_how, "bytes_per_launch": spmv_bytes, "ms_per_launch": spmv_ms},
               "parity": parity, "cpu_baseline": None, "e2e": e2e, "extra_512": x512}
        print(json.dumps(out), flush=True)
    m.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
