#!/usr/bin/env python
"""Benchmark of the theta-step hot path on BASELINE.json's batched configuration (cfg 3):
65,536 independent 1,024-cell problems per GPU, exponential fitting, zero-flux Robin ends,
Crank-Nicolson, per-problem a/d/dt.

    python bench.py [--gpus N] [--steps K] [--warmup W]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference      # the reference's numpy/scipy loop on the host cores

One bench "step" = cfg 3's whole run for this GPU's problems: ONE launch of the fused kernel
advances every problem by --inner = 10,000 theta steps and writes the state out every
--checkpoint-every = 1,000 steps (10 checkpoints of 512 MiB).  Metric: fp64 cell-updates/s (one
cell advanced by one theta step), whole job over all ranks.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fp64 cell-updates/sec (theta-step)"
UNIT = "cell-updates/s"
# algorithmic work per cell-update of the fused on-chip variant
FLOP_PER_CELL_UPDATE = 11.0          # RHS 3 FMA, forward 1 FMA, backward 1 FMA + 1 MUL
FP64_INSTR_PER_CELL_UPDATE = 167.0 / 32.0   # what the on-chip Crank-Nicolson solver issues per lane and step (SASS of the loop): 3 per
                                            # cell for the sweeps + 2 spike-correction FMAs (31 cells), 10 for the separator system, 2 for its rhs
DFMA_DISTINCT_OPERANDS_PER_CLK_SM = 44.0   # measured: profiles/microbench/r01_fp64_operands_b200.txt
ONCHIP_B_PER_CELL_UPDATE = 40.0      # five factored coefficients read once per step (32 B TMEM + 8 B registers / shared memory)
HBM_B_PER_CELL_UPDATE_THOMAS = 40.0  # three per-cell-per-problem arrays streamed (24 B) + state in/out (16 B)
HBM_B_MOVED_THOMAS = 64.0            # what thomas_steps_kernel really moves: w, l, kappa, s in, t out (40) + t, s in, w out (24)
HBM_B_PER_CELL_UPDATE_SPIKE = 40.0   # variant 3: w, p, q, b in; g out (three per-cell arrays streamed)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--problems", type=int, default=65536, help="problems per GPU")
    ap.add_argument("--cells", type=int, default=1024)
    ap.add_argument("--inner", type=int, default=None,
                    help="theta steps per bench step; default: the workload's own run length "
                         "(cfg 3: 10,000 in one launch; cfg 4: 10; cfg 5: 20)")
    ap.add_argument("--checkpoint-every", type=int, default=1000,
                    help="cfg 3: the state is written out every this many steps (0: never)")
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg4", "cfg5"],
                    help="cfg3 (default, the headline): batched on-chip; cfg4: batched long systems; "
                         "cfg5: one mesh of 2^27 cells per GPU split over the ranks")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=None, help="groups run_host cuts the batch into (default: its own)")
    ap.add_argument("--no-long-legs", action="store_true", help="skip the variant-3 legs (cfg 5 / cfg 4) of the default run")
    return ap.parse_args()


def config(args, extra=None):
    cfg = {"workload": "cfg3 batched sweep: %d problems/GPU x %d cells, exponential fitting, zero-flux "
                       "Robin, Crank-Nicolson theta=0.5, per-problem a/d/dt" % (args.problems, args.cells),
           "problems_per_gpu": args.problems, "cells": args.cells, "theta": 0.5,
           "theta_steps_per_bench_step": args.inner, "checkpoint_every": getattr(args, "checkpoint_every", 0),
           "sharding": "independent problems, no communication",
           "l2": "state (%.0f MiB/GPU) exceeds the 126 MB L2; no flush needed" %
                 (args.problems * args.cells * 8 / 2 ** 20)}
    if extra:
        cfg.update(extra)
    return cfg


# ------------------------------------------------------------------------------------------
# CPU side: the reference's algorithm (oracle port = literal numpy/scipy loop of fvm_classes.py:408-411)
def _cpu_worker(job):
    os.environ["OMP_NUM_THREADS"] = "1"
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    from oracle import fvm_oracle as O
    from fvm_b200 import workloads as W
    first, count, cells, nsteps = job
    wl = W.batched_sweep(65536, cells, first=first, count=count)
    t0 = time.perf_counter()
    for p in range(count):
        asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", **wl["bc"])
        O.theta_run_scipy(asm, wl["k"][p], 0.5, wl["w0"][p], nsteps, every=nsteps)
    return time.perf_counter() - t0


def cpu_baseline_single(cells, budget_s=12.0):
    """1 core, bounded sample of cfg 3: first problems of the batch, 200 steps each."""
    from oracle import fvm_oracle as O
    from fvm_b200 import workloads as W
    wl = W.batched_sweep(65536, cells, first=0, count=8)
    asm = O.assemble(wl["faces"], wl["a"][0], wl["d"][0], "exponential", **wl["bc"])
    O.theta_run_scipy(asm, wl["k"][0], 0.5, wl["w0"][0], 3)             # warm-up, 3 steps
    nsteps, done, t_total = 200, 0, 0.0
    for p in range(8):
        asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", **wl["bc"])
        t0 = time.perf_counter()
        O.theta_run_scipy(asm, wl["k"][p], 0.5, wl["w0"][p], nsteps, every=nsteps)
        t_total += time.perf_counter() - t0
        done += 1
        if t_total > budget_s:
            break
    return {"value": done * nsteps * cells / t_total, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "%d problems x %d cells x %d steps of cfg3 (oracle port of the reference's numpy/scipy "
                      "spsolve loop, fvm_classes.py:408-411; assembly excluded)" % (done, cells, nsteps)}


def run_reference_arm(args):
    """--impl reference: the reference's CPU algorithm on all host cores (one process per core,
    one problem each, single-threaded BLAS), bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    if args.inner is None:
        args.inner = 10000
    cores = os.cpu_count() or 1
    nsteps = 200
    jobs = [(i, 1, args.cells, nsteps) for i in range(cores)]
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_worker, jobs)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_worker, jobs)
        dt = time.perf_counter() - t0
    value = args.steps * cores * nsteps * args.cells / dt
    sample = "per step: %d problems (one per core) x %d cells x %d theta steps of cfg3" % (cores, args.cells, nsteps)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config(args, {"sample": sample}),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc = index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        out = self.proc.communicate()[0]
        sm, mx, reasons = [], [], set()
        for ln in out.strip().splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


FUSED_SUMMARY = "r02b_fused_cfg3_full.summary.txt"     # ncu --set full of the default command's dominant kernel


def ncu_traffic(summary_name):
    """dram__bytes_read.sum + dram__bytes_write.sum (bytes per launch) of a committed `ncu --set full`
    capture, parsed from profiles/<summary_name> (written by profiles/ncu_summary.py); None if absent."""
    path = os.path.join(ROOT, "profiles", summary_name)
    if not os.path.exists(path):
        return None
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    total, seen = 0.0, 0
    for ln in open(path):
        f = ln.split()
        if len(f) >= 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and f[1] in scale:
            total += float(f[2]) * scale[f[1]]
            seen += 1
    return total if seen == 2 else None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


def _slab_rows(faces, a, d, dt, theta, J, j0, j1, f_lo):
    """CHECKER ONLY (not the product path): rows of A = I - dt*theta*M and B = I + dt*(1-theta)*M for the
    global cells [j0, j1) of a graded mesh, scalar a and d, exponential fitting, zero-flux Robin ends,
    restated in torch from SURVEY.md 8(a)'s closed form [fvm_classes.py:117-131,149-177,376-377].  Used
    for the full-size residual ||A w_new - B w_old|| of the cfg 5 leg, where no oracle can run."""
    import torch
    f = faces                                  # faces[i] = global face f_lo + i
    o = j0 - f_lo                              # local index of face j0
    n = j1 - j0
    h = f[1:] - f[:-1]                         # widths of global cells f_lo ...
    c = 0.5 * (f[1:] + f[:-1])
    mu = a * h / d
    D = d * (mu / torch.expm1(mu) + 0.5 * mu)
    dx = c[1:] - c[:-1]                        # face between local cells i, i+1
    af = h[1:] / (2 * dx) * a + h[:-1] / (2 * dx) * a   # NOT a: at 2^27+ cells dx carries ~1e-7 of rounding, as in the reference's arithmetic
    Df = h[1:] / (2 * dx) * D[:-1] + h[:-1] / (2 * dx) * D[1:]
    A_hi, A_lo, Dd = af * h[1:] / (2 * dx), af * h[:-1] / (2 * dx), Df / dx
    z = torch.zeros(1, dtype=f.dtype, device=f.device)
    # face arrays indexed by local cell: left face of cell i = face i-1, right face = face i; pad outside the mesh
    def left(x):
        return torch.cat([z, x])[o:o + n] if f_lo == 0 else x[o - 1:o - 1 + n]
    def right(x):
        full = torch.cat([x, z]) if f_lo + f.numel() - 1 == J else x
        return full[o:o + n]
    hh = h[o:o + n]
    lo = (left(A_hi) + left(Dd)) / hh
    di = (left(A_lo) - right(A_hi) - left(Dd) - right(Dd)) / hh
    up = (-right(A_lo) + right(Dd)) / hh
    sA, sB = dt * theta, dt * (1 - theta)
    return (-sA * lo, 1 - sA * di, -sA * up), (sB * lo, 1 + sB * di, sB * up)


def _timed_steps(m, dev, world, local, min_seconds=2.0):
    """Warm up, find a step count that lasts `min_seconds`, time it with CUDA events and the clock sampler;
    returns (ms per theta step = max over ranks, steps, clocks)."""
    import torch
    import torch.distributed as dist
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    m.step(3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(); m.step(10); e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = int(min(5000, max(20, np.ceil(min_seconds * 1e3 / float(t.item())))))
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0.record(); m.step(n); e1.record()
    barrier()
    clocks = sampler.stop()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), n, clocks

def long_system_legs(world, rank, local, dev, peaks, peak_src, group):
    """Variant 3 inside the default (driver-run) command, on ALL ranks:
    cfg 5 -- ONE graded mesh of world x 2^27 cells split over the ranks (SPIKE, one 4-double NCCL all-gather
    per theta step), timed for >= 2 s with clocks, next to the same kernels on 2^27 cells per rank with no
    communication (the N = 1 figure the weak efficiency is quoted against), the all-gather's own duration,
    agreement with a single-GPU solve of the whole mesh after 20 steps, and the residual of one step at
    full size; cfg 4 -- 512 problems x 2^20 cells per GPU (BASELINE's 4,096 on 8 GPUs), theta = 1."""
    import torch
    import torch.distributed as dist
    from fvm_b200 import PartitionedAdvectionDiffusionModel, workloads as W
    from fvm_b200.partitioned import plan_partition
    out = {}
    from fvm_b200 import _lib
    T, Jr = int(_lib.lib().fvm_spike_tile_cells(2)), 1 << 27       # the default tile of PartitionedAdvectionDiffusionModel

    def model_cfg5(J, j0, j1, grp):
        sl = W.graded_slab(J, j0, j1, dev)
        m = PartitionedAdvectionDiffusionModel(sl["faces"], 1.0, sl["d_pe2"], sl["dt"], "exponential", P=1, J=J,
                                               group=grp, device=dev)
        m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
        m.prepare(0.5)
        m.set_state(sl["w0"])
        return m, sl

    # ---- cfg 5, every rank alone on 2^27 cells (no collective): the N = 1 reference of the weak scaling
    m1, _ = model_cfg5(Jr, 0, Jr, None)
    ms1, n1, clocks1 = _timed_steps(m1, dev, world, local)
    extents = [int(m1.extents[0]), int(m1.extents[1])]
    del m1
    torch.cuda.empty_cache()
    leg = {"kernel": "tile_solve_kernel + reduced + separators (variant 3), cfg5: one graded mesh of %d x 2^27 cells"
                     % world, "bound": "hbm", "algorithmic_bytes_per_cell_update": HBM_B_PER_CELL_UPDATE_SPIKE,
           "peak": peaks["hbm_gbs"], "unit": "GB/s", "peak_source": peak_src, "tile_cells": T,
           "spike_extents_cells": extents,
           "single_gpu": {"cells": Jr, "ms_per_theta_step": ms1, "steps_timed": n1, "clocks": clocks1,
                          "cell_updates_per_s": Jr / (ms1 * 1e-3),
                          "achieved": HBM_B_PER_CELL_UPDATE_SPIKE * Jr / (ms1 * 1e-3) / 1e9},
           "traffic": ncu_traffic("r02b_tile_solve_cfg5.summary.txt"),
           "traffic_note": "tile_solve_kernel only, one launch at 2^27 cells: algorithmic 40 B x 2^27 = 5.37e9, plus 16 B of "
                           "spike vectors for the 34 % of the cells inside the spike extents = 0.73e9"}
    # ---- cfg 5 proper: one mesh over all ranks
    J = world * Jr
    j0, j1 = plan_partition(J, T, world)["ranges"][rank]
    m, sl = model_cfg5(J, j0, j1, group)
    msN, nN, clocksN = _timed_steps(m, dev, world, local)
    gather_us = None
    if world > 1:
        m.comm.timing = []
        m.step(200)
        torch.cuda.synchronize()
        gather_us = 1e3 * float(np.mean([a.elapsed_time(b) for a, b in m.comm.timing]))
        m.comm.timing = None
    gbs = HBM_B_PER_CELL_UPDATE_SPIKE * Jr / (msN * 1e-3) / 1e9
    leg.update({"cells": J, "n_gpus": world, "nccl_ranks": world if world > 1 else 0,
                "ms_per_theta_step": msN, "steps_timed": nN, "clocks": clocksN,
                "cell_updates_per_s": J / (msN * 1e-3), "achieved": gbs, "frac": gbs / peaks["hbm_gbs"],
                "achieved_note": "per GPU: 40 B x 2^27 cells / ms_per_theta_step (whole step incl. the collective)",
                "weak_efficiency_vs_n1": ms1 / msN,
                "collective": {"name": "ncclAllGather (torch.distributed.all_gather_into_tensor), 4 doubles per block of the "
                                       "separator system and step: %d blocks of %d separators per rank, %d B per rank "
                                       "(fvm_b200/partitioned.py: _step_rank_level)"
                                       % (m.B, m.nt // m.B, 32 * m.B),
                               "us_per_step": gather_us,
                               "share_of_step": None if gather_us is None else gather_us * 1e-3 / msN}})
    # ---- residual of ONE theta step at full size: || A w_new - B w_old ||_inf / || B w_old ||_inf
    m.set_state(sl["w0"])
    m.step(19)
    w_old = m.get_state()[0]
    m.step(1)
    w_new = m.get_state()[0]
    gl = lambda w: torch.stack([w[0], w[-1]])
    if world > 1:
        e_old = [torch.empty(2, dtype=w_old.dtype, device=dev) for _ in range(world)]
        e_new = [torch.empty(2, dtype=w_old.dtype, device=dev) for _ in range(world)]
        dist.all_gather(e_old, gl(w_old)); dist.all_gather(e_new, gl(w_new))
    zero = torch.zeros((), dtype=w_old.dtype, device=dev)
    def ghosted(w, edges):
        left = edges[rank - 1][1] if rank > 0 else zero
        right = edges[rank + 1][0] if rank + 1 < world else zero
        return torch.cat([left[None], w, right[None]])
    go = ghosted(w_old, e_old if world > 1 else None)
    gn = ghosted(w_new, e_new if world > 1 else None)
    del w_old
    (Al, Ad, Au), (Bl, Bd, Bu) = _slab_rows(m.faces, 1.0, sl["d_pe2"], sl["dt"], 0.5, J, m.j0, m.j1, m.f_lo)
    rhs = Bl * go[:-2] + Bd * go[1:-1] + Bu * go[2:]
    res = Al * gn[:-2] + Ad * gn[1:-1] + Au * gn[2:] - rhs
    r = torch.stack([res.abs().max(), rhs.abs().max()])
    del Al, Ad, Au, Bl, Bd, Bu, rhs, res, go, gn
    if world > 1:
        dist.all_reduce(r, op=dist.ReduceOp.MAX)
    leg["residual_after_20_steps"] = {"value": float(r[0] / r[1]), "what": "max |A w_21 - B w_20| / max |B w_20| over all "
                                      "%d cells, A and B restated in torch from the mesh (bench.py: _slab_rows)" % J}
    # ---- N ranks against ONE GPU on the same mesh (rank 0 solves all of it; it fits: 7 arrays of J doubles)
    if world > 1:
        full = torch.empty(J, dtype=torch.float64, device=dev)
        if rank == 0:
            ms_, sl_ = model_cfg5(J, 0, J, None)
            ms_.step(20)
            full.copy_(ms_.get_state()[0])
            del ms_, sl_
            torch.cuda.empty_cache()
        dist.broadcast(full, 0)
        d = torch.stack([(w_new - full[m.j0:m.j1]).abs().max(), full[m.j0:m.j1].abs().max()])
        # w_new is the state after 20 steps of the distributed run
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        mv = (w_new - sl["w0"][0]).abs().max()[None]
        dist.all_reduce(mv, op=dist.ReduceOp.MAX)
        leg["agreement_1_vs_n_ranks"] = {"value": float(d[0] / d[1]), "steps": 20,
                                         "state_change_over_the_20_steps": float(mv[0] / d[1]),
                                         "what": "max |w(%d ranks) - w(1 GPU)| / max |w| over all %d cells" % (world, J)}
        del full
    del m, sl, w_new
    torch.cuda.empty_cache()
    out["roofline_long_system_variant"] = leg
    # ---- cfg 4: 512 problems x 2^20 cells per GPU, theta = 1, dt = 100 h_min / a, reaction term; no communication
    Pn, J4 = 512, 1 << 20
    wl = W.stiff_reaction(P=world * Pn, J=J4, first=rank * Pn, count=Pn)
    m4 = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=Pn,
                                            reaction=wl["reaction"], device=dev)
    m4.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
    m4.prepare(1.0)
    m4.set_state(torch.as_tensor(wl["w0"], device=dev))
    ms4, n4, clocks4 = _timed_steps(m4, dev, world, local)
    finite = bool(torch.isfinite(m4.get_state()).all())
    g4 = HBM_B_PER_CELL_UPDATE_SPIKE * Pn * J4 / (ms4 * 1e-3) / 1e9
    out["cfg4_long_systems"] = {
        "workload": "cfg4 stiff reaction-diffusion: %d problems/GPU x 2^20 cells (BASELINE: 4,096 = 512 per GPU on 8 GPUs), "
                    "per-problem a/d/reaction, theta = 1, dt = 100 h_min/a, sharded, no communication" % Pn,
        "n_gpus": world, "cell_updates_per_s": world * Pn * J4 / (ms4 * 1e-3), "ms_per_theta_step": ms4,
        "steps_timed": n4, "clocks": clocks4, "bound": "hbm", "achieved": g4, "peak": peaks["hbm_gbs"], "unit": "GB/s",
        "frac": g4 / peaks["hbm_gbs"], "algorithmic_bytes_per_cell_update": HBM_B_PER_CELL_UPDATE_SPIKE,
        "note": "per GPU; every cell lies inside the spike extents here (dt = 100 h_min/a: the spikes do not decay), so "
                "the kernels really move 56 B per cell-update", "finite": finite,
        "spike_extents_cells": [int(m4.extents[0]), int(m4.extents[1])]}
    del m4, wl
    torch.cuda.empty_cache()
    return out


def run_partitioned(args):
    """--workload cfg5: ONE graded mesh of world * 2^27 cells (2^30 at 8 GPUs), contiguous slabs per rank,
    exponential fitting, zero-flux Robin, Crank-Nicolson, dt = 4 h_min/a; the reduced separator system is
    all-gathered over NCCL once per theta step.
    --workload cfg4: 256 problems per GPU x 2^20 cells, per-problem a/d/reaction, theta = 1, dt = 100 h_min/a
    (independent problems: sharded, no communication).  One bench step = --inner theta steps (default 20/10)."""
    import torch
    import torch.distributed as dist
    from fvm_b200 import PartitionedAdvectionDiffusionModel, workloads as W

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K, Wm = args.steps, max(args.warmup, 3)
    if args.workload == "cfg5":
        inner = 20 if args.inner is None else args.inner
        J = world << 27
        from fvm_b200.partitioned import plan_partition
        from fvm_b200 import _lib
        T = int(_lib.lib().fvm_spike_tile_cells(2))
        j0, j1 = plan_partition(J, T, world)["ranges"][rank]
        sl = W.graded_slab(J, j0, j1, dev)
        m = PartitionedAdvectionDiffusionModel(sl["faces"], 1.0, sl["d_pe2"], sl["dt"], "exponential", P=1, J=J,
                                               group=dist.group.WORLD if world > 1 else None, device=dev)
        theta, w0, cells_rank = 0.5, sl["w0"], j1 - j0
        desc = "cfg5 one graded mesh of %d x 2^27 = %d cells split over %d GPU(s), exponential, zero-flux Robin, " \
               "Crank-Nicolson, dt = 4 h_min/a, SPIKE reduced system all-gathered over NCCL each step" % (world, J, world)
    else:
        inner = 10 if args.inner is None else args.inner
        Pn, J = 512, 1 << 20
        wl = W.stiff_reaction(P=world * Pn, J=J, first=rank * Pn, count=Pn)
        m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=Pn,
                                               reaction=wl["reaction"], device=dev)
        theta, w0, cells_rank = 1.0, torch.as_tensor(wl["w0"], device=dev), Pn * J
        desc = "cfg4 stiff reaction-diffusion: %d problems/GPU x 2^20 cells (BASELINE: 4,096 problems = 512 per GPU on 8 GPUs), " \
               "per-problem a/d/reaction, theta = 1, dt = 100 h_min/a" % Pn
    m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
    m.prepare(theta)
    m.set_state(w0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(Wm):
        m.step(inner)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(K):
        m.step(inner)
    t1.record()
    barrier()
    clocks = sampler.stop()
    t = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    cu_step = float(cells_rank) * inner
    value = world * cu_step * K / (ms_total * 1e-3)
    finite = bool(torch.isfinite(m.get_state()).all())
    # end to end: pinned host state in, `inner` steps, pinned host state out
    n_real = m.j1 - m.j0
    h_in = torch.empty((m.P, n_real), dtype=torch.float64).pin_memory()
    h_out = torch.empty((m.P, n_real), dtype=torch.float64).pin_memory()
    h_in.copy_(w0 if w0.shape[1] == n_real else w0[:, :n_real])
    stage = torch.empty((m.P, n_real), dtype=torch.float64, device=dev)

    def e2e_step():
        stage.copy_(h_in, non_blocking=True)
        m.set_state(stage)
        m.step(inner)
        h_out.copy_(m.get_state(), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    e2e_step()
    barrier()
    t0.record()
    for _ in range(K):
        e2e_step()
    t1.record()
    barrier()
    te = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    if rank == 0:
        peaks, peak_src = measured_peaks()
        gbs = HBM_B_PER_CELL_UPDATE_SPIKE * cu_step * K / (ms_total * 1e-3) / 1e9      # per GPU
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": desc, "theta_steps_per_bench_step": inner, "cells_per_gpu": cells_rank,
                           "tile_cells": m.T, "spike_extents_cells": [int(m.extents[0]), int(m.extents[1])],
                           "l2": "state and factors (%.1f GiB/GPU) exceed the 126 MB L2; no flush needed"
                                 % (7 * cells_rank * 8 / 2 ** 30)},
                "e2e": {"value": world * cu_step * K / (float(te.item()) * 1e-3), "unit": UNIT,
                        "h2d_bytes_per_step": m.P * n_real * 8, "d2h_bytes_per_step": m.P * n_real * 8,
                        "api": "PartitionedAdvectionDiffusionModel.set_state/step/get_state with pinned host state"},
                "gpu_launches": K * inner * 5, "clocks": clocks,
                "roofline": {"kernel": "tile_solve_kernel (variant 3)", "bound": "hbm", "achieved": gbs,
                             "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
                             "peak_source": peak_src, "algorithmic_bytes_per_cell_update": HBM_B_PER_CELL_UPDATE_SPIKE,
                             "traffic": None, "note": "per GPU; whole theta step incl. reduced solve, correction and all-gather"},
                "cpu_baseline": None if args.no_cpu_baseline else cpu_baseline_single(1024), "finite": finite}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_b200(args):
    import torch
    import torch.distributed as dist
    from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from fvm_b200 import _lib
    numa_node = _lib.bind_to_gpu_numa_node(local)      # host buffers next to this rank's GPU
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    if args.inner is None:
        args.inner = 10000
    P, J, inner, K, Wm = args.problems, args.cells, args.inner, args.steps, max(args.warmup, 3)
    # weak scaling: every rank owns its own shard of `problems` problems of a (world*P)-problem sweep
    wl = W.batched_sweep(world * P, J, first=rank * P, count=P)
    model = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", device=dev)
    model.set_boundary_conditions(**wl["bc"])
    w0 = torch.as_tensor(wl["w0"], device=dev)
    w = w0.clone()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ce = args.checkpoint_every if 0 < args.checkpoint_every <= inner else 0
    n_ck = inner // ce if ce else 0
    ckpt = torch.empty((n_ck, P, J), dtype=torch.float64, device=dev) if n_ck else None

    # ---- device-resident throughput ("value"): one launch = the whole run of `inner` steps,
    # checkpoints written to HBM on the way
    for _ in range(Wm):
        model.run_fused_(w, 0.5, inner, ce, ckpt)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t_all0 = torch.cuda.Event(enable_timing=True); t_all1 = torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for i in range(K):
        ev[i][0].record()
        model.run_fused_(w, 0.5, inner, ce, ckpt)
        ev[i][1].record()
    t_all1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = t_all0.elapsed_time(t_all1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    cu_per_step_rank = float(P) * J * inner
    value = world * cu_per_step_rank * K / (ms_total * 1e-3)
    finite = bool(torch.isfinite(w).all()) and (ckpt is None or bool(torch.isfinite(ckpt[-1]).all()))
    del ckpt

    # ---- end to end through the host-facing call: pinned host state in, pinned host state out
    e2e = None
    if not args.no_e2e:
        h_in = torch.empty((P, J), dtype=torch.float64).pin_memory()
        h_out = torch.empty((P, J), dtype=torch.float64).pin_memory()
        h_ck = torch.empty((n_ck, P, J), dtype=torch.float64).pin_memory() if n_ck else None
        h_in.copy_(w0)
        e2e_kw = {} if args.e2e_chunks is None else {"chunks": args.e2e_chunks}
        for _ in range(2):
            model.run_host(h_in, 0.5, inner, out=h_out, checkpoint_every=ce, ckpt_out=h_ck, **e2e_kw)
        barrier()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(K):
            model.run_host(h_in, 0.5, inner, out=h_out, checkpoint_every=ce, ckpt_out=h_ck, **e2e_kw)
        t1.record()
        barrier()
        te = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": world * cu_per_step_rank * K / (float(te.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": P * J * 8, "d2h_bytes_per_step": (1 + n_ck) * P * J * 8,
               "api": "BatchedAdvectionDiffusionModel.run_host (pinned host state in, final state and every "
                      "checkpoint out to pinned host memory, every step)", "host_numa_node": numa_node}
        del h_ck

    # ---- streaming variant (HBM roofline evidence), rank 0 only, small step count
    hbm = None
    peaks, peak_src = measured_peaks()
    if rank == 0:
        try:
            w_t = w0.t().contiguous()
            model.factor(0.5)
            for _ in range(3):
                model.step_thomas_(w_t, 0.5, 1)
            a0 = torch.cuda.Event(enable_timing=True); a1 = torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            n_th = 20
            a0.record()
            for _ in range(n_th):
                model.step_thomas_(w_t, 0.5, 1)
            a1.record()
            torch.cuda.synchronize()
            ms = a0.elapsed_time(a1) / n_th
            gbs = HBM_B_PER_CELL_UPDATE_THOMAS * P * J / (ms * 1e-3) / 1e9
            hbm = {"kernel": "thomas_steps_kernel (variant 1, one theta step per launch)", "bound": "hbm",
                   "cell_updates_per_s": P * J / (ms * 1e-3), "achieved": gbs, "peak": peaks["hbm_gbs"],
                   "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"], "peak_source": peak_src,
                   "algorithmic_bytes_per_cell_update": HBM_B_PER_CELL_UPDATE_THOMAS,
                   "traffic_model": {"bytes_per_cell_update": HBM_B_MOVED_THOMAS,
                                     "achieved": HBM_B_MOVED_THOMAS * P * J / (ms * 1e-3) / 1e9,
                                     "note": "the two-pass kernel parks the eliminated value in HBM and re-reads s: "
                                             "64 B moved for 40 B of algorithmic traffic"},
                   "traffic": ncu_traffic("r02_thomas_v1.summary.txt")}
            del w_t
        except Exception as exc:  # noqa: BLE001
            hbm = {"error": repr(exc)}

    # ---- variant 3 (cfg 5 over all ranks with its NCCL all-gather, cfg 4 sharded), every rank takes part
    del w, w0, model
    torch.cuda.empty_cache()
    long_legs = {"roofline_long_system_variant": None, "cfg4_long_systems": None}
    if not args.no_long_legs:
        long_legs = long_system_legs(world, rank, local, dev, peaks, peak_src, dist.group.WORLD if world > 1 else None)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (fused_tmem_kernel): it touches HBM only at launch
    # boundaries, so the bound is on chip.  Coefficients come from tensor memory and registers
    # (not shared memory), which leaves the fp64 pipe as the binding unit: 64 DFMA/clk/SM,
    # measured on this GPU type by profiles/microbench/fp64_peak.cu (r01_fp64_peak_b200.txt).
    f_clk = (clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)) * 1e6
    f_max = peaks.get("sm_max_mhz", 1965.0) * 1e6
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    cu_s_kernel = cu_per_step_rank / (kernel_ms * 1e-3)
    fp64_peak = sms * 64.0 * 2.0 * f_max / 1e12                # TFLOP/s at max clock
    fp64_ach = FLOP_PER_CELL_UPDATE * cu_s_kernel / 1e12
    issue_frac = FP64_INSTR_PER_CELL_UPDATE * cu_s_kernel / (sms * 64.0 * f_clk)   # thread-level DFMAs issued / (64 lanes per clk per SM)
    smem_peak = sms * 128.0 * f_max / 1e9
    onchip = ONCHIP_B_PER_CELL_UPDATE * cu_s_kernel / 1e9
    roofline = {"kernel": "fused_tmem_pipe_kernel<1, 0> (variant 2, Crank-Nicolson: state in registers, coefficients in TMEM)",
                "bound": "fp64", "achieved": fp64_ach, "peak": fp64_peak, "unit": "TFLOP/s",
                "frac": fp64_ach / fp64_peak, "traffic": ncu_traffic(FUSED_SUMMARY),
                "traffic_note": "DRAM bytes per launch from profiles/%s (ncu --set full, same command); algorithmic "
                                "HBM bytes per launch = (16 + 8 x %d checkpoints) B x %d x %d = %.3e"
                                % (FUSED_SUMMARY, n_ck, P, J, (16.0 + 8.0 * n_ck) * P * J),
                "peak_source": "%d SMs x 64 DFMA/clk x 2 flop x %.0f MHz (max SM clock, %s); the 64 DFMA/clk/SM rate is "
                               "measured (profiles/microbench/r01_fp64_peak_b200.txt)" % (sms, f_max / 1e6, peak_src),
                "register_operand_ceiling": {
                    "dfma_per_clk_per_sm": DFMA_DISTINCT_OPERANDS_PER_CLK_SM,
                    "issued_dfma_per_clk_per_sm": issue_frac * 64.0,
                    "frac": issue_frac * 64.0 / DFMA_DISTINCT_OPERANDS_PER_CLK_SM,
                    "note": "a DFMA whose three operands are distinct registers (every FMA of the sweeps) sustains "
                            "42-45 per clk per SM on this GPU, not 64: register-file operand bandwidth "
                            "(profiles/microbench/fp64_operands.cu, r01_fp64_operands_b200.txt); roofline.frac above is "
                            "still quoted against the 64/clk/SM pipe peak and SURVEY's 11 flop per cell-update, "
                            "although the loop issues 5.2 DFMA (10.4 flop) per cell-update"},
                "flop_per_cell_update": FLOP_PER_CELL_UPDATE,
                "fp64_pipe_issue_frac": issue_frac,
                "fp64_instr_per_cell_update": FP64_INSTR_PER_CELL_UPDATE,
                "onchip_coefficient_traffic": {"achieved": onchip, "unit": "GB/s", "bytes_per_cell_update": ONCHIP_B_PER_CELL_UPDATE,
                                               "shared_memory_peak": smem_peak,
                                               "note": "coefficient reads served by TMEM (tcgen05.ld) and registers; exceeds what "
                                                       "the 128 B/clk/SM shared-memory pipe could deliver when frac > 1",
                                               "frac_of_shared_memory_peak": onchip / smem_peak},
                "hbm": {"achieved": (16.0 + 8.0 * n_ck) * P * J / (kernel_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                        "unit": "GB/s", "note": "state read once, written once, plus %d checkpoints, per launch of "
                                                "%d steps" % (n_ck, inner)},
                "kernel_ms_per_launch": kernel_ms, "cell_updates_per_launch": cu_per_step_rank}

    cpu = None if args.no_cpu_baseline else cpu_baseline_single(J)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config(args), "e2e": e2e, "gpu_launches": K,
            "clocks": clocks, "roofline": roofline, "roofline_streaming_variant": hbm,
            "roofline_long_system_variant": long_legs["roofline_long_system_variant"],
            "cfg4_long_systems": long_legs["cfg4_long_systems"], "cpu_baseline": cpu, "finite": finite}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference_arm(a)
    elif a.workload == "cfg3":
        run_b200(a)
    else:
        run_partitioned(a)
