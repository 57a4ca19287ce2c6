"""Parity over cfg 3's FULL run length (10,000 Crank-Nicolson steps) on the stiffest problems of the
sweep (smallest Peclet number: largest dt*d/h^2, where round-off survives longest), against the
reference's scipy loop (oracle) and against the longdouble Thomas arbiter.
    python profiles/microbench/long_run_parity.py [fused|thomas] > gpurun_out/long_run_parity.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from oracle import fvm_oracle as O
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

J, ns, pool, pick = 1024, 10000, 4096, 6
wl = W.batched_sweep(65536, J, first=0, count=pool)
pe = wl["a"] / wl["d"]
idx = np.concatenate([np.argsort(pe)[:pick - 2], np.argsort(pe)[-2:]])     # the stiffest, and two advective ones
m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"][idx], wl["d"][idx], wl["k"][idx], "exponential")
m.set_boundary_conditions(**wl["bc"])
variant = sys.argv[1] if len(sys.argv) > 1 else "fused"
print("variant", variant)
w, _ = m.run(wl["w0"][idx], 0.5, ns, variant=variant)
w = w.cpu().numpy()
As, Bs, betas = [], [], []
for n, p in enumerate(idx):
    asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", **wl["bc"])
    A, B = O.theta_operators(asm, wl["k"][p], 0.5)
    As.append(A); Bs.append(B); betas.append(asm["beta"])
    ref = O.theta_run_scipy(asm, wl["k"][p], 0.5, wl["w0"][p], ns, every=ns)[0]
    sc = max(np.max(np.abs(ref)), 1e-6 * np.max(np.abs(wl["w0"][p])))
    print("problem %5d  a/d %.3e  gpu-vs-scipy %.2e" % (p, pe[p], np.max(np.abs(w[n] - ref)) / sc), flush=True)
stack = lambda L, k: np.stack([x[k] for x in L])
A = {k: stack(As, k) for k in ("lo", "di", "up")}
B = {k: stack(Bs, k) for k in ("lo", "di", "up")}
arb = O.theta_run_thomas(A, B, np.stack(betas), wl["w0"][idx], ns, every=ns)[0].astype(float)
for n, p in enumerate(idx):
    sc = max(np.max(np.abs(arb[n])), 1e-6 * np.max(np.abs(wl["w0"][p])))
    print("problem %5d  gpu-vs-longdouble %.2e" % (p, np.max(np.abs(w[n] - arb[n])) / sc), flush=True)
