"""cfg 2 through the on-chip CTA-per-system kernel: 29 Peclet numbers x 3 schemes on the 4,096-cell graded
mesh, 2,000 Crank-Nicolson steps (the command the r02 ncu capture of fused_cta_kernel profiles).
    python profiles/microbench/cfg2_run.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

wl = W.peclet_sweep(4096)
for scheme in ("central", "upwind", "exponential"):
    m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], scheme)
    m.set_boundary_conditions(**wl["bc"])
    w0 = torch.as_tensor(wl["w0"], device="cuda")
    m.run(w0, 0.5, 20, variant="fused")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); w, _ = m.run(w0, 0.5, 2000, variant="fused"); e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("%-11s 29 problems x 4096 cells x 2000 steps: %.2f ms, %.3e cell-updates/s, finite %s"
          % (scheme, ms, 29 * 4096 * 2000 / ms * 1e3, bool(torch.isfinite(w).all())), flush=True)

# the same kernel with every SM busy: two systems per SM
wl = W.peclet_sweep(4096)
rep = -(-296 // 29)
tile = lambda x: np.concatenate([x] * rep)[:296]
m = BatchedAdvectionDiffusionModel(wl["faces"], tile(wl["a"]), tile(wl["d"]), tile(wl["k"]), "exponential")
m.set_boundary_conditions(**wl["bc"])
w0 = torch.as_tensor(np.concatenate([wl["w0"]] * rep)[:296], device="cuda")
m.run(w0, 0.5, 20, variant="fused")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); w, _ = m.run(w0, 0.5, 2000, variant="fused"); e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("exponential 296 problems x 4096 cells x 2000 steps (2 per SM): %.2f ms, %.3e cell-updates/s" % (ms, 296 * 4096 * 2000 / ms * 1e3))
