"""cfg 3 sizes, fused on-chip variant, for the three code paths of theta: Crank-Nicolson (4 fp64
operations per cell-update in the sweeps), implicit Euler (3) and a general theta (5).
    python profiles/microbench/theta_modes.py > gpurun_out/theta_modes.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

P, J, n = 65536, 1024, 2000
wl = W.batched_sweep(P, J)
m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential")
m.set_boundary_conditions(**wl["bc"])
w0 = torch.as_tensor(wl["w0"], device="cuda")
for theta in (0.5, 1.0, 0.7):
    w = w0.clone()
    m.run_fused_(w, theta, n)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        m.run_fused_(w, theta, n)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print("theta %.1f: %.2f ms per %d steps, %.3e cell-updates/s, finite=%s"
          % (theta, ms, n, P * J * n / (ms * 1e-3), bool(torch.isfinite(w).all())))
