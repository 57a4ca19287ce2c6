"""Variant 1 on tiny meshes (north_star: "interleaved thread-per-system Thomas for tiny meshes"): the
register-resident kernel, J = 64 (and 32) cells, 2^20 problems, many steps per launch.
    python profiles/microbench/thomas_tiny.py [J] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

J = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 50
P = 1 << 20
wl = W.batched_sweep(P, J)
m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential")
m.set_boundary_conditions(**wl["bc"])
w_t = torch.as_tensor(wl["w0"], device="cuda").t().contiguous()
m.step_thomas_(w_t, 0.5, n)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); m.step_thomas_(w_t, 0.5, n); e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
cu = P * J * n / ms * 1e3
print("J = %d, P = 2^20, %d steps in one launch: %.2f ms, %.3e cell-updates/s = %.0f GB/s at 40 B (%.2f of 6548 GB/s), %.0f GB/s at the 24 B the kernel reads"
      % (J, n, ms, cu, cu * 40 / 1e9, cu * 40 / 1e9 / 6548.2, cu * 24 / 1e9))
print("finite", bool(torch.isfinite(w_t).all()))
