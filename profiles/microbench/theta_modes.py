"""cfg-3-shaped throughput of the on-chip kernel for theta = 1/2 (6 fp64 operations per cell-update),
theta = 1 (5: no "- c g" term) and a general theta (MODE 0).
    python profiles/microbench/theta_modes.py > gpurun_out/theta_modes.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

P, J, n = 65536, 1024, 2000
wl = W.batched_sweep(P, J)
m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential")
m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
w0 = torch.as_tensor(wl["w0"], device="cuda")


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

for theta in (0.5, 1.0, 0.7):
    t = timed(lambda: m.run(w0, theta, n, variant="fused"))
    print("theta = %.1f: %.2f ms per %d steps, %.3e cell-updates/s" % (theta, t, n, P * J * n / t * 1e3))
