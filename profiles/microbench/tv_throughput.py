"""cfg-3-shaped throughput of the on-chip kernel with a boundary-flux history and with a time-step
history, against the constant-coefficient launch.
    python profiles/microbench/tv_throughput.py > gpurun_out/tv_throughput.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
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

base = timed(lambda: m.run(w0, 0.5, n, variant="fused"))
print("constant beta, constant dt      : %.2f ms per %d steps, %.3e cell-updates/s" % (base, n, P * J * n / base * 1e3))
lf = 0.1 * np.sin(0.01 * np.arange(n))
m.set_boundary_conditions(left_flux=0.1, right_flux=0.0)
t = timed(lambda: m.run_time_varying(w0, 0.5, left_flux_t=lf, right_flux_t=-lf))
print("flux history (shared, both ends): %.2f ms, %.3e cu/s, %.3f of the constant launch" % (t, P * J * n / t * 1e3, base / t))
lfp = torch.as_tensor(lf, device="cuda")[:, None] * torch.linspace(0.5, 1.5, P, device="cuda", dtype=torch.float64)[None, :]
t = timed(lambda: m.run_time_varying(w0, 0.5, left_flux_t=lfp))
print("flux history (per problem, left): %.2f ms, %.3e cu/s, %.3f of the constant launch" % (t, P * J * n / t * 1e3, base / t))
m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
for runs in (1, 4, 20):
    dt = torch.as_tensor(wl["k"], device="cuda")[None, :].repeat(n, 1)
    for r in range(1, runs):
        dt[r * n // runs:] *= 1.01
    t = timed(lambda: m.run_variable_dt(w0, 0.5, dt))
    print("time-step history, %2d runs      : %.2f ms, %.3e cu/s, %.3f of the constant launch" % (runs, t, P * J * n / t * 1e3, base / t))
