"""CTA-per-system kernel with more systems than SMs: J = 2048 (128 threads x 16 cells, four CTAs per SM) and
J = 4096 (256 x 16, two per SM) against the one-CTA-per-SM configurations that fewer systems get.
    python profiles/microbench/cta_many_systems.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

sms = torch.cuda.get_device_properties(0).multi_processor_count
for J, per_sm in ((2048, 4), (4096, 2)):
    for P in (sms, per_sm * sms):
        wl = W.batched_sweep(P, J, seed=1)
        m = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential")
        m.set_boundary_conditions(**wl["bc"])
        w0 = torch.as_tensor(wl["w0"], device="cuda")
        m.run(w0, 0.5, 20, variant="fused")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); m.run(w0, 0.5, 2000, variant="fused"); e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print("J = %d, %4d systems (%d per SM): %.2f ms per 2000 steps, %.3e cell-updates/s"
              % (J, P, P // sms, ms, P * J * 2000 / ms * 1e3), flush=True)
