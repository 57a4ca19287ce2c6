"""Scale test of variant 3 on one GPU: one mesh of 2^k cells (closed-form graded mesh generated
on the device), cfg 5 parameters.  Prints ms/step and algorithmic GB/s."""
import sys, time, json
import numpy as np, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__)))))
from fvm_b200 import PartitionedAdvectionDiffusionModel
from fvm_b200.workloads import graded_slab

k = int(sys.argv[1]) if len(sys.argv) > 1 else 26
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1
J = 1 << k
dev = torch.device('cuda')
sl = graded_slab(J, 0, J, dev)
m = PartitionedAdvectionDiffusionModel(sl['faces'], 1.0, sl['d_pe2'], sl['dt'], 'exponential', P=P, J=J, config=0)
m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
t0 = time.time(); m.prepare(0.5); torch.cuda.synchronize(); print('setup s', time.time() - t0, 'extents', list(m.extents), 'tiles', m.nt)
w0 = sl['w0'].expand(P, -1).contiguous()
m.set_state(w0)
m.step(3); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 10
e0.record(); m.step(n); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
w = m.get_state()
h = sl['faces'][1:] - sl['faces'][:-1]
print(json.dumps(dict(J=J, P=P, ms_per_step=ms, cell_updates_per_s=P * J / ms * 1e3, gbs_40B=40 * P * J / ms / 1e6,
                      mass0=float((w0[0] * h).sum()), mass1=float((w[0] * h).sum()), finite=bool(torch.isfinite(w).all()))))
