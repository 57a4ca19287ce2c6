"""cfg 5 on one GPU (2^27 cells, variant 3): a few theta steps, for the per-kernel launch list.
    ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file X.csv python profiles/microbench/cfg5_steps.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from fvm_b200 import PartitionedAdvectionDiffusionModel, workloads as W

J = 1 << 27
sl = W.graded_slab(J, 0, J, "cuda")
m = PartitionedAdvectionDiffusionModel(sl["faces"], 1.0, sl["d_pe2"], sl["dt"], "exponential", P=1, J=J, config=int(os.environ.get("TILE_CFG", 2)))
m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
m.prepare(0.5)
m.set_state(sl["w0"])
m.step(3)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); m.step(10); e1.record()
torch.cuda.synchronize()
print("2^27 cells: %.4f ms per theta step" % (e0.elapsed_time(e1) / 10))
