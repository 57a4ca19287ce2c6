"""One long system of 10^5 .. 10^7 cells through variant 3: time per theta step as a CUDA graph of two steps
(fvm_spike_graph_*, what step() uses at these sizes), with the whole loop in one host call (fvm_spike_steps) and
with the entry points driven one by one from Python.
    python profiles/microbench/mid_mesh_steps.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from fvm_b200 import PartitionedAdvectionDiffusionModel, workloads as W

for J in (100_000, 1_000_000, 10_000_000):
    sl = W.graded_slab(J, 0, J, "cuda")
    out = []
    for mode in ("graph", "one call", "host loop"):
        host_loop = mode == "host loop"
        m = PartitionedAdvectionDiffusionModel(sl["faces"], 1.0, sl["d_pe2"], sl["dt"], "exponential", P=1, J=J)
        m.use_graph = mode == "graph"
        m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
        m.prepare(0.5)
        m.set_state(sl["w0"])
        if host_loop:
            m.comm.timing = []
        m.step(50)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); m.step(1000); e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1))
    print("J = %9d: %.2f us per step as a CUDA graph, %.2f us in one host call, %.2f us step by step from Python"
          % (J, out[0], out[1], out[2]))
