"""Small run of every kernel family, for compute-sanitizer memcheck:
   compute-sanitizer --tool memcheck python profiles/microbench/sanitize_small.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from fvm_b200 import BatchedAdvectionDiffusionModel, PartitionedAdvectionDiffusionModel, workloads as W

def batch(J, P, variant, theta, bc):
    f = W.graded_faces(J, 10.0); h = np.diff(f); rng = np.random.default_rng(J)
    a = rng.uniform(0.5, 2, P); d = a * h.max() / rng.uniform(0.3, 1.5, P); k = rng.uniform(1, 4, P) * h.min()
    m = BatchedAdvectionDiffusionModel(f, a, d, k, "exponential"); m.set_boundary_conditions(**bc)
    w, ck = m.run(rng.random((P, J)) + 0.1, theta, 12, checkpoint_every=4, variant=variant)
    m.diagnostics(w)
    return float(w.sum())

bcs = [dict(left_flux=0.1, right_flux=-0.2), dict(left_value=1.0, right_value=0.0)]
out = []
for J in (3, 33, 100, 257, 500):                       # warp-per-system, shared-memory coefficients
    out.append(batch(J, 5, "fused", 0.5, bcs[J % 2]))
for J in (600, 1024):                                  # warp-per-system, coefficients in tensor memory
    out.append(batch(J, 9, "fused", (0.5, 1.0)[J % 2], bcs[J % 3 % 2]))
for J in (1500, 3000, 5000):                           # CTA-per-system
    out.append(batch(J, 2, "fused", 0.7, bcs[J % 2]))
for J in (7, 100, 2000):                               # streaming Thomas
    out.append(batch(J, 37, "thomas", 1.0, bcs[0]))
out.append(batch(64, 3, "auto", 0.0, bcs[0]))          # explicit step
out.append(batch(3000, torch.cuda.get_device_properties(0).multi_processor_count + 3, "fused", 0.5, bcs[0]))   # CTA kernel, 256 x 16
for J, cfg, blocks in ((200, 1, None), (257, 1, None), (1000, 1, 2), (3000, 1, 3), (9000, 0, None), (9000, 2, None)):   # partitioned, lazy and eager
    f = W.graded_faces(J, 10.0); h = np.diff(f)
    for lazy in (True, False):
        m = PartitionedAdvectionDiffusionModel(f, np.array([1.0, -0.5]), np.array([0.05, 0.02]), 3 * h.min(), "upwind",
                                               P=2, config=cfg, lazy=lazy, blocks=blocks)
        m.set_boundary_conditions(**bcs[0]); m.check_mesh()
        out.append(float(m.run(np.random.default_rng(1).random((2, J)), 0.5, 5).sum()))
torch.cuda.synchronize()
print("sanitize_small ok", len(out), all(np.isfinite(out)))
