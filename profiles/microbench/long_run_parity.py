"""Parity over cfg 3's FULL run length (10,000 Crank-Nicolson steps) on the 16 stiffest problems of
the whole 65,536-problem sweep (tests/golden/cfg3_long_run.npz): the CUDA path with the reference's
own modified diffusivity against the reference's scipy loop and the exact solution of its
operators, and with its own Bernoulli-form fitting against the exact solution of those operators.
    python profiles/microbench/long_run_parity.py [fused|thomas] > gpurun_out/long_run_parity.txt"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

g = np.load(os.path.join(ROOT, "tests", "golden", "cfg3_long_run.npz"))
variant = sys.argv[1] if len(sys.argv) > 1 else "fused"
sc0 = 1e-6 * np.max(np.abs(g["w0"]), axis=-1)
print("variant", variant)
print("problems", " ".join(str(int(p)) for p in g["problems"]))
print("a/d", " ".join("%.3g" % x for x in g["a"] / g["d"]), " k*d", " ".join("%.3g" % x for x in g["k"] * g["d"]))


def err(got, want):
    return np.max(np.abs(got - want), axis=-1) / np.maximum(np.max(np.abs(want), axis=-1), sc0)


for label, scheme, d, targets in (("reference D (central scheme on D_reference)", "central", g["D_reference"], ("reference", "exact")),
                                  ("own fitting (exponential scheme)", "exponential", g["d"], ("exact_fitted", "reference"))):
    m = BatchedAdvectionDiffusionModel(W.graded_faces(1024), g["a"], d, g["k"], scheme)
    m.set_boundary_conditions(left_flux=0.0, right_flux=0.0)
    w, ck = m.run(g["w0"], 0.5, 10000, checkpoint_every=1000, variant=variant)
    ck = ck.cpu().numpy()
    print(label)
    for i, s in enumerate(g["snaps"]):
        for name in targets:
            e = err(ck[int(s) // 1000 - 1], g[name][:, i])
            print("  step %5d vs %-12s max %.2e   per problem: %s" % (s, name, e.max(), " ".join("%.1e" % x for x in e)))
for a, b in (("reference", "exact"), ("exact_fitted", "exact")):
    e = np.max(np.abs(g[a] - g[b]), axis=-1) / np.max(np.abs(g[b]), axis=-1)
    print("%s vs %s at steps %s: max %s" % (a, b, [int(s) for s in g["snaps"]], " ".join("%.1e" % x for x in e.max(axis=0))))
