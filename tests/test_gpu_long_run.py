"""Parity over cfg 3's FULL run length (north_star: "max relative error <= 1e-10 in fp64 over the
full run").  The loop to match is fvm_classes.py:408-411, 10,000 Crank-Nicolson steps.

Fixture: tests/golden/cfg3_long_run.npz (tests/golden/make_golden.py cfg3_long_run) -- the
UNMODIFIED reference's scipy loop on the 16 stiffest problems of the whole 65,536-problem sweep
(8 smallest a/d, 8 largest dt*d), snapshots after 1,000 / 5,000 / 10,000 steps, the reference's
own modified diffusivity D of every cell, and the exact solutions (oracle/thomas_arbiter.c,
__float128) of the reference's fp64 operators and of the operators with the fitting evaluated
in the cancellation-safe Bernoulli form.

What the fixture shows, before any GPU is involved (tests/test_oracle_golden.py pins it):
  * the reference's SuperLU loop ends 2e-12 ... 9e-11 from the exact solution of ITS OWN operators;
  * those operators carry the cancellation noise of kappa = (e^mu+1)/(e^mu-1) - 2/mu
    (fvm_classes.py:117) -- ~1e-12 relative in D per cell at these cell Peclet numbers, different
    on every libm (numpy's exp is not correctly rounded) -- and 10,000 stiff steps amplify it to
    as much as 7e-10 between `exact` and `exact_fitted`.
So "within 1e-10 of the reference after 10,000 steps" is only defined for a solver that is handed
the reference's D.  The tests therefore are:
  1. reference D passed in as the per-cell diffusivity with the central scheme (which is literally
     what the reference does after fvm_classes.py:131: every later line reads self.d):
     <= 1e-10 vs the reference's loop, <= 1e-11 vs exact.  Covers assembly of the rows,
     theta operators and the 10,000 solves; kappa itself is pinned in tests/test_gpu_assembly.py.
  2. the CUDA path's own exponential fitting (Bernoulli form, within 4 ulp of its longdouble
     evaluation): <= 1e-11 vs the exact solution of the operators IT assembled (computed here with
     the C arbiter from fvm_assemble's diagonals), and <= 1e-9 vs the reference (the kappa noise,
     not the solver: even two correct roundings of the same D, a few ulp apart, end 1.5e-10 apart
     after 10,000 of these steps -- `exact_fitted` in the fixture is such a neighbour).
"""
import os

import numpy as np
import pytest

from fvm_b200 import workloads as W

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cfg3_long_run.npz")
BC = dict(left_flux=0.0, right_flux=0.0)
EVERY = 1000


def _err(got, want, w0):
    scale = np.maximum(np.max(np.abs(want), axis=-1), 1e-6 * np.max(np.abs(w0), axis=-1))
    return np.max(np.abs(got - want), axis=-1) / scale


def _run(g, variant, scheme, d):
    from fvm_b200 import BatchedAdvectionDiffusionModel
    m = BatchedAdvectionDiffusionModel(W.graded_faces(1024), g["a"], d, g["k"], scheme)
    m.set_boundary_conditions(**BC)
    w, ck = m.run(g["w0"], 0.5, int(g["snaps"][-1]), checkpoint_every=EVERY, variant=variant)
    ck = ck.cpu().numpy()
    assert np.array_equal(ck[-1], w.cpu().numpy())
    return [ck[int(s) // EVERY - 1] for s in g["snaps"]]


@pytest.mark.parametrize("variant", ["fused", "thomas"])
def test_full_run_length_with_the_reference_diffusivity(variant):
    g = np.load(GOLD)
    assert int(g["snaps"][-1]) == 10000
    worst = [0.0, 0.0]
    for i, got in enumerate(_run(g, variant, "central", g["D_reference"])):
        e_ref = _err(got, g["reference"][:, i], g["w0"])
        e_exact = _err(got, g["exact"][:, i], g["w0"])
        worst = [max(worst[0], e_ref.max()), max(worst[1], e_exact.max())]
        assert e_ref.max() <= 1e-10, (variant, g["snaps"][i], g["problems"][e_ref.argmax()], e_ref.max())
        assert e_exact.max() <= 1e-11, (variant, g["snaps"][i], g["problems"][e_exact.argmax()], e_exact.max())
    print("%s, reference D: worst vs reference %.2e, vs exact %.2e" % (variant, worst[0], worst[1]))


_OWN = {}


def _exact_of_own_operators(g, picks):
    """Exact solution (oracle/thomas_arbiter.c, __float128) of the operators the CUDA path itself
    assembles with its Bernoulli-form fitting: fvm_assemble's diagonals of M, then A and B rounded
    as the reference rounds them (fvm_classes.py:376-377).  ~7 s of host time per problem."""
    if "exact" not in _OWN:
        from fvm_b200 import BatchedAdvectionDiffusionModel
        from oracle import fvm_oracle as O
        m = BatchedAdvectionDiffusionModel(W.graded_faces(1024), g["a"][picks], g["d"][picks], g["k"][picks], "exponential")
        m.set_boundary_conditions(**BC)
        asm = m.assemble()
        out = []
        for q, i in enumerate(picks):
            d = dict(lo=asm["lo"][q].cpu().numpy(), di=asm["di"][q].cpu().numpy(), up=asm["up"][q].cpu().numpy(),
                     alpha=np.ones(1024))
            A, B = O.theta_operators(d, g["k"][i], 0.5)
            out.append(O.theta_run_arbiter(A, B, np.zeros(1024), g["w0"][i], int(g["snaps"][-1]), g["snaps"]))
        _OWN["exact"] = np.array(out)
        _OWN["D"] = asm["D"].cpu().numpy()
    return _OWN["exact"], _OWN["D"]


@pytest.mark.parametrize("variant", ["fused", "thomas"])
def test_full_run_length_with_the_fitted_diffusivity(variant):
    from oracle import fvm_oracle as O
    g = np.load(GOLD)
    picks = [2, 7, 9, 15]                         # incl. the two problems the reference's kappa noise moves most
    exact_own, D_own = _exact_of_own_operators(g, picks)
    # the fitting itself: within 4 ulp of the longdouble evaluation, ~1e-12 from the reference's noisy D
    h = np.diff(W.graded_faces(1024))
    for q, i in enumerate(picks):
        D_ld = O.fitted_D(np.full(1024, g["a"][i]), np.full(1024, g["d"][i]), h)
        assert np.max(np.abs(D_own[q] - D_ld) / D_ld) <= 4 * np.finfo(float).eps
        assert np.max(np.abs(D_own[q] - g["D_reference"][i]) / D_ld) <= 5e-12
    got = _run(g, variant, "exponential", g["d"])
    worst = [0.0, 0.0]
    for s_i in range(len(g["snaps"])):
        e_own = _err(got[s_i][picks], exact_own[:, s_i], g["w0"][picks])
        e_ref = _err(got[s_i], g["reference"][:, s_i], g["w0"])
        worst = [max(worst[0], e_own.max()), max(worst[1], e_ref.max())]
        assert e_own.max() <= 1e-11, (variant, g["snaps"][s_i], e_own)
        assert e_ref.max() <= 1e-9, (variant, g["snaps"][s_i], g["problems"][e_ref.argmax()], e_ref.max())
    print("%s, own fitting: worst vs exact of own operators %.2e, vs reference %.2e" % (variant, worst[0], worst[1]))


@pytest.mark.parametrize("J,nprob", [(256, 3), (500, 3), (2048, 1)])
def test_full_run_length_on_the_other_on_chip_kernels(J, nprob):
    """The other on-chip kernels -- `fused_kernel` (J <= 512: one warp per system, coefficients in
    shared memory) and `fused_cta_kernel` (J > 1024: one CTA per system) -- over 10,000 stiff
    Crank-Nicolson steps, against the exact solution of the operators the oracle assembles (central
    scheme: assembly is bit-identical to the reference's, so these ARE the reference's operators).
    Same bar as the cfg 3 kernels: 1e-11."""
    from fvm_b200 import BatchedAdvectionDiffusionModel
    from oracle import fvm_oracle as O
    wl = W.batched_sweep(64, J, seed=3)
    pick = np.argsort(wl["a"] / wl["d"])[:nprob]                # the most diffusive (stiffest) problems
    a, d, k, w0 = wl["a"][pick], wl["d"][pick], wl["k"][pick], wl["w0"][pick]
    m = BatchedAdvectionDiffusionModel(wl["faces"], a, d, k, "central")
    m.set_boundary_conditions(**BC)
    got = {v: m.run(w0, 0.5, 10000, variant=v)[0].cpu().numpy() for v in ("fused", "thomas")}
    for q in range(len(pick)):
        asm = O.assemble(wl["faces"], a[q], d[q], "central", **BC)
        A, B = O.theta_operators(asm, k[q], 0.5)
        exact = O.theta_run_arbiter(A, B, asm["beta"], w0[q], 10000)[0]
        for v in got:
            err = _err(got[v][q], exact, w0[q])
            assert err <= 1e-11, (J, v, q, err)


@pytest.mark.parametrize("config", [1, 2])
def test_full_run_length_on_the_partitioned_variant(config):
    """Variant 3 (tile SPIKE, the kernels cfg 4 and cfg 5 run on) over the same 10,000 Crank-Nicolson steps
    with the reference's D: 256-cell tiles cut each 1,024-cell system into four (every step goes through
    the reduced system and the lazy spike correction), the 4,096-cell production tile holds it whole with
    3,072 padded cells behind it.  Same bars as the other variants."""
    from fvm_b200 import PartitionedAdvectionDiffusionModel
    g = np.load(GOLD)
    P = len(g["problems"])
    m = PartitionedAdvectionDiffusionModel(W.graded_faces(1024), g["a"], g["D_reference"], g["k"], "central", P=P,
                                           config=config)
    m.set_boundary_conditions(**BC)
    m.prepare(0.5)
    m.set_state(g["w0"])
    done, worst = 0, [0.0, 0.0]
    for i, snap in enumerate(g["snaps"]):
        m.step(int(snap) - done)
        done = int(snap)
        got = m.get_state().cpu().numpy()
        e_ref = _err(got, g["reference"][:, i], g["w0"])
        e_exact = _err(got, g["exact"][:, i], g["w0"])
        worst = [max(worst[0], e_ref.max()), max(worst[1], e_exact.max())]
        assert e_ref.max() <= 1e-10, (config, snap, g["problems"][e_ref.argmax()], e_ref.max())
        assert e_exact.max() <= 1e-11, (config, snap, g["problems"][e_exact.argmax()], e_exact.max())
    print("partitioned (config %d), reference D: worst vs reference %.2e, vs exact %.2e" % (config, worst[0], worst[1]))
