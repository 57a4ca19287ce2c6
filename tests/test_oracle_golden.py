"""The oracle restatement against outputs of the reference itself.

tests/golden/*.npz were produced by tests/golden/make_golden.py from the UNMODIFIED
reference (loaded through oracle/ref_loader.py).  These tests pin oracle/fvm_oracle.py
to them; where /root/reference is present (build container) the oracle is also checked
against the live reference on fresh random inputs.
"""
import numpy as np
import pytest

from conftest import bc_from_array, load_golden, step_error
from oracle import fvm_oracle as O
from oracle import ref_loader as R
from fvm_b200 import workloads as W


def test_assembly_bit_identical_to_golden(golden_assembly):
    g = golden_assembly
    n = int(g["n_cases"])
    assert n == 120
    for i in range(n):
        pre = "c%03d_" % i
        a, d = g[pre + "a"], g[pre + "d"]
        a = float(a) if a.ndim == 0 else a
        d = float(d) if d.ndim == 0 else d
        asm = O.assemble(g[pre + "faces"], a, d, O.SCHEMES[int(g[pre + "scheme"])],
                         **bc_from_array(g[pre + "bc"]))
        for key in ("lo", "di", "up", "alpha", "beta", "kappa", "D"):
            assert np.array_equal(asm[key], g[pre + key], equal_nan=True), (i, key)


@pytest.mark.parametrize("kind", ["uniform", "random", "geometric"])
def test_example_time_loop_matches_golden(kind):
    g = load_golden("cfg1_examples")
    f = O.example_faces(kind)
    assert np.array_equal(f, g[kind + "_faces"])           # incl. the seeded random mesh
    w0 = g[kind + "_w0"]
    for scheme in O.SCHEMES:
        asm = O.assemble(f, 1, 1e-3, scheme, left_value=1., right_value=0.)
        got = O.theta_run_scipy(asm, 0.01, 0.5, w0, 251)
        assert step_error(got, g["%s_%s_theta0.5" % (kind, scheme)], np.max(np.abs(w0))) < 1e-12
        got1 = O.theta_run_scipy(asm, 0.01, 1.0, w0, 251, every=50)
        assert step_error(got1, g["%s_%s_theta1.0" % (kind, scheme)], np.max(np.abs(w0))) < 1e-12


def test_demo_matches_golden_and_survey_value():
    g = load_golden("demo_J149")
    f = O.demo_faces()
    assert np.array_equal(f, g["faces"])
    asm = O.assemble(f, 1, 1e-3, "exponential", left_value=1., right_value=0.)
    assert np.array_equal(asm["kappa"], g["kappa"])
    got = O.theta_run_scipy(asm, 0.01, 1.0, g["w0"], 201)
    assert step_error(got, g["states"], 1.0) < 1e-12
    # conserved "area" printed by the demo (fvm_classes.py:420-421); value from SURVEY.md 8(c)
    assert abs(np.sum(got[-1] * asm["h"]) - 0.9989593187683591) < 1e-12


def test_cfg3_subset_matches_golden():
    g = load_golden("cfg3_subset")
    wl = W.batched_sweep(65536, 1024, first=0, count=8)
    assert np.array_equal(wl["a"], g["a"]) and np.array_equal(wl["d"], g["d"])
    snaps = list(g["snaps"])
    for p in (0, 5):
        asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", **wl["bc"])
        got = O.theta_run_scipy(asm, wl["k"][p], 0.5, wl["w0"][p], 100)
        want = g["states"][p]
        for s in (1, 10, 100):
            assert step_error(got[s - 1], want[snaps.index(s)], 1.1) < 1e-12


def test_thomas_arbiter_agrees_with_scipy_loop():
    """The longdouble no-pivot Thomas loop (not in the reference) against the reference's
    spsolve loop on a well-conditioned graded mesh: agreement far below the 1e-10 target."""
    g = load_golden("cfg3_subset")
    wl = W.batched_sweep(65536, 1024, first=0, count=8)
    asm = [O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", **wl["bc"]) for p in range(8)]
    ops = [O.theta_operators(asm[p], wl["k"][p], 0.5) for p in range(8)]
    A = {k: np.array([o[0][k] for o in ops]) for k in ("lo", "di", "up")}
    B = {k: np.array([o[1][k] for o in ops]) for k in ("lo", "di", "up")}
    beta = np.array([a["beta"] for a in asm])
    got = O.theta_run_thomas(A, B, beta, wl["w0"], 100, every=100)[0]
    want = g["states"][:, list(g["snaps"]).index(100)]
    assert step_error(got, want, 1.1) < 1e-12


def test_conservation_robin_robin():
    """h^T M = 0 for zero-outflow Robin ends (mass conservation)."""
    f = W.graded_faces(200, 50.0)
    for scheme in O.SCHEMES:
        asm = O.assemble(f, 0.7, 0.01, scheme, left_flux=0., right_flux=0.)
        M = O.coefficient_matrix(asm).toarray()
        col = asm["h"] @ M
        assert np.max(np.abs(col)) <= 1e-12 * np.max(np.abs(M))


def test_kappa_special_cases_and_overflow():
    mu = np.array([0.0, np.inf, -np.inf, 1.0, 709.7, 709.8, 1e3])
    with np.errstate(all="ignore"):
        k = O.kappa_reference(mu, np.sign(mu), "exponential")
    assert k[0] == 0 and k[1] == 1 and k[2] == -1                    # fvm_classes.py:118-120
    assert abs(k[4] - 0.99718) < 1e-4 and np.isnan(k[5]) and np.isnan(k[6])
    kf = O.kappa_fitted(mu)
    assert np.allclose(kf[:5], k[:5], rtol=1e-13, atol=0) and np.all(np.isfinite(kf))
    # fitted D equals the reference D where the reference is accurate
    a, d, h = 1.0, 1e-2, np.logspace(-4, np.log10(7.0), 60)        # mu from 1e-2 to 700
    D_ref = d + 0.5 * a * h * O.kappa_reference(a * h / d, np.full(60, a), "exponential")
    D_fit = O.fitted_D(np.full(60, a), np.full(60, d), h)
    assert np.max(np.abs(D_fit - D_ref) / D_ref) < 1e-13
    # ... and stays finite (the upwind limit a h / 2) where the reference overflows to NaN
    assert O.fitted_D(np.array([1.0]), np.array([1e-3]), np.array([1.0]))[0] == pytest.approx(0.5, rel=1e-15)


def test_error_paths():
    with pytest.raises(ValueError):
        O.mesh_arrays([0, 0.5, 0.5, 1])                                   # fvm_classes.py:36-37
    with pytest.raises(UnboundLocalError):
        O.assemble(np.linspace(0, 1, 6), 1, 1, "central")                 # all-None BCs, :260
    with pytest.raises(IndexError):
        O.assemble([0, 0.5, 1], 1, 1, "central", left_value=1, right_value=0)   # J = 2, :292
    with pytest.raises(AttributeError):
        O.assemble(np.linspace(0, 1, 6), 1, 1, "quick", left_value=1, right_value=0)  # :128,131
    # alpha follows left_flux at BOTH ends (:237-238)
    assert O.alpha_ends(left_value=1., right_flux=0.) == (0.0, 0.0)
    assert O.alpha_ends(left_flux=0., right_value=1.) == (1.0, 1.0)
    assert O.alpha_ends(left_flux=0., right_value=1., strict_bc=True) == (1.0, 0.0)


@pytest.mark.skipif(not R.reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_against_live_reference():
    rng = np.random.default_rng(7)
    bcs = [dict(left_value=1., right_value=0.), dict(left_flux=0.3, right_flux=-0.2),
           dict(left_value=1., right_flux=0.1), dict(left_flux=0.2, right_value=0.5)]
    for t in range(24):
        J = int(rng.integers(3, 120))
        f = np.concatenate([[0], np.sort(rng.uniform(0, 1, J - 1)), [1]])
        a = float(rng.normal()) if t % 2 else rng.normal(size=J)
        d = float(10 ** rng.uniform(-3, 0)) if t % 3 else 10 ** rng.uniform(-3, 0, size=J)
        scheme = O.SCHEMES[t % 3]
        bc = bcs[t % 4]
        ref = R.reference_assembly(f, a, d, 0.01, scheme, bc)
        asm = O.assemble(f, a, d, scheme, **bc)
        for key in ("lo", "di", "up", "alpha", "beta", "kappa", "D"):
            assert np.array_equal(ref[key], asm[key], equal_nan=True), (t, key)
    f = W.graded_faces(64, 20.0)
    w0 = W.pulse(0.5 * (f[1:] + f[:-1]))
    ref = R.reference_theta_run(f, 1.0, 5e-3, 0.01, "exponential", bcs[1], 0.5, w0, 20)
    got = O.theta_run_scipy(O.assemble(f, 1.0, 5e-3, "exponential", **bcs[1]), 0.01, 0.5, w0, 20)
    assert np.array_equal(ref, got)


def test_cfg3_long_run_fixture_says_what_the_gpu_tests_rely_on():
    """tests/golden/cfg3_long_run.npz (10,000 Crank-Nicolson steps, the 16 stiffest problems of
    cfg 3): pins the oracle and the C arbiter to it on one problem, and checks the three facts
    tests/test_gpu_long_run.py builds on."""
    g = load_golden("cfg3_long_run")
    faces = W.graded_faces(1024)
    bc = dict(left_flux=0.0, right_flux=0.0)
    rel = lambda x, y: np.max(np.abs(x - y), axis=-1) / np.max(np.abs(y), axis=-1)
    # (1) the reference's own loop is within 1e-10 of the exact solution of its operators, barely
    e = rel(g["reference"], g["exact"])
    assert 5e-11 < e.max() < 1e-10
    # (2) the reference's kappa formula is noisy at these cell Peclet numbers: its D differs from the
    # cancellation-safe evaluation by ~1e-12 relative ...
    i = 2
    asm_ref = O.assemble(faces, g["a"][i], g["d"][i], "exponential", **bc)
    asm_fit = O.assemble(faces, g["a"][i], g["d"][i], "exponential", exponential="fitted", **bc)
    assert np.array_equal(asm_ref["D"], g["D_reference"][i])
    dD = np.abs(asm_ref["D"] - asm_fit["D"]) / asm_fit["D"]
    assert 1e-13 < dD.max() < 5e-12
    # ... which 10,000 stiff steps amplify beyond the 1e-10 parity tolerance
    assert rel(g["exact_fitted"], g["exact"]).max() > 3e-10
    # (3) oracle restatement + arbiter reproduce the fixture (first snapshot: 1,000 steps)
    A, B = O.theta_operators(asm_ref, g["k"][i], 0.5)
    n = int(g["snaps"][0])
    got = O.theta_run_arbiter(A, B, asm_ref["beta"], g["w0"][i], n)[0]
    assert np.array_equal(got, g["exact"][i, 0])
    got = O.theta_run_scipy(asm_ref, g["k"][i], 0.5, g["w0"][i], n, every=n)[0]
    assert rel(got, g["reference"][i, 0]) < 1e-12
    # the longdouble numpy arbiter and the C one agree (short run: the numpy loop is slow)
    short = O.theta_run_arbiter(A, B, asm_ref["beta"], g["w0"][i], 20)[0]
    slow = O.theta_run_thomas({k: v[None] for k, v in A.items()}, {k: v[None] for k, v in B.items()},
                              asm_ref["beta"][None], g["w0"][i][None], 20, every=20)[0][0]
    assert rel(slow, short) < 1e-15
