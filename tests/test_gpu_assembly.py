"""fvm_assemble (CUDA) against the oracle and the reference-generated goldens."""
import numpy as np
import pytest

from conftest import bc_from_array
from oracle import fvm_oracle as O
from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W

pytestmark = pytest.mark.gpu
KEYS = ("lo", "di", "up")


def gpu_assemble(faces, a, d, scheme, bc, interleaved=False, **kw):
    m = BatchedAdvectionDiffusionModel(faces, a, d, 0.01, scheme, **kw)
    m.set_boundary_conditions(**bc)
    out = m.assemble(interleaved=interleaved)
    res = {k: out[k].cpu().numpy() for k in KEYS + ("kappa", "D", "beta_left", "beta_right")}
    if interleaved:
        for k in KEYS + ("kappa", "D"):
            res[k] = res[k].T
    res["alpha"] = (out["alpha_left"], out["alpha_right"])
    return res


def rel(x, y):
    return np.max(np.abs(x - y) / np.maximum(np.abs(y), 1e-300))


def test_golden_cases_central_upwind_bit_exact_exponential_close(golden_assembly):
    """120 reference-generated cases: central/upwind rows, kappa and D are bit-identical to
    the reference; exponential differs only through the Bernoulli form of D (checked against
    the reference where its own formula is accurate, |mu| >= 1e-3)."""
    g = golden_assembly
    for i in range(int(g["n_cases"])):
        pre = "c%03d_" % i
        a, d = g[pre + "a"], g[pre + "d"]
        a = float(a) if a.ndim == 0 else a[None, :]
        d = float(d) if d.ndim == 0 else d[None, :]
        scheme = O.SCHEMES[int(g[pre + "scheme"])]
        bc = bc_from_array(g[pre + "bc"])
        out = gpu_assemble(g[pre + "faces"], a, d, scheme, bc, P=1)
        J = len(g[pre + "di"])
        assert out["alpha"] == (int(g[pre + "alpha"][0]), int(g[pre + "alpha"][-1]))
        assert out["beta_left"][0] == g[pre + "beta"][0] and out["beta_right"][0] == g[pre + "beta"][J - 1]
        if scheme != "exponential":
            for k in KEYS + ("kappa", "D"):
                assert np.array_equal(out[k][0], g[pre + k]), (i, scheme, k)
        else:
            h = np.diff(g[pre + "faces"])
            mu = np.broadcast_to(np.asarray(a), (1, J))[0] * h / np.broadcast_to(np.asarray(d), (1, J))[0]
            ok = np.abs(mu) >= 1e-3
            assert rel(out["D"][0][ok], g[pre + "D"][ok]) < 1e-12, (i, "D")
            assert np.max(np.abs(out["kappa"][0][ok] - g[pre + "kappa"][ok])) < 1e-9, (i, "kappa")
            # the rows against the oracle fed with the fitted D (same formulas, same order)
            asm = O.assemble(g[pre + "faces"], np.squeeze(a), np.squeeze(d), scheme, exponential="fitted", **bc)
            assert rel(out["D"][0], asm["D"]) < 4e-16 * 8, (i, "D fitted")
            scale = np.max(np.abs(asm["di"]))
            for k in KEYS:
                assert np.max(np.abs(out[k][0] - asm[k])) <= 1e-14 * scale, (i, k)


def test_batched_layouts_and_strides():
    """(P,) scalars, (J,) shared arrays and (P, J) arrays, problem-major and interleaved
    outputs all give the oracle's rows."""
    rng = np.random.default_rng(3)
    J, P = 37, 5
    f = W.graded_faces(J, 20.0)
    a_p, d_p = rng.normal(size=P), 10 ** rng.uniform(-2, 0, P)
    a_j, d_j = rng.normal(size=J), 10 ** rng.uniform(-2, 0, J)
    a_pj, d_pj = rng.normal(size=(P, J)), 10 ** rng.uniform(-2, 0, (P, J))
    bc = dict(left_flux=0.25, right_flux=-0.5)
    for a, d in ((a_p, d_p), (a_j, d_j), (a_pj, d_pj), (a_p, d_pj), (0.3, d_j)):
        for inter in (False, True):
            out = gpu_assemble(f, a, d, "upwind", bc, interleaved=inter, P=P)
            for p in range(P):
                ap = a if np.ndim(a) == 0 else (a[p] if np.shape(a) in ((P,), (P, J)) else a)
                dp = d if np.ndim(d) == 0 else (d[p] if np.shape(d) in ((P,), (P, J)) else d)
                asm = O.assemble(f, ap, dp, "upwind", **bc)
                for k in KEYS + ("D",):
                    assert np.array_equal(out[k][p], asm[k]), (k, p, inter)
                assert out["beta_left"][p] == asm["beta"][0] and out["beta_right"][p] == asm["beta"][-1]


def test_per_problem_flux_and_mixed_bcs():
    J, P = 12, 4
    f = np.linspace(0, 2, J + 1)
    lf = np.array([0.0, 0.5, -1.0, 2.0])
    out = gpu_assemble(f, 1.0, 0.1, "central", dict(left_flux=lf, right_value=0.0), P=P)
    for p in range(P):
        asm = O.assemble(f, 1.0, 0.1, "central", left_flux=lf[p], right_value=0.0)
        for k in KEYS:
            assert np.array_equal(out[k][p], asm[k])
        assert out["beta_left"][p] == asm["beta"][0] and out["beta_right"][p] == 0.0
    assert out["alpha"] == (1, 1)                       # the reference's quirk: both ends follow left_flux
    out = gpu_assemble(f, 1.0, 0.1, "central", dict(left_flux=lf, right_value=0.0), P=P, strict_bc=True)
    assert out["alpha"] == (1, 0)


def test_reaction_term_on_non_dirichlet_rows():
    J = 20
    f = W.graded_faces(J, 5.0)
    r = np.linspace(1.0, 3.0, J)
    for bc in (dict(left_value=1.0, right_flux=0.0), dict(left_flux=0.0, right_flux=0.0)):
        out = gpu_assemble(f, 0.5, 0.05, "upwind", bc, P=1, reaction=r)
        asm = O.assemble(f, 0.5, 0.05, "upwind", reaction=r, **bc)
        for k in KEYS:
            assert np.array_equal(out[k][0], asm[k])


def test_exponential_special_cases_and_large_peclet():
    """mu = 0, huge |mu| (where the reference returns NaN) and both signs of a."""
    J = 16
    f = np.linspace(0, 1, J + 1)
    h = 1.0 / J
    a = np.array([0.0, 1.0, -1.0, 1.0, -1.0, 1.0])
    d = np.array([1.0, 1.0, 1.0, 1e-6, 1e-6, 0.0])
    out = gpu_assemble(f, a, d, "exponential", dict(left_value=1.0, right_value=0.0), P=6)
    assert np.all(out["kappa"][0] == 0) and np.all(out["D"][0] == 1.0)             # mu == 0, :118
    mu = a[1] * h / d[1]
    assert np.allclose(out["kappa"][1], 1 / np.tanh(mu / 2) - 2 / mu, rtol=1e-12)
    assert np.allclose(out["kappa"][2], -out["kappa"][1], rtol=0, atol=1e-15)
    # mu = 62500 >= 709.8: reference NaN; fitted form gives the upwind limit D = |a| h / 2
    assert np.allclose(out["D"][3], 0.5 * h, rtol=1e-4) and np.allclose(out["D"][4], 0.5 * h, rtol=1e-4)
    assert np.all(np.isfinite(out["di"][3])) and np.all(np.isfinite(out["di"][4]))
    assert np.allclose(out["kappa"][3], 1.0, atol=1e-4) and np.allclose(out["kappa"][4], -1.0, atol=1e-4)
    # d == 0: mu = +inf -> kappa = 1 (:119), D = a h / 2
    assert np.all(out["kappa"][5] == 1.0) and np.allclose(out["D"][5], 0.5 * h, rtol=1e-15)
    fit = O.fitted_D(np.full(J, a[3]), np.full(J, d[3]), np.full(J, h))
    assert rel(out["D"][3], fit) < 1e-15


def test_peclet_sweep_rows_against_oracle():
    """cfg 2, all 29 Peclet numbers x 3 schemes on the 4096-cell graded mesh."""
    wl = W.peclet_sweep(4096)
    for scheme in O.SCHEMES:
        out = gpu_assemble(wl["faces"], wl["a"], wl["d"], scheme, wl["bc"], P=29)
        for p in range(29):
            exact = scheme != "exponential"
            asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], scheme,
                             exponential="reference" if exact else "fitted", **wl["bc"])
            scale = np.max(np.abs(asm["di"]))
            for k in KEYS:
                if exact:
                    assert np.array_equal(out[k][p], asm[k]), (scheme, p, k)
                else:
                    assert np.max(np.abs(out[k][p] - asm[k])) <= 1e-14 * scale, (scheme, p, k)
            if not exact and wl["pe"][p] < 700:
                ref = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], scheme, **wl["bc"])
                # the reference's own cancellation error in D/d grows as mu -> 0
                assert rel(out["D"][p], ref["D"]) < 2e-10, (p, rel(out["D"][p], ref["D"]))


def test_conservation_of_assembled_rows():
    f = W.graded_faces(300, 100.0)
    out = gpu_assemble(f, np.array([0.5, 2.0]), np.array([1e-3, 1.0]), "exponential",
                       dict(left_flux=0.0, right_flux=0.0), P=2)
    h = np.diff(f)
    for p in range(2):
        col = h * out["di"][p]
        col[:-1] += h[1:] * out["lo"][p][1:]
        col[1:] += h[:-1] * out["up"][p][:-1]
        assert np.max(np.abs(col)) <= 1e-12 * np.max(np.abs(out["di"][p]))
