"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED
reference (/root/reference/fvm_classes.py, loaded through oracle/ref_loader.py)
in the build container.  The reference ships no tests or golden vectors of its
own, so these files are what pins the oracle -- and,
through it, the CUDA path -- to the reference's behaviour.

    python tests/golden/make_golden.py        # needs /root/reference; ~5 min

Every array below comes from the reference's own classes and the literal loop
of fvm_classes.py:373-377,408-411 (scipy spsolve); nothing is computed by the
oracle restatement or by the product.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_loader as R            # noqa: E402
from oracle import fvm_oracle as O            # noqa: E402  (mesh generators only)
from fvm_b200 import workloads as W           # noqa: E402

BCS = [dict(left_value=1., right_value=0.), dict(left_flux=0.3, right_flux=-0.2),
       dict(left_value=1., right_flux=0.1), dict(left_flux=0.2, right_value=0.5),
       dict(left_flux=0.0, right_flux=0.0)]


def bc_arrays(bc):
    """[left_flux, right_flux, left_value, right_value] with NaN for None."""
    return np.array([np.nan if bc.get(key) is None else bc[key]
                     for key in ("left_flux", "right_flux", "left_value", "right_value")])


def assembly_cases():
    rng = np.random.default_rng(1234)
    out = {}
    n = 0
    for t in range(40):
        J = int(rng.integers(3, 64))
        kind = t % 3
        if kind == 0:
            f = np.linspace(0, 1, J + 1)
        elif kind == 1:
            f = W.graded_faces(J, 30.0)
        else:
            f = np.concatenate([[0], np.sort(rng.uniform(0, 1, J - 1)), [1]])
        scal = t % 2 == 0
        a = float(rng.normal()) if scal else rng.normal(size=J)
        d = float(10 ** rng.uniform(-3, 0)) if scal else 10 ** rng.uniform(-3, 0, size=J)
        for scheme in O.SCHEMES:
            bc = BCS[n % len(BCS)]
            r = R.reference_assembly(f, a, d, 0.01, scheme, bc)
            pre = "c%03d_" % n
            out[pre + "faces"] = f
            out[pre + "a"] = np.asarray(a, dtype=float)
            out[pre + "d"] = np.asarray(d, dtype=float)
            out[pre + "scheme"] = np.array(O.SCHEMES.index(scheme))
            out[pre + "bc"] = bc_arrays(bc)
            for key in ("lo", "di", "up", "alpha", "beta", "kappa", "D"):
                out[pre + key] = r[key]
            n += 1
    out["n_cases"] = np.array(n)
    return out


def cfg1_examples():
    """fvm_examples.py flow on its three meshes (old-API shim):
    a=1, d=1e-3, k=0.01, Dirichlet flags, w_init=sin(pi c)^100, 251 steps."""
    out = {}
    bc = dict(left_value=1., right_value=0.)
    for kind in ("uniform", "random", "geometric"):
        f = O.example_faces(kind)
        c = 0.5 * (f[1:] + f[:-1])
        w0 = np.sin(np.pi * c) ** 100
        out[kind + "_faces"] = f
        out[kind + "_w0"] = w0
        for scheme in O.SCHEMES:
            out["%s_%s_theta0.5" % (kind, scheme)] = R.reference_theta_run(
                f, 1, 1e-3, 0.01, scheme, bc, 0.5, w0, 251, every=1)
            out["%s_%s_theta1.0" % (kind, scheme)] = R.reference_theta_run(
                f, 1, 1e-3, 0.01, scheme, bc, 1.0, w0, 251, every=50)
    return out


def demo_J149():
    """The shipped demo, fvm_classes.py:346-411: J=149, exponential, theta=1, 201 steps."""
    f = O.demo_faces()
    c = 0.5 * (f[1:] + f[:-1])
    w0 = np.sin(np.pi * c) ** 100
    w0[0] = 1.0
    bc = dict(left_value=1., right_value=0.)
    states = R.reference_theta_run(f, 1, 1e-3, 0.01, "exponential", bc, 1.0, w0, 201, every=1)
    asm = R.reference_assembly(f, 1, 1e-3, 0.01, "exponential", bc)
    return dict(faces=f, w0=w0, states=states, kappa=asm["kappa"], D=asm["D"],
                lo=asm["lo"], di=asm["di"], up=asm["up"])


def cfg3_subset():
    """cfg 3 on its first 8 problems: J=1024 graded mesh, exponential, zero-flux Robin,
    Crank-Nicolson, 2000 steps; snapshots after 1, 10, 100, 1000, 2000."""
    wl = W.batched_sweep(65536, 1024, first=0, count=8)
    snaps = [1, 10, 100, 1000, 2000]
    states = []
    for p in range(8):
        st = R.reference_theta_run(wl["faces"], wl["a"][p], wl["d"][p], wl["k"][p], "exponential",
                                   wl["bc"], 0.5, wl["w0"][p], 2000, every=1)
        states.append(st[[s - 1 for s in snaps]])
        print("cfg3", p, flush=True)
    return dict(snaps=np.array(snaps), states=np.array(states), a=wl["a"], d=wl["d"], k=wl["k"])


def _cfg2_one(args):
    faces, a, d, k, scheme, bc, w0, snaps = args
    st = R.reference_theta_run(faces, a, d, k, scheme, bc, 0.5, w0, snaps[-1], every=1)
    return st[[s - 1 for s in snaps]]


def cfg2_full():
    """cfg 2 in full: J=4096 graded mesh, all 29 Peclet numbers x 3 schemes = 87 problems, 2000
    Crank-Nicolson steps with the reference's loop; snapshots after 100 and 2000 steps.  For the five
    Peclet numbers >= 709.8 the reference's exponential kappa is NaN (fvm_classes.py:117: inf/inf) and
    so is its solution; `exponential_nan` lists them, and `exponential_fitted_XX` holds, for those, the
    exact solution (oracle/thomas_arbiter.c) of the operators assembled with the cancellation- and
    overflow-safe form D = d (B(mu) + mu/2) (oracle.fitted_D) -- what the CUDA path is checked against
    where the reference has no answer."""
    from concurrent.futures import ProcessPoolExecutor
    wl = W.peclet_sweep(4096)
    snaps = [100, 2000]
    out = dict(snaps=np.array(snaps), pe=wl["pe"])
    jobs, keys = [], []
    for scheme in O.SCHEMES:
        for j in range(len(wl["pe"])):
            jobs.append((wl["faces"], wl["a"][j], wl["d"][j], wl["k"][j], scheme, wl["bc"], wl["w0"][j], snaps))
            keys.append("%s_%02d" % (scheme, j))
    with ProcessPoolExecutor(min(8, os.cpu_count() or 1)) as ex:
        for key, st in zip(keys, ex.map(_cfg2_one, jobs)):
            out[key] = st
    nan = [j for j in range(len(wl["pe"])) if not np.all(np.isfinite(out["exponential_%02d" % j]))]
    out["exponential_nan"] = np.array(nan)
    for j in nan:
        del out["exponential_%02d" % j]
        asm = O.assemble(wl["faces"], wl["a"][j], wl["d"][j], "exponential", exponential="fitted", **wl["bc"])
        A, B = O.theta_operators(asm, wl["k"][j], 0.5)
        out["exponential_fitted_%02d" % j] = O.theta_run_arbiter(A, B, asm["beta"], wl["w0"][j], snaps[-1], snaps)
    return out


LONG_RUN_STEPS = 10000
LONG_RUN_SNAPS = [1000, 5000, 10000]


def long_run_problems(n_each=8):
    """The problems of cfg 3's 65,536 where round-off survives longest: the n_each smallest a/d
    (most diffusive: largest dt*d/h^2 everywhere) and the n_each largest dt*d (stiffest fine end)."""
    wl = W.batched_sweep(65536, 1024)
    pe = wl["a"] / wl["d"]
    kd = wl["k"] * wl["d"]
    first = list(np.argsort(pe)[:n_each])
    rest = [p for p in np.argsort(-kd) if p not in first][:n_each]
    return wl, np.array(sorted(first + rest))


def _long_run_one(args):
    faces, a, d, k, bc, w0 = args
    return R.reference_theta_run(faces, a, d, k, "exponential", bc, 0.5, w0, LONG_RUN_STEPS,
                                 every=1)[[s - 1 for s in LONG_RUN_SNAPS]]


def cfg3_long_run():
    """cfg 3 at its FULL run length: the reference's literal scipy loop for all 10,000
    Crank-Nicolson steps on the 16 stiffest problems of the 65,536-problem sweep, snapshots after
    1,000 / 5,000 / 10,000 steps (`reference`), with the reference's own modified diffusivity of
    every cell (`D_reference`, fvm_classes.py:117-131).  Beside it:
      exact        -- the exact solution of the reference's fp64 operators (oracle/thomas_arbiter.c,
                      __float128 Thomas).  The reference's SuperLU loop is 2e-12 ... 9e-11 from it
                      after 10,000 steps: that is what a tolerance of 1e-10 has to absorb.
      exact_fitted -- the same with the exponential fitting evaluated in the cancellation-safe
                      form D = d (B(mu) + mu/2) (oracle.fitted_D; what north_star asks the CUDA path
                      to use).  It is up to 7e-10 from `exact`: at these cell Peclet numbers
                      (1e-4 ... 1e-2) the reference's kappa formula (:117) carries a cancellation
                      error of ~1e-12 relative in D, per cell and depending on the libm behind
                      numpy.exp, and 10,000 steps of a stiff operator amplify it that far."""
    from concurrent.futures import ProcessPoolExecutor
    wl, idx = long_run_problems()
    jobs = [(wl["faces"], wl["a"][p], wl["d"][p], wl["k"][p], wl["bc"], wl["w0"][p]) for p in idx]
    with ProcessPoolExecutor(min(8, os.cpu_count() or 1)) as ex:
        ref = np.array(list(ex.map(_long_run_one, jobs)))
    exact, exact_fitted, D_ref = [], [], []
    for p in idx:
        D_ref.append(R.reference_assembly(wl["faces"], wl["a"][p], wl["d"][p], wl["k"][p], "exponential",
                                          wl["bc"])["D"])
        for mode, out in (("reference", exact), ("fitted", exact_fitted)):
            asm = O.assemble(wl["faces"], wl["a"][p], wl["d"][p], "exponential", exponential=mode, **wl["bc"])
            if mode == "reference":
                assert np.array_equal(asm["D"], D_ref[-1])
            A, B = O.theta_operators(asm, wl["k"][p], 0.5)
            out.append(O.theta_run_arbiter(A, B, asm["beta"], wl["w0"][p], LONG_RUN_STEPS, LONG_RUN_SNAPS))
        print("long run exact", p, flush=True)
    return dict(problems=idx, snaps=np.array(LONG_RUN_SNAPS), reference=ref, exact=np.array(exact),
                exact_fitted=np.array(exact_fitted), D_reference=np.array(D_ref),
                a=wl["a"][idx], d=wl["d"][idx], k=wl["k"][idx], w0=wl["w0"][idx])


if __name__ == "__main__":
    only = sys.argv[1:]
    jobs = dict(assembly_cases=assembly_cases, cfg1_examples=cfg1_examples, demo_J149=demo_J149,
                cfg3_subset=cfg3_subset, cfg3_long_run=cfg3_long_run, cfg2_full=cfg2_full)
    for name, fn in jobs.items():
        if only and name not in only:
            continue
        t0 = time.time()
        data = fn()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **data)
        print(name, "%.1fs" % (time.time() - t0), flush=True)
