"""Synthetic inputs of the BASELINE.json configs.

Input generators only -- no solver arithmetic lives here.  Everything is seeded
and host-side numpy so that the CUDA path, the oracle and the committed golden
fixtures see identical bits.
"""
from __future__ import annotations

import numpy as np


def graded_faces(J, ratio=100.0):
    """Smoothly graded mesh: h_i proportional to r^i, h_0/h_{J-1} = ratio, faces on [0, 1].
    (Wild width ratios make the 1e-10 parity tolerance meaningless)"""
    r = ratio ** (-1.0 / (J - 1))
    h = r ** np.arange(J)
    f = np.concatenate([[0.0], np.cumsum(h)])
    return f / f[-1]


def pulse(cells):
    """Initial condition of the example driver, sin(pi x)^100 [fvm_examples.py:75]."""
    return np.sin(np.pi * np.asarray(cells)) ** 100


def peclet_sweep(J=4096, n_pe=29, pe_lo=-3.0, pe_hi=4.0):
    """cfg 2: Peclet sweep on a graded mesh, a = 1, d = a h_max / Pe, dt = 4 h_min / a,
    Dirichlet flags, w_init = sin(pi c)^100 with the ends set to (1, 0)."""
    faces = graded_faces(J)
    h = np.diff(faces)
    c = 0.5 * (faces[1:] + faces[:-1])
    pe = np.logspace(pe_lo, pe_hi, n_pe)
    a = np.ones(n_pe)
    d = a * h.max() / pe
    k = np.full(n_pe, 4.0 * h.min())
    w0 = pulse(c)
    w0[0], w0[-1] = 1.0, 0.0
    return dict(faces=faces, a=a, d=d, k=k, pe=pe, w0=np.tile(w0, (n_pe, 1)), theta=0.5,
                bc=dict(left_value=1.0, right_value=0.0))


def batched_sweep(P=65536, J=1024, seed=0, first=0, count=None):
    """cfg 3: P independent problems on one graded mesh with per-problem scalars
    a ~ U(0.5, 2), d = a h_max / Pe with Pe ~ logU(1e-2, 1e2), dt = CFL h_min / a with
    CFL ~ U(1, 8); exponential fitting, zero-flux Robin ends, Crank-Nicolson.
    ``first``/``count`` select a contiguous shard of the P problems (multi-GPU: the
    scalars of problem p do not depend on how the batch is sharded).
    Initial state: a Gaussian pulse whose centre varies with the problem, on a floor
    of 0.1 so that every cell carries mass."""
    faces = graded_faces(J)
    h = np.diff(faces)
    c = 0.5 * (faces[1:] + faces[:-1])
    rng = np.random.default_rng(seed)
    a = rng.uniform(0.5, 2.0, P)
    pe = 10.0 ** rng.uniform(-2.0, 2.0, P)
    cfl = rng.uniform(1.0, 8.0, P)
    x0 = rng.uniform(0.2, 0.8, P)
    d = a * h.max() / pe
    k = cfl * h.min() / a
    count = P - first if count is None else count
    sl = slice(first, first + count)
    a, d, k, x0 = a[sl], d[sl], k[sl], x0[sl]
    w0 = 0.1 + np.exp(-((c[None, :] - x0[:, None]) / 0.05) ** 2)
    return dict(faces=faces, a=a, d=d, k=k, w0=w0, theta=0.5,
                bc=dict(left_flux=0.0, right_flux=0.0))


def stiff_reaction(P=4096, J=1 << 20, seed=0, first=0, count=None):
    """cfg 4: long systems, per-problem a, d as cfg 3 plus reaction r ~ logU(1, 1e4),
    fully implicit (theta = 1), dt = 100 h_min / a, zero-flux Robin ends."""
    faces = graded_faces(J)
    h = np.diff(faces)
    c = 0.5 * (faces[1:] + faces[:-1])
    rng = np.random.default_rng(seed)
    a = rng.uniform(0.5, 2.0, P)
    pe = 10.0 ** rng.uniform(-2.0, 2.0, P)
    r = 10.0 ** rng.uniform(0.0, 4.0, P)
    x0 = rng.uniform(0.2, 0.8, P)
    d = a * h.max() / pe
    k = 100.0 * h.min() / a
    count = P - first if count is None else count
    sl = slice(first, first + count)
    a, d, k, r, x0 = a[sl], d[sl], k[sl], r[sl], x0[sl]
    w0 = 0.1 + np.exp(-((c[None, :] - x0[:, None]) / 0.05) ** 2)
    return dict(faces=faces, a=a, d=d, k=k, reaction=r, w0=w0, theta=1.0,
                bc=dict(left_flux=0.0, right_flux=0.0))


def graded_slab(J, j0, j1, device, ratio=100.0):
    """cfg 5: the cells [j0, j1) of a J-cell graded mesh (h_0/h_{J-1} = ratio) generated ON THE
    DEVICE from the closed form of the geometric series, faces_j = (1 - r^j)/(1 - r^J), so that
    no rank ever holds the whole 2^30-cell mesh and every rank computes identical values for the
    faces it shares with its neighbours.  Returns the local arrays in the layout
    PartitionedAdvectionDiffusionModel(J=...) expects (one ghost face/cell on each side where the
    mesh continues), the scalar d for Pe_max = 2 with a = 1, dt = 4 h_min / a, and the initial
    state sin(pi c)^100 + 0.1 for the owned cells."""
    import math
    import torch
    lr = -math.log(ratio) / (J - 1)
    f_lo, f_hi = max(j0 - 1, 0), min(j1 + 1, J)
    j = torch.arange(f_lo, f_hi + 1, device=device, dtype=torch.float64)
    faces = torch.expm1(j * lr) / math.expm1(J * lr)
    h_max = math.expm1(lr) / math.expm1(J * lr)              # width of cell 0
    h_min = h_max / ratio
    own = faces[(j0 - f_lo):(j1 - f_lo + 1)]
    c = 0.5 * (own[1:] + own[:-1])
    w0 = torch.sin(math.pi * c) ** 100 + 0.1
    return dict(faces=faces, d_pe2=h_max / 2.0, dt=4.0 * h_min, w0=w0[None, :], h_min=h_min, h_max=h_max)
