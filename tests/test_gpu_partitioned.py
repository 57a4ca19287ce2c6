"""Variant 3 (partitioned SPIKE/Thomas for long systems, fvm_spike_*) against the oracle's
spsolve loop, against the fused kernel, and with the cells split over two ranks."""
import os
import socket

import numpy as np
import pytest
import torch

from conftest import step_error
from oracle import fvm_oracle as O
from fvm_b200 import (BatchedAdvectionDiffusionModel, PartitionedAdvectionDiffusionModel, plan_partition,
                      workloads as W)

pytestmark = pytest.mark.gpu
TOL = 1e-10


def oracle_run(f, a, d, k, scheme, bc, theta, w0, nsteps, reaction=None):
    asm = O.assemble(f, a, d, scheme, reaction=reaction, **bc)
    return O.theta_run_scipy(asm, k, theta, w0, nsteps, every=nsteps)[0]


@pytest.mark.parametrize("J", [200, 255, 256, 257, 511, 1000, 3001])
def test_small_tiles_against_oracle(J):
    """Tile configuration 1 (256 cells per tile): one to twelve tiles, ragged ends, every BC
    combination with nonzero fluxes, three values of theta."""
    rng = np.random.default_rng(J)
    f = W.graded_faces(J, 20.0)
    c = 0.5 * (f[1:] + f[:-1])
    h = np.diff(f)
    P = 2
    a = rng.uniform(0.5, 2.0, P) * np.array([1.0, -1.0])
    d = np.abs(a) * h.max() / rng.uniform(0.3, 1.5, P)
    k = rng.uniform(1.0, 6.0, P) * h.min()
    w0 = 0.2 + np.exp(-((c[None, :] - rng.uniform(0.3, 0.7, (P, 1))) / 0.1) ** 2)
    bcs = [dict(left_flux=0.3, right_flux=-0.2), dict(left_value=1., right_value=0.),
           dict(left_value=1., right_flux=0.1), dict(left_flux=0.2, right_value=0.5)]
    for bc, theta, scheme in zip(bcs, (0.5, 1.0, 0.7, 0.5), ("exponential", "upwind", "central", "exponential")):
        m = PartitionedAdvectionDiffusionModel(f, a, d, k, scheme, P=P, config=1)
        m.set_boundary_conditions(**bc)
        got = m.run(w0, theta, 25).cpu().numpy()
        for p in range(P):
            want = oracle_run(f, a[p], d[p], k[p], scheme, bc, theta, w0[p], 25)
            assert step_error(got[p], want, np.max(np.abs(w0[p]))) < TOL, (J, bc, theta, p)


@pytest.mark.parametrize("config", [0, 2])
def test_production_tiles_against_oracle(config):
    """Tile configurations 0 (8192 cells per tile, three tiles) and 2 (4096 cells, two CTAs per SM, the
    default; five tiles), per-cell coefficients, large dt."""
    J = 20000
    f = W.graded_faces(J, 100.0)
    c = 0.5 * (f[1:] + f[:-1])
    h = np.diff(f)
    a = 1.0 + 0.3 * np.sin(2 * np.pi * c)
    d = np.full(J, h.max() / 2.0)
    k = 40.0 * h.min()
    w0 = W.pulse(c) + 0.1
    bc = dict(left_flux=0.0, right_flux=0.0)
    m = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", config=config)
    m.set_boundary_conditions(**bc)
    got = m.run(w0[None, :], 0.5, 20).cpu().numpy()[0]
    want = oracle_run(f, a, d, k, "exponential", bc, 0.5, w0, 20)
    assert step_error(got, want, 1.1) < TOL
    assert abs(np.sum(got * h) - np.sum(w0 * h)) < 1e-12          # zero-flux Robin ends conserve mass


def test_batched_stiff_reaction_cfg4_shape():
    """cfg 4 in miniature: P long problems, per-problem a/d/reaction, fully implicit, dt = 100 h_min/a."""
    wl = W.stiff_reaction(P=3, J=12000, seed=3)
    m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=3,
                                           reaction=wl["reaction"], config=0)
    m.set_boundary_conditions(**wl["bc"])
    got = m.run(wl["w0"], 1.0, 10).cpu().numpy()
    for p in range(3):
        want = oracle_run(wl["faces"], wl["a"][p], wl["d"][p], wl["k"][p], "exponential", wl["bc"], 1.0,
                          wl["w0"][p], 10, reaction=wl["reaction"][p])
        assert step_error(got[p], want, 1.1) < TOL


def test_agrees_with_fused_kernel_and_untruncated_spikes():
    wl = W.batched_sweep(65536, 1024, first=7, count=5)
    fused = BatchedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential")
    fused.set_boundary_conditions(**wl["bc"])
    wf, _ = fused.run(wl["w0"], 0.5, 100, variant="fused")
    res = {}
    for tol in (1e-30, 0.0):
        m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=5,
                                               config=1, spike_tol=tol)
        m.set_boundary_conditions(**wl["bc"])
        res[tol] = m.run(wl["w0"], 0.5, 100)
        assert step_error(res[tol].cpu().numpy(), wf.cpu().numpy(), 1.1) < 1e-11
    # truncating the spikes at 1e-30 cannot be seen in fp64
    assert float((res[1e-30] - res[0.0]).abs().max()) < 1e-15


def test_plan_partition():
    pl = plan_partition(20000, 8192, 1)
    assert pl["tiles_per_rank"] == 3 and pl["ranges"] == [(0, 20000)]
    pl = plan_partition(20000, 8192, 2)
    assert pl["tiles_per_rank"] == 2 and pl["tiles_global"] == 4 and pl["ranges"] == [(0, 16384), (16384, 20000)]
    pl = plan_partition(1000, 256, 8)                    # more ranks than tiles: trailing ranks hold padding only
    assert pl["tiles_per_rank"] == 1 and pl["ranges"][3] == (768, 1000) and pl["ranges"][5] == (1000, 1000)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _two_rank_worker(rank, world, port, J, out_dir, blocks):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(0)
    f = W.graded_faces(J, 30.0)
    c = 0.5 * (f[1:] + f[:-1])
    h = np.diff(f)
    a = np.array([1.0, -0.7])
    d = np.abs(a) * h.max() / 1.2
    k = np.array([5.0, 3.0]) * h.min()
    w0 = 0.2 + np.exp(-((c[None, :] - np.array([[0.4], [0.6]])) / 0.1) ** 2)
    bc = dict(left_flux=0.3, right_flux=-0.2)
    m = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", P=2, config=1, group=dist.group.WORLD,
                                           blocks=blocks)
    m.set_boundary_conditions(**bc)
    got = m.run(w0, 0.5, 30).cpu().numpy()
    np.save(os.path.join(out_dir, "rank%d.npy" % rank), got)
    np.save(os.path.join(out_dir, "range%d.npy" % rank), np.array([m.j0, m.j1]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,J,blocks", [(2, 2000, None), (3, 1500, None), (2, 2000, 2), (2, 3000, 3)])
def test_cells_split_over_ranks(tmp_path, world, J, blocks):
    """The multi-GPU path with the ranks sharing this one GPU (collectives over gloo): the slabs'
    results, put side by side, equal the single-rank run and the oracle.  `blocks`: every rank's part of
    the separator system cut into that many blocks (4 or 6 tiles per rank here: 2 unknowns per block)."""
    import torch.multiprocessing as mp
    mp.spawn(_two_rank_worker, args=(world, _free_port(), J, str(tmp_path), blocks), nprocs=world, join=True)
    f = W.graded_faces(J, 30.0)
    c = 0.5 * (f[1:] + f[:-1])
    h = np.diff(f)
    a = np.array([1.0, -0.7]); d = np.abs(a) * h.max() / 1.2; k = np.array([5.0, 3.0]) * h.min()
    w0 = 0.2 + np.exp(-((c[None, :] - np.array([[0.4], [0.6]])) / 0.1) ** 2)
    bc = dict(left_flux=0.3, right_flux=-0.2)
    parts = [np.load(tmp_path / ("rank%d.npy" % r)) for r in range(world)]
    ranges = [tuple(np.load(tmp_path / ("range%d.npy" % r))) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == J and all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
    got = np.concatenate(parts, axis=1)
    single = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", P=2, config=1)
    single.set_boundary_conditions(**bc)
    ws = single.run(w0, 0.5, 30).cpu().numpy()
    assert np.max(np.abs(got - ws)) < 1e-13
    for p in range(2):
        want = oracle_run(f, a[p], d[p], k[p], "exponential", bc, 0.5, w0[p], 30)
        assert step_error(got[p], want, 1.3) < TOL


def test_device_mesh_check():
    f = W.graded_faces(5000, 10.0)
    m = PartitionedAdvectionDiffusionModel(f, 1.0, 0.1, 1e-3, "central", config=1)
    assert m.check_mesh() == 0
    g = f.copy(); g[1234] = g[1233]                           # duplicate face -> zero-width cell
    with pytest.raises(ValueError):
        PartitionedAdvectionDiffusionModel(g, 1.0, 0.1, 1e-3, "central", config=1).check_mesh()
    g = f.copy(); g[[100, 101]] = g[[101, 100]]               # unsorted: accepted, reported
    assert PartitionedAdvectionDiffusionModel(g, 1.0, 0.1, 1e-3, "central", config=1).check_mesh() == 1


def test_lazy_and_eager_correction_agree():
    """The spike correction applied inside the next tile solve (lazy, the default) and as a separate
    pass after every step (lazy=False) are the same arithmetic on the same numbers."""
    wl = W.stiff_reaction(P=2, J=9000, seed=5)          # non-decaying spikes: every cell is corrected
    res = {}
    for lazy in (True, False):
        m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=2,
                                               reaction=wl["reaction"], config=1, lazy=lazy)
        m.set_boundary_conditions(**wl["bc"])
        m.prepare(1.0)
        m.set_state(wl["w0"])
        m.step(3)
        mid = m.get_state()                               # materialises, then the loop goes on
        m.step(4)
        res[lazy] = (mid, m.get_state())
    for a, b in zip(res[True], res[False]):
        assert float((a - b).abs().max()) <= 1e-14 * float(b.abs().max())


@pytest.mark.parametrize("blocks", [2, 3, 6])
def test_separator_system_in_blocks(blocks):
    """One GPU, the separator system cut into blocks that are solved in parallel (what cfg 5 runs with
    128 blocks of 256 separators): same result as the one-CTA solve of the whole system and as the oracle,
    with spikes that do not decay (every block separator is coupled to its neighbours)."""
    wl = W.stiff_reaction(P=3, J=12 * 256 - 7, seed=11)
    res = {}
    for b in (1, blocks):
        m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=3,
                                               reaction=wl["reaction"], config=1, blocks=b)
        m.set_boundary_conditions(**wl["bc"])
        res[b] = m.run(wl["w0"], 1.0, 12).cpu().numpy()
        assert m.hier == (b > 1) and m.nt == 12
    assert np.max(np.abs(res[blocks] - res[1])) <= 1e-13 * np.max(np.abs(res[1]))
    for p in range(3):
        want = oracle_run(wl["faces"], wl["a"][p], wl["d"][p], wl["k"][p], "exponential", wl["bc"], 1.0,
                          wl["w0"][p], 12, reaction=wl["reaction"][p])
        assert step_error(res[blocks][p], want, 1.1) < TOL
    bad = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=3,
                                             reaction=wl["reaction"], config=1, blocks=5)
    bad.set_boundary_conditions(**wl["bc"])
    with pytest.raises(ValueError):
        bad.prepare(1.0)


def test_one_call_time_loop_equals_the_step_by_step_host_loop():
    """fvm_spike_steps (the whole loop in one host call, what step() uses) against the same kernels driven
    entry point by entry point from Python (the path a gloo group takes): identical bits, with and without
    the block level, lazy and eager."""
    wl = W.stiff_reaction(P=2, J=12 * 256 - 7, seed=7)
    for blocks, lazy in ((1, True), (3, True), (3, False), (1, False)):
        res = []
        for host_loop in (False, True):
            m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=2,
                                                   reaction=wl["reaction"], config=1, blocks=blocks, lazy=lazy)
            m.set_boundary_conditions(**wl["bc"])
            m.prepare(1.0)
            m.set_state(wl["w0"])
            if host_loop:
                m.comm.timing = []              # a timing request sends step() through the Python loop
            m.step(3)
            mid = m.get_state()
            m.step(4)
            res.append((mid, m.get_state()))
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1]), (blocks, lazy)


@pytest.mark.parametrize("config,J,blocks", [(1, 3000, 3), (2, 9000, None), (0, 20000, None)])
def test_no_stray_writes_around_the_state_buffers(config, J, blocks):
    """compute-sanitizer is not available on the GPU pool: guard bands of NaN-free canaries around both state
    buffers and the per-tile work arrays must survive a time loop (lazy steps, a materialisation, more steps)."""
    wl = W.stiff_reaction(P=2, J=J, seed=J)
    m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=2,
                                           reaction=wl["reaction"], config=config, blocks=blocks)
    m.set_boundary_conditions(**wl["bc"])
    m.prepare(0.5)
    pad, canary = 4096, -777.25
    guarded = {}

    def guard(t):
        big = torch.full((t.numel() + 2 * pad,), canary, dtype=torch.float64, device=t.device)
        view = big[pad:pad + t.numel()].view(t.shape)
        view.copy_(t)
        guarded[id(view)] = big
        return view

    m._wbuf = [guard(b) for b in m._wbuf]
    m.rs, m.rr, m.g_first = guard(m.rs), guard(m.rr), guard(m.g_first)
    if m.hier:
        m.xs_ext, m._gr, m.packet = guard(m.xs_ext), guard(m._gr), guard(m.packet)
    else:
        m.xs = guard(m.xs)
    m._plan = m._make_plan()
    m.set_state(wl["w0"])
    m.step(5)
    mid = m.get_state()
    m.step(4)
    out = m.get_state()
    assert bool(torch.isfinite(mid).all()) and bool(torch.isfinite(out).all())
    for big in guarded.values():
        assert bool((big[:pad] == canary).all()) and bool((big[-pad:] == canary).all())


@pytest.mark.parametrize("blocks", [1, 3])
def test_cuda_graph_time_loop_is_the_same_loop(blocks):
    """Launch-bound sizes replay two captured steps as a CUDA graph (fvm_spike_graph_*): same bits as the plain
    one-call loop, from every starting condition (fresh state, pending state, odd and even buffer)."""
    wl = W.stiff_reaction(P=2, J=12 * 256 - 7, seed=9)
    res = []
    for use_graph in (True, False):
        m = PartitionedAdvectionDiffusionModel(wl["faces"], wl["a"], wl["d"], wl["k"], "exponential", P=2,
                                               reaction=wl["reaction"], config=1, blocks=blocks)
        m.use_graph = use_graph
        m.set_boundary_conditions(**wl["bc"])
        m.prepare(0.5)
        m.set_state(wl["w0"])
        out = []
        for n in (37, 16, 3, 21):          # fresh -> graph; pending in buffer 1 -> graph; no graph; pending again
            m.step(n)
            out.append(m.get_state().clone())
        m.set_state(out[0])                 # a true state again, graph kept
        m.step(40)
        out.append(m.get_state().clone())
        assert (m._graph is not None) == use_graph
        res.append(out)
    for a, b in zip(*res):
        assert torch.equal(a, b)
