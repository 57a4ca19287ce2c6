"""Long systems: one huge mesh (or a batch of long ones) partitioned into tiles, optionally
with the cells split in contiguous slabs over the ranks of a torch.distributed group
(BASELINE.json cfg 4 and cfg 5).  Drives variant 3 of the C-ABI
(fvm_spike_*): tile-local Thomas solves, a reduced tridiagonal over the tile separators (itself
partitioned into blocks, ONE all-gather of 4 doubles per block and step across ranks), and a spike correction
near the tile edges.  A step has no right-hand side (w' = A^-1(w/theta + beta) - c w), so a tile
needs nothing from its neighbours and the state carries no ghost cells.

Semantics are those of the reference's loop (fvm_classes.py:373-377,408-411).
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from .batched import _alpha_ends, _bc_kinds

SEP_LO, SEP_UP, SEP_RL, SEP_RDA, SEP_V0, SEP_W0, SEP_STRIDE = range(7)


def plan_partition(J, tile_cells, world):
    """How a J-cell mesh is cut: global tiles (padded so that every rank owns the same number),
    tiles per rank, and for each rank the global cell range [j0, j1) it really holds."""
    J, T, world = int(J), int(tile_cells), int(world)
    n_tiles = -(-J // T)
    per_rank = -(-n_tiles // world)
    ranges = []
    for r in range(world):
        j0 = r * per_rank * T
        ranges.append((min(j0, J), min(j0 + per_rank * T, J)))
    return dict(tiles_per_rank=per_rank, tiles_global=per_rank * world, ranges=ranges)


def pick_blocks(nt, world, wanted=None):
    """Blocks per rank of the separator system (nt unknowns per rank): as many as keep a block at >= 256
    unknowns (one pass of the CTA-wide solver), divide the rank's tiles evenly and keep the top system
    (world * blocks unknowns) within the 1,024 one CTA solves.  ``wanted``: the caller's choice, checked."""
    nt, cap = int(nt), 1024 // int(world)
    if wanted is not None:
        B = int(wanted)
        if B < 1 or nt % B or nt // B < 2 or B > cap:
            raise ValueError("blocks=%d does not cut %d tiles per rank into equal blocks of >= 2 (at most %d)"
                             % (B, nt, cap))
        return B
    best = 1
    for B in range(2, min(nt // 256, cap) + 1):
        if nt % B == 0:
            best = B
    return best


class _NcclUniqueId(ctypes.Structure):
    _fields_ = [("internal", ctypes.c_byte * 128)]


class OwnNccl(object):
    """An NCCL communicator of our own over the ranks of a torch.distributed NCCL group, made with the NCCL
    library the process already has loaded.  fvm_spike_steps enqueues the per-step all-gather itself and
    needs the address of ncclAllGather and an ncclComm_t for that; torch does not hand out its own."""

    def __init__(self, group, rank, world, device):
        import torch
        import torch.distributed as dist
        self.lib = ctypes.CDLL("libnccl.so.2")
        uid = _NcclUniqueId()
        if rank == 0:
            self._check(self.lib.ncclGetUniqueId(ctypes.byref(uid)), "ncclGetUniqueId")
        t = torch.tensor(list(bytes(uid)), dtype=torch.uint8, device=device)
        dist.broadcast(t, src=dist.get_global_rank(group, 0), group=group)
        ctypes.memmove(ctypes.byref(uid), bytes(t.cpu().tolist()), 128)
        self.comm = ctypes.c_void_p()
        self.lib.ncclCommInitRank.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, _NcclUniqueId, ctypes.c_int]
        with torch.cuda.device(device):
            self._check(self.lib.ncclCommInitRank(ctypes.byref(self.comm), world, uid, rank), "ncclCommInitRank")
        self.allgather = ctypes.cast(self.lib.ncclAllGather, ctypes.c_void_p)

    @staticmethod
    def _check(rc, what):
        if rc != 0:
            raise RuntimeError("%s failed with ncclResult %d" % (what, rc))


class Comm(object):
    """The one collective the data path needs: all-gather of a small tensor over the ranks.
    world == 1: no-op.  NCCL groups gather on the device; a gloo group (tests: several ranks
    sharing one GPU) stages through host memory."""

    def __init__(self, group=None):
        self.group = group
        self.timing = None          # set to a list to collect (start, end) CUDA events of every gather
        self.own = None             # OwnNccl, made on first use (NCCL groups only)
        if group is None:
            self.rank, self.world, self.backend = 0, 1, None
        else:
            import torch.distributed as dist
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
            self.backend = dist.get_backend(group)

    def all_gather(self, t, out=None):
        """t: tensor of identical shape on every rank -> tensor [world, *t.shape] (``out`` if given: NCCL only)."""
        import torch
        if self.world == 1:
            return t.unsqueeze(0)
        import torch.distributed as dist
        if self.backend == "gloo":
            host = t.detach().cpu().contiguous()
            parts = [torch.empty_like(host) for _ in range(self.world)]
            dist.all_gather(parts, host, group=self.group)
            return torch.stack(parts).to(t.device)
        t = t.contiguous()
        if out is None:
            out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        if self.timing is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.all_gather_into_tensor(out, t, group=self.group)
            e1.record()
            self.timing.append((e0, e1))
            return out
        dist.all_gather_into_tensor(out, t, group=self.group)
        return out

    def own_nccl(self, device):
        """Our own communicator over the group's ranks (None unless the group's backend is NCCL)."""
        if self.backend != "nccl":
            return None
        if self.own is None:
            import torch
            import torch.distributed as dist
            try:
                own = OwnNccl(self.group, self.rank, self.world, device)
            except (OSError, RuntimeError, AttributeError):
                own = None
            # every rank or none: a rank that could not make its communicator sends all of them to the host loop
            ok = torch.tensor([1 if own is not None else 0], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            self.own = own if int(ok.item()) == 1 else False
        return self.own or None

    def all_reduce_max_int(self, values):
        import torch
        if self.world == 1:
            return [int(v) for v in values]
        import torch.distributed as dist
        t = torch.tensor([int(v) for v in values], dtype=torch.int64)
        if self.backend == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return [int(v) for v in t.cpu()]


class PartitionedAdvectionDiffusionModel(object):
    """P long problems on one mesh of ``J`` cells, this rank holding the cells
    ``plan_partition(J, T, world)['ranges'][rank]``.

    Two ways to pass the mesh and the coefficients:
      * global (default): ``faces`` has J+1 entries, per-cell ``a``/``d`` have J entries
        ((J,) or (P, J)); every rank slices what it needs.  Fine while J fits host memory.
      * local (``J=...`` given): ``faces`` covers the global faces
        [max(j0-1, 0), min(j1+1, J)] of this rank's range [j0, j1), per-cell coefficients
        cover the cells [max(j0-1, 0), min(j1, J-1)].
    ``a``, ``d``, ``reaction`` may also be scalars or (P,) per-problem scalars; ``k`` (the time
    step) is a scalar or (P,).  ``config`` picks the tile: 2 (default) = 4,096 cells, 256 threads, two CTAs
    per SM, so one tile's scans overlap the other's loads; 0 = 8,192 cells, 512 threads; 1 = 256 cells (tests).
    """

    def __init__(self, faces, a, d, k, discretisation="central", P=1, reaction=None, strict_bc=False,
                 J=None, group=None, config=2, spike_tol=1e-30, device=None, hierarchical=True, lazy=True,
                 blocks=None):
        torch = _lib.require_cuda()
        lib = _lib.lib()
        if discretisation not in _lib.SCHEMES:
            raise AttributeError("discretisation must be one of `upwind`, `central` or `exponential`")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.comm = Comm(group)
        self.config = int(config)
        self.T = int(lib.fvm_spike_tile_cells(self.config))
        if self.T == 0:
            raise ValueError("unknown tile configuration %r" % (config,))
        self.TPB = int(lib.fvm_spike_tile_threads(self.config))
        self.CPT = self.T // self.TPB
        self.local_input = J is not None
        self.J = int(J) if J is not None else int(np.shape(faces)[-1]) - 1
        if self.J < 3:
            raise IndexError("at least 3 cells are needed (the reference fails at fvm_classes.py:292)")
        self.P = int(P)
        self.discretisation, self.strict_bc, self.spike_tol = discretisation, strict_bc, float(spike_tol)
        self.hierarchical = bool(hierarchical)
        self._blocks_wanted = blocks    # blocks per rank of the separator system (None: chosen from its size)
        self.use_graph = True           # replay launch-bound time loops as a CUDA graph (see step)
        self._graph, self._gstream = None, None
        self.lazy = bool(lazy)          # apply the spike correction inside the next step's tile solve
        self._pending = False           # the current buffer holds uncorrected tile solutions
        plan = plan_partition(self.J, self.T, self.comm.world)
        self.nt = plan["tiles_per_rank"]
        self.nt_global = plan["tiles_global"]
        self.tile_offset = self.comm.rank * self.nt
        self.j0, self.j1 = plan["ranges"][self.comm.rank]
        self.n_local = self.nt * self.T                      # padded slab length
        # global index of the first face / cell this rank holds in its local arrays
        self.f_lo = max(self.j0 - 1, 0) if self.j1 > self.j0 else min(self.j0, self.J)
        self.c_lo = self.f_lo
        f_hi = min(self.j1 + 1, self.J)
        c_hi = min(self.j1, self.J - 1)
        self.faces = self._slab(faces, self.f_lo, f_hi + 1, self.J + 1)
        self.a, self._a_str = self._coef(a, c_hi)
        self.d, self._d_str = self._coef(d, c_hi)
        self.reaction, self._r_str = (None, (0, 0)) if reaction is None else self._coef(reaction, c_hi)
        kk = self._dev(k).reshape(-1)
        self.k = (kk.expand(self.P) if kk.numel() == 1 else kk).contiguous()
        self._ready = None

    # ------------------------------------------------------------------ inputs
    def _dev(self, x):
        torch = _lib.require_cuda()
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=torch.float64)
        return torch.as_tensor(np.asarray(x, dtype=np.float64), device=self.device)

    def _slab(self, x, lo, hi, n_global):
        """Entries [lo, hi) of a global per-face/per-cell array (last axis), or the local array as is."""
        t = self._dev(x)
        if self.local_input:
            if t.shape[-1] != max(hi - lo, 0):
                raise ValueError("local array has %d entries, expected %d" % (t.shape[-1], hi - lo))
            return t.contiguous()
        if t.shape[-1] != n_global:
            raise ValueError("global array has %d entries, expected %d" % (t.shape[-1], n_global))
        return t[..., lo:hi].contiguous()

    def _coef(self, x, c_hi):
        """Tensor, (problem stride, cell stride) and the global cell its element 0 refers to."""
        t = self._dev(x)
        if t.ndim == 0:
            return t.reshape(1).contiguous(), (0, 0)
        if t.ndim == 1 and t.numel() == self.P and (self.local_input or t.numel() != self.J):
            return t.contiguous(), (1, 0)
        s = self._slab(t, self.c_lo, c_hi + 1, self.J)
        if s.ndim == 1:
            return s, (0, 1)
        if s.ndim == 2 and s.shape[0] == self.P:
            return s, (s.shape[1], 1)
        raise ValueError("coefficient must be scalar, (P,), per-cell or (P, per-cell)")

    def check_mesh(self):
        """The duplicate-face check of Mesh.__init__ (fvm_classes.py:36-37) on the device, for this
        rank's slab: raises ValueError on zero-width cells; returns the number of negative-width
        cells (unsorted faces, which the reference accepts silently)."""
        torch = _lib.require_cuda()
        counts = torch.zeros(2, dtype=torch.int64, device=self.device)
        n = int(self.faces.numel()) - 1
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_mesh_check(_lib.ptr(self.faces), n, ctypes.c_void_p(counts.data_ptr()),
                                           _lib.stream_ptr())
        _lib.check(st, "fvm_mesh_check")
        zero, neg = (int(v) for v in counts.cpu())
        if zero:
            raise ValueError("The faces array contains duplicated positions (%d zero-width cells). No cell can "
                             "have zero volume so please update with unique face positions." % zero)
        return neg

    def set_boundary_conditions(self, left_flux=None, right_flux=None, left_value=None, right_value=None):
        """Same keywords as the reference (fvm_classes.py:141-147); fluxes may be per problem."""
        self.left_flux, self.right_flux = left_flux, right_flux
        self.left_value, self.right_value = left_value, right_value
        self._ready = None

    def _flux(self, x):
        t = self._dev(x).reshape(-1)
        return (t.expand(self.P) if t.numel() == 1 else t).contiguous()

    def _problem(self):
        if not hasattr(self, "left_flux"):
            raise AttributeError("boundary conditions are not set: call set_boundary_conditions first")
        lk, rk = _bc_kinds(self.left_flux, self.right_flux, self.left_value, self.right_value)
        al, ar = _alpha_ends(self.left_flux, lk, rk, self.strict_bc)
        self._lf = self._flux(self.left_flux) if lk == _lib.BC_ROBIN else None
        self._rf = self._flux(self.right_flux) if rk == _lib.BC_ROBIN else None
        pr = _lib.FvmProblem()
        pr.J, pr.P = self.J, self.P
        # the kernels index by GLOBAL cell: shift the base pointers by what this rank does not hold
        pr.faces = self.faces.data_ptr() - 8 * self.f_lo
        pr.a = self.a.data_ptr() - 8 * self.c_lo * self._a_str[1]
        pr.a_ps, pr.a_cs = self._a_str
        pr.d = self.d.data_ptr() - 8 * self.c_lo * self._d_str[1]
        pr.d_ps, pr.d_cs = self._d_str
        if self.reaction is not None:
            pr.reaction = self.reaction.data_ptr() - 8 * self.c_lo * self._r_str[1]
            pr.r_ps, pr.r_cs = self._r_str
        pr.scheme = _lib.SCHEMES[self.discretisation]
        pr.left_kind, pr.right_kind, pr.alpha_left, pr.alpha_right = lk, rk, al, ar
        self._has_lf = self._lf is not None and bool((self._lf != 0).any())
        self._has_rf = self._rf is not None and bool((self._rf != 0).any())
        pr.left_flux = self._lf.data_ptr() if self._has_lf else None
        pr.right_flux = self._rf.data_ptr() if self._has_rf else None
        return pr

    # ------------------------------------------------------------------ set-up
    def prepare(self, theta):
        """Assembly, tile factorisation, spikes and the factored reduced system for this theta."""
        torch = _lib.require_cuda()
        theta = float(theta)
        if self._ready == theta:
            return
        lib, dev, P, nt, T = _lib.lib(), self.device, self.P, self.nt, self.T
        pr = self._problem()
        new = lambda *shape: torch.empty(shape, dtype=torch.float64, device=dev)
        self.p, self.q, self.b = new(P, self.n_local), new(P, self.n_local), new(P, self.n_local)
        self.Vt, self.Wt = new(P, self.n_local), new(P, self.n_local)
        self.sep = new(P, nt, SEP_STRIDE)
        self.bends = new(P, 2)
        ext = torch.zeros(2, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            st = lib.fvm_spike_setup(ctypes.byref(pr), _lib.ptr(self.k), theta, self.tile_offset, nt, self.config,
                                     _lib.ptr(self.p), _lib.ptr(self.q), _lib.ptr(self.b), _lib.ptr(self.Vt),
                                     _lib.ptr(self.Wt), _lib.ptr(self.sep), _lib.ptr(self.bends),
                                     ctypes.c_void_p(ext.data_ptr()), self.spike_tol, _lib.stream_ptr())
        _lib.check(st, "fvm_spike_setup")
        # reduced tridiagonal over the separators of every rank
        sep_all = self.comm.all_gather(self.sep).permute(1, 0, 2, 3).reshape(P, self.nt_global, SEP_STRIDE)
        self.sup_all = sep_all[..., SEP_UP].contiguous()
        self.V0_all = sep_all[..., SEP_V0].contiguous()
        self.W0_all = sep_all[..., SEP_W0].contiguous()
        nxt = lambda x: torch.cat([x[:, 1:], torch.zeros_like(x[:, :1])], dim=1)
        RL = sep_all[..., SEP_RL].contiguous()
        RD = (sep_all[..., SEP_RDA] - self.sup_all * nxt(self.V0_all)).contiguous()
        RU = (-self.sup_all * nxt(self.W0_all)).contiguous()
        self.B = self._pick_blocks() if (self.hierarchical and nt >= 2) else 1
        self.hier = self.hierarchical and nt >= 2 and (self.comm.world > 1 or self.B > 1)
        if not self.hier:
            # every rank factors (and later solves) the whole separator system
            self.rp, self.rq, self.rinv = new(P, self.nt_global), new(P, self.nt_global), new(P, self.nt_global)
            with torch.cuda.device(dev):
                st = lib.fvm_spike_reduced_factor(self.nt_global, P, _lib.ptr(RL), _lib.ptr(RD), _lib.ptr(RU),
                                                  _lib.ptr(self.rp), _lib.ptr(self.rq), _lib.ptr(self.rinv),
                                                  _lib.stream_ptr())
            _lib.check(st, "fvm_spike_reduced_factor")
        else:
            self._prepare_block_level(RL, RD, RU)
        e = self.comm.all_reduce_max_int(ext.cpu().tolist())
        self.extents = (ctypes.c_int32 * 2)(min(e[0], T - 1), min(e[1], T - 1))
        # work arrays of a step
        self.rs, self.rr, self.g_first = new(P, nt, 2), new(P, nt), new(P)
        self.xs = new(P, self.nt_global)
        self.w_stride = self.n_local
        self._wbuf = [torch.zeros((P, self.w_stride), dtype=torch.float64, device=dev) for _ in range(2)]
        jl = self.J - 1 - self.j0
        self._last_local = jl if (self._has_rf and 0 <= jl < self.n_local) else -1
        self._theta = theta
        self._drop_graph()                  # a graph of an earlier prepare() points at arrays that are gone
        self._plan = self._make_plan()
        self._ready = theta

    def _graph_worthwhile(self, n):
        """A CUDA graph of the step pays where the step is launch-bound: one GPU, lazy correction, a slab of
        at most 2^22 cells in all, and enough steps to amortise the capture."""
        return (self.use_graph and self.comm.world == 1 and self.lazy and n >= 16
                and self.P * self.n_local <= (1 << 22))

    def _drop_graph(self):
        if getattr(self, "_graph", None) is not None:
            _lib.lib().fvm_spike_graph_destroy(self._graph)
        self._graph = None

    def __del__(self):
        try:
            self._drop_graph()
        except Exception:
            pass

    def _make_plan(self):
        """struct fvm_spike_plan: everything fvm_spike_steps needs to run whole steps in one host call."""
        pl = _lib.FvmSpikePlan()
        pl.n_tiles, pl.P, pl.config, pl.lazy = self.nt, self.P, self.config, int(self.lazy)
        pl.has_first_cell, pl.world, pl.rank = int(self.comm.rank == 0), self.comm.world, self.comm.rank
        pl.extents[0], pl.extents[1] = self.extents[0], self.extents[1]
        pl.last_cell_local, pl.theta, pl.w_stride = self._last_local, self._theta, self.w_stride
        d = lambda t: t.data_ptr()
        pl.p, pl.q, pl.b, pl.Vt, pl.Wt, pl.sep, pl.bends = (d(self.p), d(self.q), d(self.b), d(self.Vt), d(self.Wt),
                                                            d(self.sep), d(self.bends))
        pl.rp, pl.rq, pl.rinv = d(self.rp), d(self.rq), d(self.rinv)
        pl.rs, pl.rr, pl.g_first = d(self.rs), d(self.rr), d(self.g_first)
        if self.hier:
            pl.B, pl.m = self.B, self.nt // self.B
            pl.Vr, pl.Wr = d(self.Vr), d(self.Wr)
            pl.tp, pl.tq, pl.tinv, pl.cst = d(self.tp), d(self.tq), d(self.tinv), d(self.top_cst)
            pl.gr, pl.packet, pl.allp, pl.xs = d(self._gr), d(self.packet), d(self._allp), d(self.xs_ext)
        elif self.comm.world == 1:
            pl.B, pl.m, pl.xs = 1, self.nt, d(self.xs)
        else:
            return None                 # hierarchical=False across ranks: the Python loop below
        return pl

    def _pick_blocks(self):
        return pick_blocks(self.nt, self.comm.world, self._blocks_wanted)

    def _reduced_local_solve(self, rhs, out=None):
        """Solve this rank's blocks of the separator system (m unknowns each, couplings to the neighbouring
        blocks cut, last unknown = the block's separator, decoupled): rhs (P, nt) -> (P, nt)."""
        torch = _lib.require_cuda()
        out = torch.empty_like(rhs) if out is None else out
        m = self.nt // self.B
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_spike_reduce_solve(m, self.P * self.B, _lib.ptr(self.rp), _lib.ptr(self.rq),
                                                   _lib.ptr(self.rinv), _lib.ptr(rhs), _lib.ptr(out),
                                                   _lib.stream_ptr())
        _lib.check(st, "fvm_spike_reduce_solve")
        return out

    def _prepare_block_level(self, RL, RD, RU):
        """SPIKE once more, one level up.  The separator system (one unknown per tile, all ranks) is
        itself partitioned into B blocks per rank of m = nt/B unknowns: the LAST tile separator of a block
        is the block's separator, the m-1 others are solved block by block in parallel (one CTA each,
        instead of one CTA for the whole system); the block separators satisfy a (world*B)-unknown
        tridiagonal, factored here once, that the last kernel of a step solves.  Across GPUs a step then
        all-gathers 4 numbers per problem and block instead of the rank's whole right-hand side."""
        torch = _lib.require_cuda()
        lib, dev, P, nt, r, W, B = _lib.lib(), self.device, self.P, self.nt, self.comm.rank, self.comm.world, self.B
        m, NB = nt // B, self.comm.world * self.B
        sl = slice(r * nt, (r + 1) * nt)
        rl, rd, ru = (x[:, sl].reshape(P, B, m).contiguous().clone() for x in (RL, RD, RU))
        cut_lo, cut_up = rl[:, :, 0].clone(), ru[:, :, m - 2].clone()   # couplings leaving the block interiors
        sep_l, sep_d, sep_u = rl[:, :, m - 1].clone(), rd[:, :, m - 1].clone(), ru[:, :, m - 1].clone()
        rl[:, :, 0] = 0.0
        ru[:, :, m - 2] = 0.0
        rl[:, :, m - 1], rd[:, :, m - 1], ru[:, :, m - 1] = 0.0, 1.0, 0.0   # the separator rows: decoupled
        new = lambda *shape: torch.empty(shape, dtype=torch.float64, device=dev)
        self.rp, self.rq, self.rinv = new(P, nt), new(P, nt), new(P, nt)
        with torch.cuda.device(dev):
            st = lib.fvm_spike_reduced_factor(m, P * B, _lib.ptr(rl), _lib.ptr(rd), _lib.ptr(ru), _lib.ptr(self.rp),
                                              _lib.ptr(self.rq), _lib.ptr(self.rinv), _lib.stream_ptr())
        _lib.check(st, "fvm_spike_reduced_factor")
        e = torch.zeros((P, B, m), dtype=torch.float64, device=dev)
        e[:, :, 0] = cut_lo
        self.Vr = self._reduced_local_solve(e.view(P, nt)).view(P, B, m)   # response to the previous block's separator
        e.zero_()
        e[:, :, m - 2] = cut_up
        self.Wr = self._reduced_local_solve(e.view(P, nt)).view(P, B, m)   # response to the block's own separator
        self.Vr[:, :, m - 1] = 0.0
        self.Wr[:, :, m - 1] = 0.0
        info = torch.stack([sep_l, sep_d, sep_u, self.Vr[:, :, 0], self.Wr[:, :, 0], self.Vr[:, :, m - 2],
                            self.Wr[:, :, m - 2]], dim=2)                 # (P, B, 7)
        allv = self.comm.all_gather(info).permute(1, 0, 2, 3).reshape(P, NB, 7)
        sl_, sd_, su_, vf, wf, vl, wlast = [allv[:, :, i].contiguous() for i in range(7)]   # (P, NB)
        nxt = lambda x: torch.cat([x[:, 1:], torch.zeros_like(x[:, :1])], dim=1)
        TL = (-sl_ * vl).contiguous()
        TD = (sd_ - sl_ * wlast - su_ * nxt(vf)).contiguous()
        TU = (-su_ * nxt(wf)).contiguous()
        self.tp, self.tq, self.tinv = new(P, NB), new(P, NB), new(P, NB)
        with torch.cuda.device(dev):
            st = lib.fvm_spike_reduced_factor(NB, P, _lib.ptr(TL), _lib.ptr(TD), _lib.ptr(TU), _lib.ptr(self.tp),
                                              _lib.ptr(self.tq), _lib.ptr(self.tinv), _lib.stream_ptr())
        _lib.check(st, "fvm_spike_reduced_factor")
        # constants of fvm_spike_rank_top, (P, 3, NB): coupling of every rank's last separator to the next
        # rank's first tile value (0 for the blocks inside a rank: reduced_rhs_kernel has that term); lo, up of
        # the block separators' rows
        c0 = torch.zeros((P, NB), dtype=torch.float64, device=dev)
        last = torch.arange(1, W + 1, device=dev) * nt - 1              # every rank's last tile
        c0[:, B - 1::B] = self.sup_all[:, last]
        self.top_cst = torch.stack([c0, sl_, su_], dim=1).contiguous()
        self.Vr, self.Wr = self.Vr.view(P, nt), self.Wr.view(P, nt)
        self.xs_ext = torch.zeros((P, nt + 1), dtype=torch.float64, device=dev)
        self.packet = torch.zeros((P, B, 4), dtype=torch.float64, device=dev)
        self._gr = torch.zeros((P, nt), dtype=torch.float64, device=dev)          # work arrays of a step: nothing is
        self._allp = torch.zeros((W, P, B, 4), dtype=torch.float64, device=dev)   # allocated inside the time loop

    # ------------------------------------------------------------------ state
    def set_state(self, w):
        """w: (P, J) global state (global input mode) or (P, j1-j0) this rank's cells."""
        torch = _lib.require_cuda()
        w = self._dev(w).reshape(self.P, -1)
        n_real = self.j1 - self.j0
        if w.shape[1] == self.J and not (self.local_input and self.comm.world > 1):
            w = w[:, self.j0:self.j1]
        if w.shape[1] != n_real:
            raise ValueError("state has %d cells per problem, expected %d" % (w.shape[1], n_real))
        self._pending = False
        buf = self._wbuf[0]
        buf.zero_()
        nat = torch.zeros((self.P, self.n_local), dtype=torch.float64, device=self.device)
        nat[:, :n_real] = w
        # natural -> tiled order inside every tile (include/fvm_b200.h: fvm_spike_tile_threads)
        buf[:, :] = nat.view(self.P, self.nt, self.TPB, self.CPT).transpose(2, 3).reshape(self.P, -1)
        self._cur = 0

    def get_state(self):
        """This rank's cells, (P, j1-j0)."""
        self._materialise()
        til = self._wbuf[self._cur].view(self.P, self.nt, self.CPT, self.TPB)
        return til.transpose(2, 3).reshape(self.P, -1)[:, :self.j1 - self.j0].contiguous()

    def _xs_index(self):
        """(tensor, tile_offset, n_tiles_global) under which the separator values are addressed."""
        if self.hier:
            return self.xs_ext, 1, self.nt + 1
        return self.xs, self.tile_offset, self.nt_global

    def _materialise(self):
        """Turn the current buffer into the true state if a lazy correction is pending."""
        if not self._pending:
            return
        torch = _lib.require_cuda()
        xs, off, ng = self._xs_index()
        buf = self._wbuf[self._cur]
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_spike_finalize(self.nt, self.P, self.config, self.extents, off, ng, _lib.ptr(self.Vt),
                                               _lib.ptr(self.Wt), _lib.ptr(xs), _lib.ptr(buf), self.w_stride,
                                               _lib.stream_ptr())
        _lib.check(st, "fvm_spike_finalize")
        self._pending = False

    def _local(self, w_in, w_out):
        """Step, part 1 (tile solves + reduced right-hand side), with the pending correction applied on read."""
        lib, P, nt, cfg, comm = _lib.lib(), self.P, self.nt, self.config, self.comm
        common = (nt, P, cfg, _lib.ptr(self.p), _lib.ptr(self.q), _lib.ptr(self.b), _lib.ptr(self.sep),
                  _lib.ptr(self.bends), int(comm.rank == 0), self._last_local, self._theta, _lib.ptr(w_in), _lib.ptr(w_out),
                  self.w_stride, _lib.ptr(self.rs), _lib.ptr(self.rr), nt, _lib.ptr(self.g_first))
        if self._pending:
            xs, off, ng = self._xs_index()
            st = lib.fvm_spike_local_lazy(*common, _lib.ptr(self.Vt), _lib.ptr(self.Wt), _lib.ptr(xs), off, ng,
                                          self.extents, _lib.stream_ptr())
            _lib.check(st, "fvm_spike_local_lazy")
        else:
            st = lib.fvm_spike_local(*common, _lib.stream_ptr())
            _lib.check(st, "fvm_spike_local")

    def _finalize(self, w_out, xs, off, ng):
        """Step, part 3: nothing when the correction is lazy (the next tile solve applies it), else the correction."""
        if self.lazy:
            self._pending = True
            return
        st = _lib.lib().fvm_spike_finalize(self.nt, self.P, self.config, self.extents, off, ng, _lib.ptr(self.Vt),
                                           _lib.ptr(self.Wt), _lib.ptr(xs), _lib.ptr(w_out), self.w_stride,
                                           _lib.stream_ptr())
        _lib.check(st, "fvm_spike_finalize")

    # ------------------------------------------------------------------ stepping
    def step(self, nsteps=1):
        """``nsteps`` theta steps of the state set by :meth:`set_state` (theta of :meth:`prepare`)."""
        torch = _lib.require_cuda()
        lib, P, nt, cfg, comm = _lib.lib(), self.P, self.nt, self.config, self.comm
        own = comm.own_nccl(self.device) if comm.world > 1 else None
        if self._plan is not None and comm.timing is None and (comm.world == 1 or own is not None):
            # the whole loop in one host call; the all-gather, if any, through our own NCCL communicator
            def steps(n, stream):
                if n <= 0:
                    return
                st = lib.fvm_spike_steps(ctypes.byref(self._plan), _lib.ptr(self._wbuf[0]), _lib.ptr(self._wbuf[1]),
                                         self._cur, int(self._pending), int(n),
                                         own.allgather if own else None, own.comm if own else None, stream)
                _lib.check(st, "fvm_spike_steps")
                self._cur = (self._cur + int(n)) % 2
                self._pending = self.lazy

            n = int(nsteps)
            with torch.cuda.device(self.device):
                if self._graph_worthwhile(n):
                    # launch-bound sizes: two steps captured once in a CUDA graph and replayed, on a stream of
                    # our own (the legacy default stream cannot capture)
                    cur_stream = torch.cuda.current_stream()
                    if self._gstream is None:
                        self._gstream = torch.cuda.Stream(device=self.device)
                    self._gstream.wait_stream(cur_stream)
                    with torch.cuda.stream(self._gstream):
                        gs = _lib.stream_ptr()
                        lead = 1 if self._cur == 1 else (0 if self._pending else 2)   # -> pending state in buffer 0
                        steps(lead, gs)
                        if self._graph is None:
                            h = ctypes.c_void_p()
                            st = lib.fvm_spike_graph_create(ctypes.byref(self._plan), _lib.ptr(self._wbuf[0]),
                                                            _lib.ptr(self._wbuf[1]), gs, ctypes.byref(h))
                            _lib.check(st, "fvm_spike_graph_create")
                            self._graph = h
                        pairs = (n - lead) // 2
                        _lib.check(lib.fvm_spike_graph_run(self._graph, pairs, gs), "fvm_spike_graph_run")
                        steps(n - lead - 2 * pairs, gs)
                    cur_stream.wait_stream(self._gstream)
                else:
                    steps(n, _lib.stream_ptr())
            return
        with torch.cuda.device(self.device):
            for _ in range(int(nsteps)):
                w_in, w_out = self._wbuf[self._cur], self._wbuf[1 - self._cur]
                self._local(w_in, w_out)
                if self.hier:
                    self._step_rank_level(w_out)
                    self._cur = 1 - self._cur
                    continue
                if comm.world == 1:
                    rr_all = self.rr
                else:
                    packet = torch.cat([self.rr, self.g_first[:, None]], dim=1)            # [P, nt+1]
                    allp = comm.all_gather(packet)                                          # [world, P, nt+1]
                    rr_all = allp[:, :, :nt].permute(1, 0, 2).reshape(P, self.nt_global).contiguous()
                    gf = allp[:, :, nt]                                                     # [world, P]
                    last = torch.arange(1, comm.world, device=self.device) * nt - 1        # last tile of ranks 0..W-2
                    rr_all[:, last] -= self.sup_all[:, last] * gf[1:].t()
                st = lib.fvm_spike_reduce_solve(self.nt_global, P, _lib.ptr(self.rp), _lib.ptr(self.rq),
                                                _lib.ptr(self.rinv), _lib.ptr(rr_all), _lib.ptr(self.xs),
                                                _lib.stream_ptr())
                _lib.check(st, "fvm_spike_reduce_solve")
                self._finalize(w_out, self.xs, self.tile_offset, self.nt_global)
                self._cur = 1 - self._cur

    def _step_rank_level(self, w_out):
        """Second half of a step with the block-level hierarchy (see _prepare_block_level): block solves of
        the separator system, ONE all-gather of 4 doubles per problem and block, top system, finalize."""
        lib, P, nt, comm, B = _lib.lib(), self.P, self.nt, self.comm, self.B
        gr = self._reduced_local_solve(self.rr, self._gr)   # block solutions; the separator rows pass their rhs through
        st = lib.fvm_spike_rank_pack(P, B, nt // B, _lib.ptr(gr), _lib.ptr(self.g_first), _lib.ptr(self.packet),
                                     _lib.stream_ptr())
        _lib.check(st, "fvm_spike_rank_pack")
        allp = comm.all_gather(self.packet, self._allp)                  # (W, P, B, 4), into a buffer kept across steps
        st = lib.fvm_spike_rank_top(P, B, nt // B, comm.world, comm.rank, _lib.ptr(allp), _lib.ptr(self.tp),
                                    _lib.ptr(self.tq), _lib.ptr(self.tinv), _lib.ptr(self.top_cst), _lib.ptr(gr),
                                    _lib.ptr(self.Vr), _lib.ptr(self.Wr), _lib.ptr(self.xs_ext), _lib.stream_ptr())
        _lib.check(st, "fvm_spike_rank_top")
        self._finalize(w_out, self.xs_ext, 1, nt + 1)

    def run(self, w, theta, nsteps):
        """set_state + prepare + step; returns this rank's cells of the final state."""
        self.prepare(theta)
        self.set_state(w)
        self.step(nsteps)
        return self.get_state()
