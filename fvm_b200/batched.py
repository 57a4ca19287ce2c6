"""Batches of independent advection-diffusion problems on one mesh, on the GPU.

This is the accelerated entry the reference's driver idiom maps to
(fvm_classes.py:367-377,408-411): construct, set boundary conditions, then
``run(w, theta, nsteps)`` instead of the ``spsolve`` loop.  State and coefficients
are torch CUDA float64 tensors; all arithmetic happens in libfvm_b200.so.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib


def _bc_kinds(left_flux, right_flux, left_value, right_value):
    """Row type of each end.  A value beats a flux when both are given
    (fvm_classes.py:277-287); an end with neither raises UnboundLocalError like the
    reference does at :304."""
    if left_value is not None:
        left = _lib.BC_DIRICHLET
    elif left_flux is not None:
        left = _lib.BC_ROBIN
    else:
        raise UnboundLocalError("left boundary condition is not set (fvm_classes.py:260,304)")
    if right_value is not None:
        right = _lib.BC_DIRICHLET
    elif right_flux is not None:
        right = _lib.BC_ROBIN
    else:
        raise UnboundLocalError("right boundary condition is not set (fvm_classes.py:260,304)")
    return left, right


def _alpha_ends(left_flux, left_kind, right_kind, strict_bc):
    """End entries of the alpha mask.  The reference keys BOTH ends off
    ``left_flux is None`` (fvm_classes.py:237-238); strict_bc keys each end off its own row."""
    if strict_bc:
        return int(left_kind == _lib.BC_ROBIN), int(right_kind == _lib.BC_ROBIN)
    a1 = 0 if left_flux is None else 1
    return a1, a1


class BatchedAdvectionDiffusionModel(object):
    """P problems on one mesh.

    faces : (J+1,) array-like.
    a, d  : scalar, (P,) per-problem scalars, (J,) per-cell shared, or (P, J).
    k     : time step (the reference's 4th positional argument, fvm_classes.py:92,351);
            scalar or (P,).
    reaction : optional decay rate, same shapes as a (not in the reference).
    """

    def __init__(self, faces, a, d, k, discretisation="central", P=None, reaction=None,
                 strict_bc=False, device=None):
        torch = _lib.require_cuda()
        _lib.lib()
        if discretisation not in _lib.SCHEMES:
            raise AttributeError("discretisation must be one of `upwind`, `central` or `exponential` "
                                 "(the reference fails on self.kappa, fvm_classes.py:128,131)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None \
            else torch.device(device)
        self.discretisation = discretisation
        self.strict_bc = strict_bc
        self.faces = self._dev(faces).reshape(-1)
        self.J = int(self.faces.numel()) - 1
        if self.J < 3:
            raise IndexError("at least 3 cells are needed (the reference fails at fvm_classes.py:292)")
        if P is None:                      # infer the batch size: k first, then 2-D, then 1-D != J
            shapes = [tuple(np.shape(x)) for x in (k, a, d)]
            P = 1
            for s in reversed(shapes[1:]):
                if len(s) == 1 and s[0] != self.J:
                    P = s[0]
            for s in reversed(shapes[1:]):
                if len(s) == 2:
                    P = s[0]
            if len(shapes[0]) == 1:
                P = shapes[0][0]
        self.P = int(P)
        self.a, self._a_str = self._cellvar(a)
        self.d, self._d_str = self._cellvar(d)
        self.reaction, self._r_str = (None, (0, 0)) if reaction is None else self._cellvar(reaction)
        kk = self._dev(k).reshape(-1)
        self.k = kk.expand(self.P).contiguous() if kk.numel() == 1 else kk.contiguous()
        if self.k.numel() != self.P:
            raise ValueError("k must be a scalar or have one entry per problem")
        self._asm = None
        self._factors = {}
        self._pr = None

    # ------------------------------------------------------------------ helpers
    def _dev(self, x):
        torch = _lib.require_cuda()
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=torch.float64)
        return torch.as_tensor(np.asarray(x, dtype=np.float64), device=self.device)

    def _cellvar(self, x):
        """Tensor and (problem stride, cell stride) of a coefficient."""
        t = self._dev(x)
        if t.ndim == 0:
            return t.reshape(1).contiguous(), (0, 0)
        if t.ndim == 1 and t.numel() == self.P:          # (when P == J a 1-D array is per problem)
            return t.contiguous(), (1, 0)
        if t.ndim == 1 and t.numel() == self.J:
            return t.contiguous(), (0, 1)
        if t.ndim == 2 and tuple(t.shape) == (self.P, self.J):
            return t.contiguous(), (self.J, 1)
        raise ValueError("coefficient must be scalar, (P,), (J,) or (P, J); got %s" % (tuple(t.shape),))

    def _flux(self, x):
        t = self._dev(x).reshape(-1)
        t = t.expand(self.P).contiguous() if t.numel() == 1 else t.contiguous()
        if t.numel() != self.P:
            raise ValueError("a boundary flux must be a scalar or have one entry per problem")
        return t

    def set_boundary_conditions(self, left_flux=None, right_flux=None, left_value=None, right_value=None):
        """Same keywords as the reference (fvm_classes.py:141-147).  Fluxes may be per
        problem; Dirichlet values are only flags, exactly as in the reference: a Dirichlet
        cell keeps whatever the initial state holds there."""
        self.left_flux, self.right_flux = left_flux, right_flux
        self.left_value, self.right_value = left_value, right_value
        self._asm = None
        self._factors = {}
        self._pr = None
        self._part_theta = None
        self._unit = None

    def _problem(self):
        if getattr(self, "_pr", None) is not None:
            return self._pr
        if not hasattr(self, "left_flux"):
            raise AttributeError("boundary conditions are not set: call set_boundary_conditions first")
        lk, rk = _bc_kinds(self.left_flux, self.right_flux, self.left_value, self.right_value)
        al, ar = _alpha_ends(self.left_flux, lk, rk, self.strict_bc)
        self._lf = self._flux(self.left_flux) if lk == _lib.BC_ROBIN else None
        self._rf = self._flux(self.right_flux) if rk == _lib.BC_ROBIN else None
        pr = _lib.FvmProblem()
        pr.J, pr.P = self.J, self.P
        pr.faces = self.faces.data_ptr()
        pr.a, (pr.a_ps, pr.a_cs) = self.a.data_ptr(), self._a_str
        pr.d, (pr.d_ps, pr.d_cs) = self.d.data_ptr(), self._d_str
        if self.reaction is not None:
            pr.reaction, (pr.r_ps, pr.r_cs) = self.reaction.data_ptr(), self._r_str
        pr.scheme = _lib.SCHEMES[self.discretisation]
        pr.left_kind, pr.right_kind = lk, rk
        pr.alpha_left, pr.alpha_right = al, ar
        pr.left_flux = self._lf.data_ptr() if self._lf is not None and bool((self._lf != 0).any()) else None
        pr.right_flux = self._rf.data_ptr() if self._rf is not None and bool((self._rf != 0).any()) else None
        self._pr = pr
        return pr

    # ------------------------------------------------------------------ assembly
    def assemble(self, interleaved=False):
        """kappa, modified D, the diagonals of M and the end entries of beta, from
        fvm_assemble.  Tensors are (P, J), or (J, P) when ``interleaved``."""
        torch = _lib.require_cuda()
        key = bool(interleaved)
        if self._asm is not None and key in self._asm:
            return self._asm[key]
        pr = self._problem()
        shape = (self.J, self.P) if interleaved else (self.P, self.J)
        ps, cs = (1, self.P) if interleaved else (self.J, 1)
        with torch.cuda.device(self.device):
            out = {n: torch.empty(shape, dtype=torch.float64, device=self.device)
                   for n in ("lo", "di", "up", "kappa", "D")}
            out["beta_left"] = torch.empty(self.P, dtype=torch.float64, device=self.device)
            out["beta_right"] = torch.empty(self.P, dtype=torch.float64, device=self.device)
            st = _lib.lib().fvm_assemble(ctypes.byref(pr), _lib.ptr(out["lo"]), _lib.ptr(out["di"]),
                                         _lib.ptr(out["up"]), _lib.ptr(out["kappa"]), _lib.ptr(out["D"]),
                                         ps, cs, _lib.ptr(out["beta_left"]), _lib.ptr(out["beta_right"]),
                                         _lib.stream_ptr())
        _lib.check(st, "fvm_assemble")
        out["alpha_left"], out["alpha_right"] = pr.alpha_left, pr.alpha_right
        if self._asm is None:
            self._asm = {}
        self._asm[key] = out
        return out

    def diagnostics(self, w=None):
        """Per problem, on the device: min/max Peclet number (a h / d with the unmodified d, what
        the reference's constructor warns about, fvm_classes.py:103-107), min/max CFL number (:110-114)
        and, when a state (P, J) is given, the conserved "area" sum_j w_j h_j (:420-421)."""
        torch = _lib.require_cuda()
        pr = self._problem() if hasattr(self, "left_flux") else self._problem_no_bc()
        out = torch.empty((self.P, 5), dtype=torch.float64, device=self.device)
        wt = None
        if w is not None:
            wt = self._dev(w).reshape(self.P, self.J).contiguous()
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_diagnostics(ctypes.byref(pr), _lib.ptr(self.k), _lib.ptr(wt), self.J, 1,
                                            _lib.ptr(out), _lib.stream_ptr())
        _lib.check(st, "fvm_diagnostics")
        res = dict(peclet_min=out[:, 0], peclet_max=out[:, 1], cfl_min=out[:, 2], cfl_max=out[:, 3])
        if w is not None:
            res["area"] = out[:, 4]
        return res

    def _problem_no_bc(self):
        """struct fvm_problem for calls that do not look at the boundary rows."""
        pr = _lib.FvmProblem()
        pr.J, pr.P = self.J, self.P
        pr.faces = self.faces.data_ptr()
        pr.a, (pr.a_ps, pr.a_cs) = self.a.data_ptr(), self._a_str
        pr.d, (pr.d_ps, pr.d_cs) = self.d.data_ptr(), self._d_str
        pr.scheme = _lib.SCHEMES[self.discretisation]
        return pr

    # ------------------------------------------------------------------ variant 1
    def factor(self, theta):
        """Factored theta operators in the interleaved layout (fvm_theta_factor)."""
        torch = _lib.require_cuda()
        theta = float(theta)
        if theta in self._factors:
            return self._factors[theta]
        asm = self.assemble(interleaved=True)
        with torch.cuda.device(self.device):
            f = {n: torch.empty((self.J, self.P), dtype=torch.float64, device=self.device)
                 for n in ("fp", "fq", "fb")}
            f["bt_left"] = torch.empty(self.P, dtype=torch.float64, device=self.device)
            f["bt_right"] = torch.empty(self.P, dtype=torch.float64, device=self.device)
            st = _lib.lib().fvm_theta_factor(
                self.J, self.P, _lib.ptr(asm["lo"]), _lib.ptr(asm["di"]), _lib.ptr(asm["up"]), 1, self.P,
                _lib.ptr(self.k), theta, asm["alpha_left"], asm["alpha_right"],
                _lib.ptr(asm["beta_left"]), _lib.ptr(asm["beta_right"]),
                _lib.ptr(f["fp"]), _lib.ptr(f["fq"]), _lib.ptr(f["fb"]), 1, self.P,
                _lib.ptr(f["bt_left"]), _lib.ptr(f["bt_right"]), _lib.stream_ptr())
        _lib.check(st, "fvm_theta_factor")
        self._factors[theta] = f
        return f

    def step_thomas_(self, w_t, theta, nsteps=1):
        """Variant 1 on an interleaved state w_t (J, P), in place."""
        torch = _lib.require_cuda()
        f = self.factor(theta)
        if tuple(w_t.shape) != (self.J, self.P) or not w_t.is_contiguous():
            raise ValueError("interleaved state must be a contiguous (J, P) tensor")
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_theta_step_thomas(
                self.J, self.P, _lib.ptr(f["fp"]), _lib.ptr(f["fq"]), _lib.ptr(f["fb"]), 1, self.P,
                float(theta), _lib.ptr(f["bt_left"]), _lib.ptr(f["bt_right"]), _lib.ptr(w_t), 1, self.P,
                int(nsteps), _lib.stream_ptr())
        _lib.check(st, "fvm_theta_step_thomas")
        return w_t

    # ------------------------------------------------------------------ variant 2
    def run_fused_(self, w, theta, nsteps, checkpoint_every=0, ckpt=None):
        """Variant 2 on a problem-major state w (P, J), in place; optional checkpoints
        tensor (nsteps // checkpoint_every, P, J)."""
        torch = _lib.require_cuda()
        if tuple(w.shape) != (self.P, self.J) or not w.is_contiguous():
            raise ValueError("state must be a contiguous (P, J) tensor")
        pr = self._problem()
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_theta_run_fused(ctypes.byref(pr), _lib.ptr(self.k), float(theta), int(nsteps),
                                                int(checkpoint_every), _lib.ptr(w), _lib.ptr(ckpt),
                                                _lib.stream_ptr())
        _lib.check(st, "fvm_theta_run_fused")
        return w

    # ------------------------------------------------------------------ public driver
    def run(self, w, theta, nsteps, checkpoint_every=0, variant="auto"):
        """Advance w (P, J) by ``nsteps`` theta steps; returns (w_final, checkpoints or None).
        Equivalent of the loop  d = (I + k*(1-theta)*alpha*M)*w + beta ; w = spsolve(A, d)
        (fvm_classes.py:408-411) for every problem of the batch.
        variant: 'fused' (on-chip, J <= fused_max_cells), 'thomas' (streaming, one thread per
        system), 'partitioned' (variant 3 on this GPU: few long systems), or 'auto' (fused when
        it fits, else thomas for batches of at least THOMAS_MIN_PROBLEMS, else partitioned)."""
        torch = _lib.require_cuda()
        w = self._dev(w).reshape(self.P, self.J).clone()
        nsteps = int(nsteps)
        ce = int(checkpoint_every or 0)
        n_ck = nsteps // ce if ce > 0 else 0
        ckpt = torch.empty((n_ck, self.P, self.J), dtype=torch.float64, device=self.device) if n_ck else None
        if float(theta) == 0.0:
            return self._run_explicit(w, nsteps, ce, ckpt), ckpt
        if variant == "auto":
            variant = self._auto_variant()
        if variant == "fused":
            self.run_fused_(w, theta, nsteps, ce, ckpt)
        elif variant == "partitioned":
            w = self._run_partitioned(w, theta, nsteps, ce, ckpt)
        elif variant == "thomas":
            w_t = w.t().contiguous()
            if n_ck:
                for i in range(n_ck):
                    self.step_thomas_(w_t, theta, ce)
                    ckpt[i].copy_(w_t.t())
                rest = nsteps - n_ck * ce
                if rest:
                    self.step_thomas_(w_t, theta, rest)
            else:
                self.step_thomas_(w_t, theta, nsteps)
            w = w_t.t().contiguous()
        else:
            raise ValueError("variant must be 'auto', 'fused', 'thomas' or 'partitioned'")
        return w, ckpt

    # a batch this wide fills the GPU with one thread per system; below it, long systems go to variant 3
    THOMAS_MIN_PROBLEMS = 16384

    def _auto_variant(self):
        if self.J <= int(_lib.lib().fvm_fused_max_cells()):
            return "fused"
        return "thomas" if self.P >= self.THOMAS_MIN_PROBLEMS else "partitioned"

    def _run_partitioned(self, w, theta, nsteps, ce, ckpt):
        """Variant 3 on this GPU alone (PartitionedAdvectionDiffusionModel with no process group)."""
        from .partitioned import PartitionedAdvectionDiffusionModel
        key = float(theta)
        m = getattr(self, "_part", None)
        if m is None:
            def glob(t, strides):          # back to the shape the constructor classifies
                if t is None:
                    return None
                return t.reshape(()) if strides == (0, 0) else t
            m = PartitionedAdvectionDiffusionModel(
                self.faces, glob(self.a, self._a_str), glob(self.d, self._d_str), self.k, self.discretisation,
                P=self.P, reaction=glob(self.reaction, self._r_str), strict_bc=self.strict_bc, device=self.device)
            self._part, self._part_theta = m, None
        if not hasattr(self, "left_flux"):
            raise AttributeError("boundary conditions are not set: call set_boundary_conditions first")
        if self._part_theta != key:
            m.set_boundary_conditions(self.left_flux, self.right_flux, self.left_value, self.right_value)
            m.prepare(theta)
            self._part_theta = key
        m.set_state(w)
        done = 0
        if ce > 0 and ckpt is not None:
            for i in range(ckpt.shape[0]):
                m.step(ce)
                ckpt[i].copy_(m.get_state())
                done += ce
        if nsteps - done:
            m.step(nsteps - done)
        return m.get_state()

    # ------------------------------------------------------------------ time-varying boundary data
    def _fused_ex(self, pr, k, w, theta, nsteps, ce, ckpt, sl=None, sr=None, sc_ps=0, sc_ss=0,
                  dt_hist=None, dt_ps=0, dt_ss=0, run_starts=None):
        """fvm_theta_run_fused_ex on a problem-major state w (P, J), in place."""
        torch = _lib.require_cuda()
        n_runs = int(run_starts.numel()) - 1 if run_starts is not None else 0
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_theta_run_fused_ex(
                ctypes.byref(pr), _lib.ptr(k), float(theta), int(nsteps), int(ce), _lib.ptr(w), _lib.ptr(ckpt),
                _lib.ptr(sl), _lib.ptr(sr), int(sc_ps), int(sc_ss), _lib.ptr(dt_hist), int(dt_ps), int(dt_ss),
                ctypes.c_void_p(run_starts.data_ptr()) if run_starts is not None else None, n_runs,
                _lib.stream_ptr())
        _lib.check(st, "fvm_theta_run_fused_ex")
        return w

    def _ex_variant(self, variant):
        if variant == "auto":
            return "fused" if self.J <= int(_lib.lib().fvm_fused_ex_max_cells()) else "thomas"
        if variant not in ("fused", "thomas"):
            raise ValueError("variant must be 'auto', 'fused' or 'thomas'")
        return variant

    def run_time_varying(self, w, theta, left_flux_t=None, right_flux_t=None, checkpoint_every=0, variant="auto"):
        """Theta steps with boundary fluxes that change from step to step: step n uses
        ``left_flux_t[n]`` / ``right_flux_t[n]`` (shape (nsteps,) or (nsteps, P)) in beta; an end
        without a history keeps the flux of set_boundary_conditions.  NOT in the reference
        (constant beta; fvm_classes.py:221,228 only mention the case) -- SURVEY.md 8(f)2.  Runs on the
        on-chip kernels (fvm_theta_run_fused_ex: the whole history in ONE launch) for J <= 1024, else on
        variant 1 (fvm_theta_step_thomas_tv); ``variant`` forces one.  Returns (w_final, checkpoints or None)."""
        torch = _lib.require_cuda()
        if not hasattr(self, "left_flux"):
            raise AttributeError("boundary conditions are not set: call set_boundary_conditions first")
        if left_flux_t is None and right_flux_t is None:
            raise ValueError("give left_flux_t and/or right_flux_t (otherwise use run)")
        if not (0.0 < float(theta) <= 1.0):
            raise ValueError("theta must be in (0, 1] for the time-varying run")
        lk, rk = _bc_kinds(self.left_flux, self.right_flux, self.left_value, self.right_value)
        if (left_flux_t is not None and lk != _lib.BC_ROBIN) or (right_flux_t is not None and rk != _lib.BC_ROBIN):
            raise ValueError("a flux history needs a flux (Robin) boundary at that end")

        variant = self._ex_variant(variant)

        def hist(x):
            if x is None:
                return None
            t = self._dev(x)
            if t.ndim == 1:
                t = t[:, None].expand(t.shape[0], self.P)
            if t.ndim != 2 or t.shape[1] != self.P:
                raise ValueError("a flux history must have shape (nsteps,) or (nsteps, P)")
            return t.contiguous()
        shared = variant == "fused" and all(x is None or np.ndim(x) == 1 for x in (left_flux_t, right_flux_t))
        if shared:          # one history for every problem: problem stride 0, nothing is expanded
            one = lambda x: None if x is None else self._dev(x).contiguous()
            sl, sr = one(left_flux_t), one(right_flux_t)
        else:
            sl, sr = hist(left_flux_t), hist(right_flux_t)
        nsteps = int((sl if sl is not None else sr).shape[0])
        if sl is not None and sr is not None and sl.shape[0] != sr.shape[0]:
            raise ValueError("the two flux histories must have the same number of steps")
        # beta is linear in the flux: factor once for a unit flux at the varying ends
        key = (sl is not None, sr is not None)
        if getattr(self, "_unit", None) is None or self._unit[0] != key:
            twin = BatchedAdvectionDiffusionModel.__new__(BatchedAdvectionDiffusionModel)
            twin.__dict__.update({k_: v for k_, v in self.__dict__.items() if k_ not in ("_unit", "_part", "_stage")})
            twin.set_boundary_conditions(left_flux=1.0 if sl is not None else self.left_flux,
                                         right_flux=1.0 if sr is not None else self.right_flux,
                                         left_value=self.left_value, right_value=self.right_value)
            self._unit = (key, twin)
        ce = int(checkpoint_every or 0)
        n_ck = nsteps // ce if ce > 0 else 0
        ckpt = torch.empty((n_ck, self.P, self.J), dtype=torch.float64, device=self.device) if n_ck else None
        if variant == "fused":
            twin = self._unit[1]
            wd = self._dev(w).reshape(self.P, self.J).clone()
            if shared:
                self._fused_ex(twin._problem(), self.k, wd, theta, nsteps, ce, ckpt, sl, sr, 0, 1)
            else:
                self._fused_ex(twin._problem(), self.k, wd, theta, nsteps, ce, ckpt, sl, sr, 1, self.P)
            return wd, ckpt
        f = self._unit[1].factor(theta)
        w_t = self._dev(w).reshape(self.P, self.J).t().contiguous()
        segments = [ce] * n_ck + ([nsteps - n_ck * ce] if nsteps - n_ck * ce else [])
        done = 0
        with torch.cuda.device(self.device):
            for i, n in enumerate(segments):
                off = done * self.P * 8
                st = _lib.lib().fvm_theta_step_thomas_tv(
                    self.J, self.P, _lib.ptr(f["fp"]), _lib.ptr(f["fq"]), _lib.ptr(f["fb"]), 1, self.P, float(theta),
                    _lib.ptr(f["bt_left"]), _lib.ptr(f["bt_right"]),
                    ctypes.c_void_p(sl.data_ptr() + off) if sl is not None else None,
                    ctypes.c_void_p(sr.data_ptr() + off) if sr is not None else None,
                    _lib.ptr(w_t), 1, self.P, int(n), _lib.stream_ptr())
                _lib.check(st, "fvm_theta_step_thomas_tv")
                done += n
                if i < n_ck:
                    ckpt[i].copy_(w_t.t())
        return w_t.t().contiguous(), ckpt

    def run_variable_dt(self, w, theta, dt_t, checkpoint_every=0, variant="auto"):
        """Theta steps whose time step changes: step n uses ``dt_t[n]`` (shape (nsteps,) or
        (nsteps, P)) in place of the constructor's k.  NOT in the reference (one k per model,
        fvm_classes.py:92,351) -- SURVEY.md 8(f)2.  The operators are re-factored whenever the time
        step differs from the previous step's; runs of equal steps share a factorisation.  J <= 1024:
        ONE launch of the on-chip kernels, which re-factor inside the kernel at the start of every run
        (fvm_theta_run_fused_ex).  Longer systems: variant 1, one fvm_theta_factor and one launch per run."""
        torch = _lib.require_cuda()
        if not (0.0 < float(theta) <= 1.0):
            raise ValueError("theta must be in (0, 1] for the variable-step run")
        dt = self._dev(dt_t)
        if dt.ndim == 1:
            dt = dt[:, None].expand(dt.shape[0], self.P)
        if dt.ndim != 2 or dt.shape[1] != self.P:
            raise ValueError("the time-step history must have shape (nsteps,) or (nsteps, P)")
        dt = dt.contiguous()
        nsteps = int(dt.shape[0])
        # boundaries of the runs of identical rows, found on the device (one small read back)
        change = torch.ones(nsteps, dtype=torch.bool, device=self.device)
        if nsteps > 1:
            change[1:] = (dt[1:] != dt[:-1]).any(dim=1)
        ce = int(checkpoint_every or 0)
        n_ck = nsteps // ce if ce > 0 else 0
        ckpt = torch.empty((n_ck, self.P, self.J), dtype=torch.float64, device=self.device) if n_ck else None
        if self._ex_variant(variant) == "fused":
            # the run boundaries never leave the device: the kernel reads them
            run_starts = torch.cat([torch.nonzero(change).flatten(),
                                    torch.tensor([nsteps], dtype=torch.int64, device=self.device)]).contiguous()
            wd = self._dev(w).reshape(self.P, self.J).clone()
            self._fused_ex(self._problem(), self.k, wd, theta, nsteps, ce, ckpt, dt_hist=dt, dt_ps=1, dt_ss=self.P,
                           run_starts=run_starts)
            return wd, ckpt
        starts = [int(i) for i in torch.nonzero(change).flatten().cpu()] + [nsteps]
        w_t = self._dev(w).reshape(self.P, self.J).t().contiguous()
        k_saved, f_saved = self.k, self._factors
        try:
            for s0, s1 in zip(starts[:-1], starts[1:]):
                self.k, self._factors = dt[s0], {}
                n = s0
                while n < s1:                         # cut the run at the checkpoints
                    stop = min(s1, (n // ce + 1) * ce) if ce > 0 else s1
                    self.step_thomas_(w_t, theta, stop - n)
                    n = stop
                    if ce > 0 and n % ce == 0 and n // ce <= n_ck:
                        ckpt[n // ce - 1].copy_(w_t.t())
        finally:
            self.k, self._factors = k_saved, f_saved
        return w_t.t().contiguous(), ckpt

    # ------------------------------------------------------------------ distances between states
    def compare(self, w_a, w_b):
        """Per problem, on the device (fvm_state_compare): max |a - b|, sum |a - b| h, max |b| and
        sum b h.  Used for steady-state detection and for the comparison with an analytic
        solution (fvm_examples.py:71)."""
        torch = _lib.require_cuda()
        a = self._dev(w_a).reshape(self.P, self.J).contiguous()
        b = self._dev(w_b).reshape(self.P, self.J).contiguous()
        out = torch.empty((self.P, 4), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            st = _lib.lib().fvm_state_compare(self.J, self.P, _lib.ptr(self.faces), _lib.ptr(a), self.J, 1,
                                              _lib.ptr(b), self.J, 1, _lib.ptr(out), _lib.stream_ptr())
        _lib.check(st, "fvm_state_compare")
        return dict(max_abs_diff=out[:, 0], l1_diff=out[:, 1], max_abs_ref=out[:, 2], area_ref=out[:, 3])

    def run_until_steady(self, w, theta, tol=1e-9, check_every=100, max_steps=1000000, variant="auto"):
        """Advance in blocks of ``check_every`` steps until, for every problem,
        max |w_n - w_{n-check_every}| <= tol * max(|w_{n-check_every}|) (or max_steps is reached).
        Returns (w, steps taken, per-problem relative change of the last block)."""
        torch = _lib.require_cuda()
        w = self._dev(w).reshape(self.P, self.J)
        steps, rel = 0, None
        while steps < int(max_steps):
            w_new, _ = self.run(w, theta, int(check_every), variant=variant)
            c = self.compare(w_new, w)
            rel = c["max_abs_diff"] / torch.clamp(c["max_abs_ref"], min=1e-300)
            w, steps = w_new, steps + int(check_every)
            done, finite = (bool(v) for v in torch.stack([(rel <= tol).all(), torch.isfinite(rel).all()]).cpu())
            if not finite:                             # diverged (Peclet / CFL violation, singular A): never steady
                raise FloatingPointError("run_until_steady: the state is no longer finite after %d steps" % steps)
            if done:                                   # one small device->host read per block
                break
        return w, steps, rel

    def _run_explicit(self, w, nsteps, ce, ckpt):
        """theta = 0: A = I, a step is w <- (I + k*alpha*M) w + beta (fvm_explicit_step)."""
        torch = _lib.require_cuda()
        asm = self.assemble()
        bufs = [w, torch.empty_like(w)]
        with torch.cuda.device(self.device):
            for n in range(nsteps):
                src, dst = bufs[n & 1], bufs[1 - (n & 1)]
                st = _lib.lib().fvm_explicit_step(
                    self.J, self.P, _lib.ptr(asm["lo"]), _lib.ptr(asm["di"]), _lib.ptr(asm["up"]), self.J, 1,
                    _lib.ptr(self.k), asm["alpha_left"], asm["alpha_right"], _lib.ptr(asm["beta_left"]),
                    _lib.ptr(asm["beta_right"]), _lib.ptr(src), _lib.ptr(dst), self.J, 1, _lib.stream_ptr())
                _lib.check(st, "fvm_explicit_step")
                if ce > 0 and (n + 1) % ce == 0 and ckpt is not None:
                    ckpt[(n + 1) // ce - 1].copy_(dst)
        return bufs[nsteps & 1]

    def run_host(self, w_host, theta, nsteps, out=None, variant="auto", chunks=None, checkpoint_every=0,
                 ckpt_out=None):
        """Host-buffer form of :meth:`run`: ``w_host`` (P, J) lives in host memory (numpy array
        or CPU tensor; pinned memory makes the copies asynchronous and fast).  Copies it to the
        device, advances it ``nsteps`` theta steps, copies the result into ``out`` (allocated
        pinned if None) and waits for it.  This is what a caller that keeps its state on the
        host -- like every script written against the reference -- pays per call.

        The batch is cut into ``chunks`` groups of problems that travel on three streams, so
        that the upload of one group, the kernel of another and the download of a third overlap
        (problems are independent: no group waits for another).  Default: groups of one wave (8 problems
        per SM), at most 64 of them -- what nothing overlaps is the first group's upload and the last
        group's download, so small groups cost least (cfg 3: 56 groups, 1.9 % faster than 14).

        ``checkpoint_every`` > 0 (fused variant): the state after every that many steps is
        copied into ``ckpt_out`` (nsteps // checkpoint_every, P, J) in host memory (allocated
        pinned if None), group by group on the same streams; returns (out, ckpt_out)."""
        torch = _lib.require_cuda()
        if isinstance(w_host, np.ndarray):
            w_host = torch.from_numpy(np.ascontiguousarray(w_host, dtype=np.float64))
        w_host = w_host.reshape(self.P, self.J)
        if getattr(self, "_stage", None) is None:
            self._stage = torch.empty((self.P, self.J), dtype=torch.float64, device=self.device)
            self._streams = [torch.cuda.Stream(device=self.device) for _ in range(3)]
        if out is None:
            out = torch.empty((self.P, self.J), dtype=torch.float64).pin_memory()
        elif isinstance(out, np.ndarray):
            out = torch.from_numpy(out)
        if variant == "auto":
            variant = self._auto_variant()
        ce = int(checkpoint_every or 0)
        n_ck = int(nsteps) // ce if ce > 0 else 0
        if n_ck and variant != "fused":
            raise ValueError("host checkpoints are written by the fused variant only; use run() otherwise")
        if n_ck:
            if ckpt_out is None:
                ckpt_out = torch.empty((n_ck, self.P, self.J), dtype=torch.float64).pin_memory()
            elif isinstance(ckpt_out, np.ndarray):
                ckpt_out = torch.from_numpy(ckpt_out)
            if tuple(ckpt_out.shape) != (n_ck, self.P, self.J) or not ckpt_out.is_contiguous():
                raise ValueError("ckpt_out must be a contiguous (%d, %d, %d) array" % (n_ck, self.P, self.J))
        with torch.cuda.device(self.device):
            if variant == "partitioned":
                self._stage.copy_(w_host, non_blocking=True)
                out.copy_(self._run_partitioned(self._stage, theta, int(nsteps), 0, None), non_blocking=True)
                torch.cuda.current_stream().synchronize()
                return out
            if variant != "fused":
                self._stage.copy_(w_host, non_blocking=True)
                w_t = self._stage.t().contiguous()
                self.step_thomas_(w_t, theta, nsteps)
                out.copy_(w_t.t(), non_blocking=True)
                torch.cuda.current_stream().synchronize()
                return out
            pr = self._problem()
            # groups of whole "waves": the on-chip kernels run 8 problems per SM at a time, so a group
            # that is a multiple of 8 x SMs leaves no SM idle in its last wave
            wave = 8 * torch.cuda.get_device_properties(self.device).multi_processor_count
            per = -(-self.P // (64 if chunks is None else max(1, int(chunks))))
            per = max(wave, -(-per // wave) * wave) if self.P >= 4 * wave else self.P
            bounds = list(range(0, self.P, per)) + [self.P]
            n_chunks = len(bounds) - 1
            main = torch.cuda.current_stream()
            for s in self._streams:
                s.wait_stream(main)
            ck_stage = None
            if n_ck:                                   # one device staging buffer per stream
                width = max(bounds[i + 1] - bounds[i] for i in range(n_chunks))
                key = (n_ck, width)
                if getattr(self, "_stage_ck", None) is None or self._stage_ck[0] != key:
                    self._stage_ck = (key, [torch.empty((n_ck, width, self.J), dtype=torch.float64, device=self.device)
                                            for _ in self._streams])
                ck_stage = self._stage_ck[1]
            for i in range(n_chunks):
                lo, hi = bounds[i], bounds[i + 1]
                st = self._streams[i % len(self._streams)]
                with torch.cuda.stream(st):
                    dst = self._stage[lo:hi]
                    dst.copy_(w_host[lo:hi], non_blocking=True)
                    sub = self._chunk_problem(pr, lo, hi)
                    ck = None
                    if n_ck:                           # the kernel's layout is (n, hi-lo, J), contiguous
                        ck = ck_stage[i % len(self._streams)].view(-1)[:n_ck * (hi - lo) * self.J].view(n_ck, hi - lo, self.J)
                    rc = _lib.lib().fvm_theta_run_fused(ctypes.byref(sub), ctypes.c_void_p(self.k.data_ptr() + 8 * lo),
                                                        float(theta), int(nsteps), ce if n_ck else 0, _lib.ptr(dst),
                                                        _lib.ptr(ck), _lib.stream_ptr())
                    _lib.check(rc, "fvm_theta_run_fused")
                    out[lo:hi].copy_(dst, non_blocking=True)
                    for n in range(n_ck):
                        ckpt_out[n, lo:hi].copy_(ck[n], non_blocking=True)
            for s in self._streams:
                main.wait_stream(s)
            main.synchronize()
        return (out, ckpt_out) if n_ck else out

    def _chunk_problem(self, pr, lo, hi):
        """struct fvm_problem of the problems [lo, hi) of this batch (same mesh, shifted pointers)."""
        sub = _lib.FvmProblem()
        ctypes.memmove(ctypes.byref(sub), ctypes.byref(pr), ctypes.sizeof(sub))
        sub.P = hi - lo
        sub.a = pr.a + 8 * lo * pr.a_ps
        sub.d = pr.d + 8 * lo * pr.d_ps
        if pr.reaction:
            sub.reaction = pr.reaction + 8 * lo * pr.r_ps
        if pr.left_flux:
            sub.left_flux = pr.left_flux + 8 * lo
        if pr.right_flux:
            sub.right_flux = pr.right_flux + 8 * lo
        return sub
