"""BatchedSpectralLDC: a batch of independent lid-driven cavities of one polynomial order on one GPU.

Thin host wrapper over the C ABI (include/ldc_b200.h).  All numerics run in the CUDA library; this class
only builds the per-order operators (operators.py), marshals NumPy / torch buffers and keeps the handle.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from . import operators as ops

_PATHS = {"auto": _lib.LDC_PATH_AUTO, "tiled": _lib.LDC_PATH_TILED, "resident": _lib.LDC_PATH_RESIDENT,
          "resident_eo": _lib.LDC_PATH_RESIDENT_EO, "stream_eo": _lib.LDC_PATH_STREAM_EO,
          "resident_eo_split": _lib.LDC_PATH_RESIDENT_EO_SPLIT, "stream_tma": _lib.LDC_PATH_STREAM_TMA}


@dataclass
class SolveResult:
    iterations: np.ndarray   # int64 [B]
    converged: np.ndarray    # bool  [B]
    diverged: np.ndarray     # bool  [B]
    rel_residual: np.ndarray  # float64 [B], last max(u_res, v_res)
    kernel_ms: float


def _per_case(value, B, name):
    a = np.ascontiguousarray(np.broadcast_to(np.asarray(value, dtype=np.float64), (B,)))
    if a.shape != (B,):
        raise ValueError(f"{name} must be a scalar or have shape ({B},)")
    return a.copy()


class BatchedSpectralLDC:
    """State + operators of `batch` cavities at order (nx, ny); see the reference's SGSolver / SpectralLevel."""

    def __init__(self, nx: int, ny: int | None = None, batch: int = 1, Lx: float = 1.0, Ly: float = 1.0,
                 basis: str = "chebyshev", device: int = 0, stream: int | None = None, path: str = "auto"):
        self.lib = _lib.load()
        ny = nx if ny is None else ny
        self.nx, self.ny, self.B = int(nx), int(ny), int(batch)
        if self.nx < 2 or self.ny < 2 or self.B < 1:
            raise ValueError("need polynomial orders nx, ny >= 2 and batch >= 1")
        self.Lx, self.Ly, self.basis = float(Lx), float(Ly), basis
        self.device = int(device)
        self.level = ops.build_level(self.nx, self.ny, self.Lx, self.Ly, basis)
        self._h = C.c_void_p()
        _lib.check(self.lib.ldc_create(C.byref(self._h), self.nx, self.ny, self.B, self.device,
                                       C.c_void_p(stream) if stream else None), "ldc_create")
        lv = self.level
        keep = [np.ascontiguousarray(a, dtype=np.float64) for a in (lv.Dx, lv.Dxx, lv.Dy, lv.Dyy, lv.Ix, lv.Iy)]
        _lib.check(self.lib.ldc_set_operators(self._h, *[_lib.ptr(a) for a in keep], lv.dx_min, lv.dy_min),
                   "ldc_set_operators")
        self.set_path(path)
        self.Re = self.beta_squared = self.CFL = self.lid_velocity = None
        self.u_lid = self.v_lid = None

    # -- lifetime ------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.ldc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- configuration -------------------------------------------------------------------------------
    def set_path(self, path: str):
        if path not in _PATHS:
            raise ValueError(f"path must be one of {sorted(_PATHS)}")
        _lib.check(self.lib.ldc_set_path(self._h, _PATHS[path]), "ldc_set_path")
        self.path = path

    @property
    def resident_supported(self) -> bool:
        return self.lib.ldc_resident_supported(self._h) == 0

    @property
    def kernel_name(self) -> str:
        """Stepping kernel ldc_step / ldc_solve will launch for this handle and path."""
        return self.lib.ldc_kernel_name(self._h).decode()

    def set_cases(self, Re, beta_squared=5.0, CFL=1.5, lid_velocity=1.0, corner_treatment="smoothing",
                  corner_smoothing=0.15, u_lid=None, v_lid=None):
        """Per-cavity physics.  Scalars broadcast over the batch; corner_treatment may be one string or a list."""
        B = self.B
        self.Re = _per_case(Re, B, "Re")
        self.beta_squared = _per_case(beta_squared, B, "beta_squared")
        self.CFL = _per_case(CFL, B, "CFL")
        self.lid_velocity = _per_case(lid_velocity, B, "lid_velocity")
        if u_lid is None:
            cs = _per_case(corner_smoothing, B, "corner_smoothing")
            methods = [corner_treatment] * B if isinstance(corner_treatment, str) else list(corner_treatment)
            if len(methods) != B:
                raise ValueError("corner_treatment must be one string or one per cavity")
            u_lid = np.empty((B, self.nx + 1))
            v_lid = np.empty((B, self.nx + 1))
            for b in range(B):
                u_lid[b], v_lid[b] = ops.lid_velocity_profile(self.level.x, methods[b], cs[b], self.lid_velocity[b],
                                                              self.Lx)
        self.u_lid = np.ascontiguousarray(u_lid, dtype=np.float64).reshape(B, self.nx + 1)
        self.v_lid = (np.zeros_like(self.u_lid) if v_lid is None
                      else np.ascontiguousarray(v_lid, dtype=np.float64).reshape(B, self.nx + 1))
        _lib.check(self.lib.ldc_set_cases(self._h, _lib.ptr(self.Re), _lib.ptr(self.beta_squared), _lib.ptr(self.CFL),
                                          _lib.ptr(self.lid_velocity), _lib.ptr(self.u_lid), _lib.ptr(self.v_lid)),
                   "ldc_set_cases")
        return self

    # -- state ---------------------------------------------------------------------------------------
    @property
    def shape_full(self):
        return (self.B, self.nx + 1, self.ny + 1)

    @property
    def shape_inner(self):
        return (self.B, self.nx - 1, self.ny - 1)

    def init_state(self):
        _lib.check(self.lib.ldc_init_state(self._h), "ldc_init_state")
        return self

    def _as_in(self, a, shape, name):
        if a is None:
            return None
        if hasattr(a, "data_ptr"):
            if tuple(a.shape) != shape or str(a.dtype) != "torch.float64":
                raise ValueError(f"{name} must be float64 of shape {shape}")
            return a.contiguous()
        a = np.ascontiguousarray(a, dtype=np.float64)
        if a.size != int(np.prod(shape)):
            raise ValueError(f"{name} must have shape {shape}")
        return a.reshape(shape)

    def set_state(self, u=None, v=None, p=None):
        u = self._as_in(u, self.shape_full, "u")
        v = self._as_in(v, self.shape_full, "v")
        p = self._as_in(p, self.shape_inner, "p")
        _lib.check(self.lib.ldc_set_state(self._h, _lib.ptr(u), _lib.ptr(v), _lib.ptr(p)), "ldc_set_state")
        return self

    def get_state(self, out=None):
        """(u, v, p) as NumPy arrays, or into the given (u, v, p) buffers (NumPy or torch, host or device)."""
        if out is None:
            out = (np.empty(self.shape_full), np.empty(self.shape_full), np.empty(self.shape_inner))
        u, v, p = out
        _lib.check(self.lib.ldc_get_state(self._h, _lib.ptr(u), _lib.ptr(v), _lib.ptr(p)), "ldc_get_state")
        return u, v, p

    # -- numerics ------------------------------------------------------------------------------------
    def residual(self):
        """(R_u, R_v, R_p, norms[B,3]) of the current state on all nodes."""
        Ru, Rv, Rp = np.empty(self.shape_full), np.empty(self.shape_full), np.empty(self.shape_inner)
        norms = np.empty((self.B, 3))
        _lib.check(self.lib.ldc_residual(self._h, _lib.ptr(Ru), _lib.ptr(Rv), _lib.ptr(Rp), _lib.ptr(norms)),
                   "ldc_residual")
        return Ru, Rv, Rp, norms

    def residual_norms(self):
        norms = np.empty((self.B, 3))
        _lib.check(self.lib.ldc_residual(self._h, None, None, None, _lib.ptr(norms)), "ldc_residual")
        return norms

    def step(self, n_steps: int = 1, freeze_pressure: bool = False, want_rel: bool = True):
        """n_steps RK4 pseudo-steps of every cavity; returns [B,2] (u_res, v_res) of the last step."""
        rel = np.empty((self.B, 2)) if want_rel else None
        _lib.check(self.lib.ldc_step(self._h, int(n_steps), int(bool(freeze_pressure)), _lib.ptr(rel)), "ldc_step")
        return rel

    def solve(self, tolerance, max_iter: int, warmup: int = 0, freeze_pressure: bool = False,
              nan_exit: bool = True) -> SolveResult:
        tol = _per_case(tolerance, self.B, "tolerance")
        iters = np.zeros(self.B, dtype=np.int64)
        conv = np.zeros(self.B, dtype=np.int32)
        divg = np.zeros(self.B, dtype=np.int32)
        rel = np.zeros(self.B)
        _lib.check(self.lib.ldc_solve(self._h, _lib.ptr(tol), int(max_iter), int(warmup), int(bool(freeze_pressure)),
                                      int(bool(nan_exit)), _lib.ptr(iters), _lib.ptr(conv), _lib.ptr(divg),
                                      _lib.ptr(rel)), "ldc_solve")
        return SolveResult(iters, conv.astype(bool), divg.astype(bool), rel, self.last_kernel_ms)

    def solve_history(self, tolerance, max_iter: int, warmup: int = 10, freeze_pressure: bool = True,
                      nan_exit: bool = False, chunk: int = 8192):
        """`solve` with the per-iteration records of base.solve (base.py:243-313).

        Returns (SolveResult, history) where history[b] is an (iterations_b, 7) array, one row per iteration taken:
        rel_iter, ||R_u||, ||R_v||, ||R_p|| (last stage's residual), energy, enstrophy, palinstrophy.  The solve is
        streamed through `ldc_solve_history` in chunks of `chunk` iterations so that host and device memory stay
        bounded whatever max_iter is.
        """
        tol = _per_case(tolerance, self.B, "tolerance")
        iters = np.zeros(self.B, dtype=np.int64)
        conv = np.zeros(self.B, dtype=np.int32)
        divg = np.zeros(self.B, dtype=np.int32)
        rel = np.zeros(self.B)
        W = np.ascontiguousarray(self.level.W)
        chunk = int(max(1, min(chunk, max_iter)))
        buf = np.zeros((self.B, chunk, 7))
        rows = [[] for _ in range(self.B)]
        start, ms = 0, 0.0
        while start < max_iter:
            _lib.check(self.lib.ldc_solve_history(self._h, _lib.ptr(tol), start, int(max_iter), int(warmup),
                                                  int(bool(freeze_pressure)), int(bool(nan_exit)), _lib.ptr(W),
                                                  _lib.ptr(buf), chunk, _lib.ptr(iters), _lib.ptr(conv), _lib.ptr(divg),
                                                  _lib.ptr(rel)), "ldc_solve_history")
            ms += self.last_kernel_ms
            for b in range(self.B):
                taken = int(min(iters[b], start + chunk)) - start
                if taken > 0:
                    rows[b].append(buf[b, :taken].copy())
            start += chunk
            if not np.any((iters >= start) & (conv == 0) & (divg == 0)):
                break
        hist = [np.concatenate(r) if r else np.zeros((0, 7)) for r in rows]
        return SolveResult(iters, conv.astype(bool), divg.astype(bool), rel, ms), hist

    @staticmethod
    def sample_iterations(n_iter: int, skip: int = 10, max_points: int = 1000):
        """The iterations whose records the reference's TimeSeries holds: records of iterations >= skip are kept
        (base.py:262-276) and thinned to max_points by np.linspace(0, L-1, max_points, dtype=int) (base.py:137-141).
        The last entry is always the final iteration (Metrics are read from it)."""
        kept = int(n_iter) - skip
        if kept <= 0:
            return np.zeros(0, dtype=np.int64)
        if kept <= max_points:
            return skip + np.arange(kept, dtype=np.int64)
        return skip + np.linspace(0, kept - 1, max_points, dtype=int).astype(np.int64)

    # replay route of solve_history_sampled: at most this many snapshot segments (= cavities of the replay batch) ...
    REPLAY_MAX_SEGMENTS = 512
    # ... first snapshot spacing (doubled whenever the segment count would pass the limit) ...
    REPLAY_SEGMENT = 512
    # ... and only for batches up to this size (larger ones take the two-pass route)
    REPLAY_MAX_BATCH = 16
    # ... whose replay needs no more than this for its records (else: two passes on this handle)
    REPLAY_MAX_BYTES = 128 << 20

    def solve_segment(self, tolerance, start: int, stop: int, max_iter: int, warmup: int = 0,
                      freeze_pressure: bool = False, nan_exit: bool = True) -> SolveResult:
        """Iterations [start, stop) of the solve `solve(tolerance, max_iter, ...)` (`ldc_solve_segment`): start = 0
        starts it, start > 0 continues the previous call.  Cutting a solve into segments does not change its steps."""
        tol = _per_case(tolerance, self.B, "tolerance")
        iters = np.zeros(self.B, dtype=np.int64)
        conv = np.zeros(self.B, dtype=np.int32)
        divg = np.zeros(self.B, dtype=np.int32)
        rel = np.zeros(self.B)
        _lib.check(self.lib.ldc_solve_segment(self._h, _lib.ptr(tol), int(start), int(stop), int(max_iter), int(warmup),
                                              int(bool(freeze_pressure)), int(bool(nan_exit)), _lib.ptr(iters),
                                              _lib.ptr(conv), _lib.ptr(divg), _lib.ptr(rel)), "ldc_solve_segment")
        return SolveResult(iters, conv.astype(bool), divg.astype(bool), rel, self.last_kernel_ms)

    def _sampled_records(self, tol, max_iter, warmup, freeze_pressure, nan_exit, rec):
        """`ldc_solve_sampled` from the current state: (iters, conv, divg, rel, buf[B, len(rec), 7])."""
        iters = np.zeros(self.B, dtype=np.int64)
        conv = np.zeros(self.B, dtype=np.int32)
        divg = np.zeros(self.B, dtype=np.int32)
        rel = np.zeros(self.B)
        W = np.ascontiguousarray(self.level.W)
        rec = np.ascontiguousarray(rec, dtype=np.int64)
        buf = np.zeros((self.B, rec.size, 7))
        _lib.check(self.lib.ldc_solve_sampled(self._h, _lib.ptr(tol), int(max_iter), int(warmup),
                                              int(bool(freeze_pressure)), int(bool(nan_exit)), _lib.ptr(W), rec.size,
                                              _lib.ptr(rec), _lib.ptr(buf), _lib.ptr(iters), _lib.ptr(conv),
                                              _lib.ptr(divg), _lib.ptr(rel)), "ldc_solve_sampled")
        return iters, conv, divg, rel, buf

    def solve_history_sampled(self, tolerance, max_iter: int, warmup: int = 10, freeze_pressure: bool = True,
                              nan_exit: bool = False, skip: int = 10, max_points: int = 1000, method: str = "auto"):
        """The down-sampled history the reference ends up with, without a record per iteration.

        Which iterations survive the reference's thinning (`sample_iterations`) is known once a cavity's iteration count
        is, so the solve comes first and the records second.  Returns (SolveResult, history, iterations): history[b] is
        (n_b, 7) in the layout of `solve_history`, iterations[b] the 0-based iteration of each row.  The SolveResult and
        the final state are those of `solve` (bitwise) on both routes:

        * "replay" (`auto` for batch <= REPLAY_MAX_BATCH): the solve is cut into segments (`ldc_solve_segment`) with a
          snapshot of the state at every segment start.  The segments are then independent of each other: ONE batch
          holding a cavity per (cavity, segment), started from the snapshots on a second handle, takes the steps of a
          single segment on the history path (`ldc_solve_history`: one launch for the lot on grids that fit an SM) and
          the kept records are picked from its rows -- the time of one segment instead of a second pass over the whole
          trajectory.
        * "two_pass": the solve is repeated from the same start state on this handle (the stepping kernels are bitwise
          reproducible) and the kept records are formed on the way.
        Either way a record is what the history path forms for that iteration; the routes agree with each other and
        with the every-iteration history of `solve_history` to rounding (a replayed segment starts from the unpacked
        snapshot and is stepped by the history kernels).
        """
        if method not in ("auto", "replay", "two_pass"):
            raise ValueError("method must be 'auto', 'replay' or 'two_pass'")
        if method == "auto":
            method = "replay" if self.B <= self.REPLAY_MAX_BATCH else "two_pass"
        if method == "replay":
            return self._sampled_by_replay(tolerance, int(max_iter), warmup, freeze_pressure, nan_exit, skip, max_points)
        state0 = self.get_state()
        first = self.solve(tolerance, max_iter, warmup=warmup, freeze_pressure=freeze_pressure, nan_exit=nan_exit)
        idx = [self.sample_iterations(n, skip, max_points) for n in first.iterations]
        rec = np.unique(np.concatenate(idx)) if self.B > 1 else idx[0]
        if rec.size == 0:
            return first, [np.zeros((0, 7)) for _ in range(self.B)], idx
        self.set_state(*state0)
        tol = _per_case(tolerance, self.B, "tolerance")
        iters, conv, divg, rel, buf = self._sampled_records(tol, max_iter, warmup, freeze_pressure, nan_exit, rec)
        res = SolveResult(iters, conv.astype(bool), divg.astype(bool), rel, first.kernel_ms + self.last_kernel_ms)
        if not np.array_equal(iters, first.iterations):
            raise RuntimeError("solve_history_sampled: the second pass took a different number of iterations "
                               f"({iters.tolist()} vs {first.iterations.tolist()}); use solve_history")
        hist = [buf[b, np.searchsorted(rec, idx[b])] for b in range(self.B)]
        return res, hist, idx

    def _sampled_by_replay(self, tolerance, max_iter, warmup, freeze_pressure, nan_exit, skip, max_points):
        B = self.B
        max_seg = max(2, self.REPLAY_MAX_SEGMENTS // B)
        seg = int(self.REPLAY_SEGMENT)
        snaps = []   # (iteration, (u, v, p)) at the start of every segment
        ms, done, res = 0.0, 0, None
        while True:
            if len(snaps) == max_seg:   # keep every other snapshot, segments twice as long from here on
                snaps = snaps[::2]
                seg *= 2
            if done % seg == 0:
                snaps.append((done, self.get_state()))
            stop = min(max_iter, (done // seg + 1) * seg)
            res = self.solve_segment(tolerance, done, stop, max_iter, warmup=warmup, freeze_pressure=freeze_pressure,
                                     nan_exit=nan_exit)
            ms += res.kernel_ms
            done = stop
            if done >= max_iter or not np.any((res.iterations >= done) & ~res.converged & ~res.diverged):
                break
        idx = [self.sample_iterations(n, skip, max_points) for n in res.iterations]
        if sum(i.size for i in idx) == 0:
            return SolveResult(res.iterations, res.converged, res.diverged, res.rel_residual, ms), \
                [np.zeros((0, 7)) for _ in range(B)], idx
        # one replay cavity per (cavity, segment that holds a kept iteration)
        starts = np.array([it for it, _ in snaps], dtype=np.int64)
        pairs = sorted({(b, int(j)) for b in range(B) for j in np.unique(np.searchsorted(starts, idx[b], "right") - 1)})
        slot = {bj: k for k, bj in enumerate(pairs)}
        offsets = np.unique(np.concatenate([idx[b] - starts[np.searchsorted(starts, idx[b], "right") - 1]
                                            for b in range(B)]))
        sel = np.array([b for b, _ in pairs])
        n_local = int(offsets[-1]) + 1
        if len(pairs) * n_local * 56 > self.REPLAY_MAX_BYTES:
            # a very long solve of several cavities: segments so long that the replay batch would need a record buffer of
            # this size; repeat the solve on this handle instead (the two-pass route from the first snapshot)
            self.set_state(*snaps[0][1])
            tol = _per_case(tolerance, B, "tolerance")
            rec = np.unique(np.concatenate(idx))
            iters, conv, divg, rel, buf = self._sampled_records(tol, max_iter, warmup, freeze_pressure, nan_exit, rec)
            if not np.array_equal(iters, res.iterations):
                raise RuntimeError("solve_history_sampled: the second pass took a different number of iterations "
                                   f"({iters.tolist()} vs {res.iterations.tolist()}); use solve_history")
            return SolveResult(iters, conv.astype(bool), divg.astype(bool), rel, ms + self.last_kernel_ms), \
                [buf[b, np.searchsorted(rec, idx[b])] for b in range(B)], idx
        with BatchedSpectralLDC(self.nx, self.ny, batch=len(pairs), Lx=self.Lx, Ly=self.Ly, basis=self.basis,
                                device=self.device) as r:
            if self.path == "tiled":
                r.set_path("tiled")
            r.set_cases(self.Re[sel], self.beta_squared[sel], self.CFL[sel], self.lid_velocity[sel],
                        u_lid=self.u_lid[sel], v_lid=self.v_lid[sel])
            r.set_state(*[np.stack([snaps[j][1][f][b] for b, j in pairs]) for f in range(3)])
            # every replay cavity takes the steps of one segment on the history path, ONE launch for the lot where the grid
            # fits an SM (k_history_fused), with a record per iteration; tolerance 0: a replay cavity never stops by
            # itself (iteration numbers are local to the segment there)
            rr, rows = r.solve_history(0.0, n_local, warmup=0, freeze_pressure=freeze_pressure, nan_exit=nan_exit,
                                       chunk=n_local)
            ms += rr.kernel_ms
        buf = np.zeros((len(pairs), n_local, 7))
        for k, a in enumerate(rows):     # a cavity that went non-finite stopped early: its later rows stay zero
            buf[k, :a.shape[0]] = a
        offsets = np.arange(n_local)
        hist = []
        for b in range(B):
            j = np.searchsorted(starts, idx[b], "right") - 1
            rows = np.array([slot[(b, int(jj))] for jj in j], dtype=np.int64)
            hist.append(buf[rows, np.searchsorted(offsets, idx[b] - starts[j])] if idx[b].size else np.zeros((0, 7)))
        return SolveResult(res.iterations, res.converged, res.diverged, res.rel_residual, ms), hist, idx

    def diagnostics(self):
        """[B,3] = (energy, enstrophy, palinstrophy)."""
        out = np.empty((self.B, 3))
        W = np.ascontiguousarray(self.level.W)
        _lib.check(self.lib.ldc_diagnostics(self._h, _lib.ptr(W), _lib.ptr(out)), "ldc_diagnostics")
        return out

    def vortex_metrics(self, operands, want_psi: bool = False):
        """[B, 19] vortex metrics (order: `postprocess.VORTEX_KEYS`) of the state on the device, psi [B, n, n] if asked;
        `operands` = `postprocess.diagonalisation_operands(level)`."""
        out = np.empty((self.B, 19))
        psi = np.empty(self.shape_full) if want_psi else None
        _lib.check(self.lib.ldc_vortex_metrics(self._h, *[_lib.ptr(a) for a in operands], _lib.ptr(out), _lib.ptr(psi)),
                   "ldc_vortex_metrics")
        return out, psi

    def validation_errors(self, plan, ref_u, ref_v):
        """[B,3] = (u_L2_error, v_L2_error, n_valid) against a stored solution at scattered points
        (compute_validation_errors, base.py:970-1054); `plan` is a `validation.BilinearPlan` for this grid."""
        ref_u = np.ascontiguousarray(ref_u, dtype=np.float64).ravel()
        ref_v = np.ascontiguousarray(ref_v, dtype=np.float64).ravel()
        nq = plan.cell.size
        if not (ref_u.size == ref_v.size == plan.tx.size == plan.ty.size == plan.mask.size == nq):
            raise ValueError("plan and reference arrays must have one entry per point")
        out = np.empty((self.B, 3))
        _lib.check(self.lib.ldc_validation_errors(self._h, nq, _lib.ptr(plan.cell), _lib.ptr(plan.tx), _lib.ptr(plan.ty),
                                                  _lib.ptr(plan.mask), _lib.ptr(ref_u), _lib.ptr(ref_v), _lib.ptr(out)),
                   "ldc_validation_errors")
        return out

    def prolongate_to(self, fine: "BatchedSpectralLDC", method: str = "fft", bc_quirk: bool = True):
        """Coarse -> fine initial guess (prolongate_solution, multigrid/fsg.py:551-614)."""
        Pxu = np.ascontiguousarray(ops.prolongation_matrix(self.nx + 1, fine.nx + 1, method))
        Pyu = np.ascontiguousarray(ops.prolongation_matrix(self.ny + 1, fine.ny + 1, method))
        Pxp = np.ascontiguousarray(ops.prolongation_matrix(self.nx - 1, fine.nx - 1, method))
        Pyp = np.ascontiguousarray(ops.prolongation_matrix(self.ny - 1, fine.ny - 1, method))
        _lib.check(self.lib.ldc_prolongate(self._h, fine._h, _lib.ptr(Pxu), _lib.ptr(Pyu), _lib.ptr(Pxp),
                                           _lib.ptr(Pyp), int(bool(bc_quirk))), "ldc_prolongate")
        return fine

    def restrict_to(self, coarse: "BatchedSpectralLDC", method: str = "fft"):
        Rxu = np.ascontiguousarray(ops.restriction_matrix(self.nx + 1, coarse.nx + 1, method))
        Ryu = np.ascontiguousarray(ops.restriction_matrix(self.ny + 1, coarse.ny + 1, method))
        Rxp = np.ascontiguousarray(ops.restriction_matrix(self.nx - 1, coarse.nx - 1, method))
        Ryp = np.ascontiguousarray(ops.restriction_matrix(self.ny - 1, coarse.ny - 1, method))
        _lib.check(self.lib.ldc_restrict(self._h, coarse._h, _lib.ptr(Rxu), _lib.ptr(Ryu), _lib.ptr(Rxp),
                                         _lib.ptr(Ryp)), "ldc_restrict")
        return coarse

    # -- accounting ----------------------------------------------------------------------------------
    @property
    def launches(self) -> int:
        return int(self.lib.ldc_launch_count(self._h))

    @property
    def last_kernel_ms(self) -> float:
        return float(self.lib.ldc_last_kernel_ms(self._h))


def fp64_peak_tflops(device: int = 0, which: str = "dmma") -> float:
    """Live register-resident FP64 ceiling of the device: 'dmma', 'dfma' or 'mixed'."""
    lib = _lib.load()
    out = C.c_double()
    _lib.check(lib.ldc_fp64_peak(device, {"dmma": 0, "dfma": 1, "mixed": 2}[which], C.byref(out)), "ldc_fp64_peak")
    return out.value
