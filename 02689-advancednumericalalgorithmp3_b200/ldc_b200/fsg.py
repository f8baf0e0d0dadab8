"""Batched drivers: single-grid solve (base.solve semantics) and the coarse-to-fine FSG sequence
(solve_fsg, spectral/multigrid/fsg.py:1053-1221) for a batch of cavities that share (N, Lx, Ly, basis)."""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import operators as ops
from .batched import BatchedSpectralLDC


@dataclass
class FSGResult:
    u: np.ndarray            # [B, n, n]
    v: np.ndarray
    p: np.ndarray            # [B, m, m] inner grid
    iterations: np.ndarray   # int64 [B]; FSG: summed over levels
    converged: np.ndarray    # bool [B]
    diverged: np.ndarray     # bool [B]
    rel_residual: np.ndarray
    per_level_iterations: list = field(default_factory=list)
    orders: list = field(default_factory=list)
    kernel_ms: float = 0.0
    launches: int = 0
    finest: BatchedSpectralLDC | None = None
    zero_mask: np.ndarray | None = None   # FSG: cavities whose finest level was never reached (fields reported as zeros)


def solve_sg_batched(N, Re, tolerance, max_iter, *, beta_squared=5.0, CFL=1.5, lid_velocity=1.0,
                     corner_treatment="smoothing", corner_smoothing=0.15, Lx=1.0, Ly=1.0, basis="chebyshev",
                     device=0, path="auto", keep_handle=False, host_fields=True) -> FSGResult:
    """`LidDrivenCavitySolver.solve` + `SGSolver.step` for a batch: zero start + lid, frozen-pressure RK4
    (SURVEY Q1), converged iff i >= 10 and rel < tol (base.py:243-313).  The reference has no NaN exit on this
    path (it would spin to max_iterations on NaNs); a non-finite residual here stops the cavity and reports
    iterations = max_iter, converged = False, which is the state the reference would end in."""
    B = int(np.broadcast(np.asarray(Re), np.asarray(corner_smoothing), np.asarray(tolerance)).size)
    s = BatchedSpectralLDC(N, N, B, Lx, Ly, basis, device, path=path)
    s.set_cases(Re, beta_squared, CFL, lid_velocity, corner_treatment, corner_smoothing)
    s.init_state()
    r = s.solve(tolerance, max_iter, warmup=10, freeze_pressure=True, nan_exit=True)
    iters = np.where(r.diverged, max_iter, r.iterations)
    u, v, p = s.get_state() if host_fields else (None, None, None)
    out = FSGResult(u, v, p, iters, r.converged, r.diverged, r.rel_residual, [r.iterations.copy()], [N], r.kernel_ms,
                    s.launches, s if keep_handle else None)
    if not keep_handle:
        s.close()
    return out


def solve_fsg_batched(N, Re, tolerance, max_iter, *, n_levels=2, coarse_tolerance_factor=1.0, prolongation="fft",
                      beta_squared=5.0, CFL=1.5, lid_velocity=1.0, corner_treatment="smoothing",
                      corner_smoothing=0.15, Lx=1.0, Ly=1.0, basis="chebyshev", device=0, path="auto",
                      keep_handle=False, host_fields=True) -> FSGResult:
    """Coarse-to-fine sequence for a batch.  Per level: tol * ctf**levels_from_finest, no warm-up, staged
    pressure, NaN/Inf exit; max_iter applies per level; iterations are summed over levels.  A cavity that
    diverges on a COARSER level leaves every finer level untouched: zero fields, converged=False (SURVEY Q8,
    multigrid/fsg.py:1200-1202); one that diverges on the finest level keeps that level's non-finite fields, as
    `solve_fsg` returns them.  host_fields=False leaves the fields on the device (keep_handle=True: `finest`;
    `zero_mask` marks the cavities to report as zeros)."""
    if prolongation not in ("fft", "polynomial"):
        raise ValueError(f"Unknown prolongation method: {prolongation}. Use 'fft' or 'polynomial'.")
    B = int(np.broadcast(np.asarray(Re), np.asarray(corner_smoothing), np.asarray(tolerance)).size)
    tol = np.broadcast_to(np.asarray(tolerance, dtype=np.float64), (B,))
    orders = ops.hierarchy_orders(N, n_levels)
    total = np.zeros(B, dtype=np.int64)
    dead = np.zeros(B, dtype=bool)       # diverged on an earlier level
    conv = np.zeros(B, dtype=bool)
    rel = np.zeros(B)
    per_level, kernel_ms, launches = [], 0.0, 0
    prev = None
    for li, n in enumerate(orders):
        lvl = BatchedSpectralLDC(n, n, B, Lx, Ly, basis, device, path=path)
        lvl.set_cases(Re, beta_squared, CFL, lid_velocity, corner_treatment, corner_smoothing)
        if prev is None:
            lvl.init_state()
        else:
            prev.prolongate_to(lvl, prolongation, bc_quirk=True)
            launches += prev.launches
            prev.close()
        r = lvl.solve(tol * coarse_tolerance_factor ** (len(orders) - 1 - li), max_iter, warmup=0,
                      freeze_pressure=False, nan_exit=True)
        kernel_ms += r.kernel_ms
        it = np.where(dead, 0, r.iterations)
        total += it
        per_level.append(it)
        conv = np.where(dead, False, r.converged)
        rel = np.where(dead, rel, r.rel_residual)
        dead_before = dead.copy()   # dead before THIS level: its fields were never legitimately stepped
        dead = dead | r.diverged
        prev = lvl
    launches += prev.launches
    u = v = p = None
    if host_fields:
        u, v, p = prev.get_state()
        u[dead_before] = 0.0
        v[dead_before] = 0.0
        p[dead_before] = 0.0
    out = FSGResult(u, v, p, total, conv & ~dead, dead, rel, per_level, orders, kernel_ms, launches,
                    prev if keep_handle else None, dead_before)
    if not keep_handle:
        prev.close()
    return out
