"""L2 error against the stored finite-volume solutions without PyVista.

Reference: `LidDrivenCavitySolver.compute_validation_errors` (src/solvers/base.py:970-1054) reads
`data/validation/{fv,fv-regu}/Re<Re>/solution.vts` with `pv.read`, evaluates the computed u, v at the stored points by
*bilinear* interpolation (`_evaluate_at_points`, base.py:1122-1160 -- SciPy's RegularGridInterpolator, not the spectral
interpolant) and takes relative L2 norms over the interior points.  Here the file is parsed by `ldc_b200.vts`, the host
builds the interpolation plan once (cell index + normalised distances, shared by every cavity on the same grid) and the
evaluation + reduction run on the device for the whole batch (`ldc_validation_errors`).
"""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass

import numpy as np

from .vts import read_vts

log = logging.getLogger(__name__)

# base.py:997-1000 -- the reference ignores its `reference_dir` argument and always looks at these two
REF_DIRS = (("data/validation/fv", ""), ("data/validation/fv-regu", "_regu"))
MARGIN = 1e-10  # base.py:1022


@dataclass
class StoredSolution:
    """Point cloud of a stored solution: what base.py:1004-1013 takes from the PyVista mesh."""

    x: np.ndarray
    y: np.ndarray
    u: np.ndarray
    v: np.ndarray

    @classmethod
    def from_vts(cls, path: str) -> "StoredSolution":
        g = read_vts(path)
        return cls(np.ascontiguousarray(g.points[:, 0]), np.ascontiguousarray(g.points[:, 1]),
                   np.ascontiguousarray(g.point_data["u"], dtype=np.float64),
                   np.ascontiguousarray(g.point_data["v"], dtype=np.float64))


@dataclass
class BilinearPlan:
    cell: np.ndarray   # int32 [nq]: ix * n1y + iy of the cell's lower node
    tx: np.ndarray     # float64 [nq]
    ty: np.ndarray
    mask: np.ndarray   # int32 [nq]: 1 = use the point


def _find_interval(grid: np.ndarray, q: np.ndarray):
    """RegularGridInterpolator._find_indices for one axis: i = searchsorted(grid, q) - 1 clipped to [0, n-2],
    t = (q - grid[i]) / (grid[i+1] - grid[i])."""
    i = np.searchsorted(grid, q) - 1
    i[i < 0] = 0
    i[i > grid.size - 2] = grid.size - 2
    t = (q - grid[i]) / (grid[i + 1] - grid[i])
    return i.astype(np.int64), t


def bilinear_plan(x_nodes, y_nodes, xq, yq, Lx: float, Ly: float) -> BilinearPlan:
    """Plan for evaluating fields stored as F[ix, iy] on (x_nodes, y_nodes) at the points (xq, yq).

    mask = inside the grid (bounds_error=False, fill_value=nan: outside points are NaN and dropped, base.py:1028)
    and strictly inside the walls by MARGIN (base.py:1022-1027)."""
    x_nodes = np.asarray(x_nodes, dtype=np.float64)
    y_nodes = np.asarray(y_nodes, dtype=np.float64)
    xq = np.asarray(xq, dtype=np.float64).ravel()
    yq = np.asarray(yq, dtype=np.float64).ravel()
    ix, tx = _find_interval(x_nodes, xq)
    iy, ty = _find_interval(y_nodes, yq)
    inside = (xq >= x_nodes[0]) & (xq <= x_nodes[-1]) & (yq >= y_nodes[0]) & (yq <= y_nodes[-1])
    interior = (xq > MARGIN) & (xq < Lx - MARGIN) & (yq > MARGIN) & (yq < Ly - MARGIN)
    cell = (ix * y_nodes.size + iy).astype(np.int32)
    return BilinearPlan(np.ascontiguousarray(cell), np.ascontiguousarray(tx), np.ascontiguousarray(ty),
                        np.ascontiguousarray((inside & interior).astype(np.int32)))


def batched_validation_errors(batch, stored: StoredSolution) -> np.ndarray:
    """[B, 3] = (u_L2_error, v_L2_error, n_valid) of every cavity of a BatchedSpectralLDC against one stored solution."""
    plan = bilinear_plan(batch.level.x, batch.level.y, stored.x, stored.y, batch.Lx, batch.Ly)
    return batch.validation_errors(plan, stored.u, stored.v)


def compute_validation_errors(batch, Re: float, root: str = ".") -> dict:
    """The reference's result dict for cavity 0 of `batch`: {u_L2_error, v_L2_error, u_L2_error_regu, v_L2_error_regu}
    for whichever of the two stored solutions exist under `root` (base.py:997-1049)."""
    out = {}
    for base, suffix in REF_DIRS:
        path = os.path.join(root, base, f"Re{int(Re)}", "solution.vts")
        if not os.path.exists(path):
            log.debug(f"No reference solution at {path}")
            continue
        stored = StoredSolution.from_vts(path)
        e = batched_validation_errors(batch, stored)[0]
        n_total = stored.u.size
        if e[2] < n_total * 0.5:
            log.warning(f"Only {int(e[2])}/{n_total} valid points for {base}")
        label = "FV-regu" if suffix else "FV"
        log.info(f"L2 errors vs {label} ({int(e[2])}/{n_total} pts): u={e[0]:.6e}, v={e[1]:.6e}")
        out[f"u_L2_error{suffix}"] = float(e[0])
        out[f"v_L2_error{suffix}"] = float(e[1])
    return out
