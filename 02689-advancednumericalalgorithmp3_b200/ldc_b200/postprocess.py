"""Post-solve quantities of the spectral solver: streamfunction and vortex metrics.

The reference solves lap(psi) = -omega with psi = 0 on the walls by assembling the sparse Kronecker Laplacian
and calling SuperLU (spectral/sg.py:556-619; 4.65 s per call at N=64, twice per solve).  The same linear system
restricted to the interior nodes is the Sylvester equation  A psi + psi B^T = -omega_int  with
A = Dxx[1:-1,1:-1], B = Dyy[1:-1,1:-1]; here it is solved by tensor-product fast diagonalisation:
two eigendecompositions per polynomial order (cached) and four m x m GEMMs per cavity.
"""
from __future__ import annotations

from functools import lru_cache

import numpy as np

from . import operators as ops


@lru_cache(maxsize=32)
def _diagonalisation(nx: int, ny: int, Lx: float, Ly: float, basis: str):
    lv = ops.build_level(nx, ny, Lx, Ly, basis)

    def eig(M):
        lam, V = np.linalg.eig(M)
        if np.abs(lam.imag).max() > 1e-8 * np.abs(lam.real).max():
            raise np.linalg.LinAlgError("second-derivative block has complex eigenvalues")
        return lam.real, V.real, np.linalg.inv(V.real)

    la, Va, Vai = eig(lv.Dxx[1:-1, 1:-1])
    if nx == ny and Lx == Ly:
        lb, Vb, Vbi = la, Va, Vai
    else:
        lb, Vb, Vbi = eig(lv.Dyy[1:-1, 1:-1])
    return la, Va, Vai, lb, Vb, Vbi


VORTEX_KEYS = ("psi_min", "psi_min_x", "psi_min_y", "omega_center", "omega_max", "omega_max_x", "omega_max_y",
               "psi_BR", "omega_BR", "psi_BR_x", "psi_BR_y", "psi_BL", "omega_BL", "psi_BL_x", "psi_BL_y",
               "psi_TL", "omega_TL", "psi_TL_x", "psi_TL_y")


def diagonalisation_operands(level):
    """The shared operands of `ldc_vortex_metrics` for this level: (Vai, VbiT, Va, VbT, inv_lam, x, y), C-contiguous."""
    la, Va, Vai, lb, Vb, Vbi = _diagonalisation(level.nx, level.ny, level.Lx, level.Ly, level.basis)
    c = np.ascontiguousarray
    return (c(Vai), c(Vbi.T), c(Va), c(Vb.T), c(1.0 / (la[:, None] + lb[None, :])), c(level.x, dtype=np.float64),
            c(level.y, dtype=np.float64))


def vortex_metrics_device(batch, want_psi: bool = False):
    """The 19 vortex entries of `Metrics` for EVERY cavity of a `BatchedSpectralLDC` handle, computed on the device from
    the state it holds (`ldc_vortex_metrics`: batched fast-diagonalisation solve + one fixed-order scan per cavity).
    Returns a list of dicts (and psi [B, n, n] when asked).  Same quantities as `streamfunction_vortex_metrics`."""
    vals, psi = batch.vortex_metrics(diagonalisation_operands(batch.level), want_psi)
    out = [{k: float(v) for k, v in zip(VORTEX_KEYS, row)} for row in vals]
    return (out, psi) if want_psi else out


def vorticity(level, u, v):
    """omega = dv/dx - du/dy on all nodes (spectral/sg.py:510-522); u, v are [..., n, n]."""
    return level.Dx @ v - u @ level.Dy.T


def streamfunction(level, u, v):
    """psi [..., n, n] with psi = 0 on the boundary (spectral/sg.py:556-619)."""
    om = vorticity(level, u, v)
    la, Va, Vai, lb, Vb, Vbi = _diagonalisation(level.nx, level.ny, level.Lx, level.Ly, level.basis)
    F = -om[..., 1:-1, 1:-1]
    G = Vai @ F @ Vbi.T
    G = G / (la[:, None] + lb[None, :])
    psi = np.zeros_like(om)
    psi[..., 1:-1, 1:-1] = Va @ G @ Vb.T
    # one step of iterative refinement against the (non-normal) interior operator
    A, B = level.Dxx[1:-1, 1:-1], level.Dyy[1:-1, 1:-1]
    R = F - (A @ psi[..., 1:-1, 1:-1] + psi[..., 1:-1, 1:-1] @ B.T)
    psi[..., 1:-1, 1:-1] += Va @ ((Vai @ R @ Vbi.T) / (la[:, None] + lb[None, :])) @ Vb.T
    return psi, om


def streamfunction_vortex_metrics(level, u, v) -> dict:
    """The 19 vortex entries of `Metrics` for one cavity (spectral/sg.py:621-743): primary vortex = global
    minimum of psi; BR/BL/TL secondary vortices = maximum of psi in the corner quadrant if positive, else
    zeros; omega_max = vorticity of largest magnitude (signed)."""
    u = np.asarray(u, dtype=np.float64).reshape(level.shape_full)
    v = np.asarray(v, dtype=np.float64).reshape(level.shape_full)
    psi, om = streamfunction(level, u, v)
    X, Y = np.meshgrid(level.x, level.y, indexing="ij")
    i0 = np.unravel_index(np.argmin(psi), psi.shape)
    im = np.unravel_index(np.argmax(np.abs(om)), om.shape)
    out = {"psi_min": float(psi[i0]), "psi_min_x": float(X[i0]), "psi_min_y": float(Y[i0]), "omega_center": float(om[i0]),
           "omega_max": float(om[im]), "omega_max_x": float(X[im]), "omega_max_y": float(Y[im])}
    regions = {"BR": (X > 0.5) & (Y < 0.5), "BL": (X < 0.5) & (Y < 0.5), "TL": (X < 0.5) & (Y > 0.5)}
    for name, mask in regions.items():
        k = np.unravel_index(np.argmax(np.where(mask, psi, -np.inf)), psi.shape)
        if psi[k] > 0:
            vals = (float(psi[k]), float(om[k]), float(X[k]), float(Y[k]))
        else:
            vals = (0.0, 0.0, 0.0, 0.0)
        out[f"psi_{name}"], out[f"omega_{name}"], out[f"psi_{name}_x"], out[f"psi_{name}_y"] = vals
    return out
