"""Import the UNMODIFIED reference (test / measurement infrastructure only -- never imported by the product package).

Looks for the staged copy `oracle/_ref/src` first (made by `oracle/build_ref.sh`; it travels to the GPU box), then for
`/root/reference/src` (authoring container).  `pyvista`, `mlflow` and `pyamg` are imported at module top by the
reference (`solvers/base.py:10-11`, `solvers/fv/linear_solvers/scipy_solver.py:4`) but are neither installed here nor
on the spectral stepping path: empty stub modules stand in for them.
"""
from __future__ import annotations

import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
CANDIDATES = (os.path.join(HERE, "_ref", "src"), os.path.join(os.environ.get("LDC_REFERENCE", "/root/reference"), "src"))


def reference_src() -> str | None:
    for c in CANDIDATES:
        if os.path.isdir(os.path.join(c, "solvers")):
            return c
    return None


def why_unavailable() -> str | None:
    """None when the reference can be imported here, else a one-line reason."""
    if reference_src() is None:
        return "no oracle/_ref (run oracle/build_ref.sh where /root/reference exists)"
    try:
        import numba  # noqa: F401
    except Exception as e:  # pragma: no cover
        return f"numba does not import: {type(e).__name__}: {e}"
    return None


def load_reference():
    """-> namespace with SGSolver, FSGSolver, LidDrivenCavitySolver, mg (multigrid.fsg), tr (transfer operators),
    create_corner_treatment.  Raises ImportError with the reason when the reference is not available."""
    why = why_unavailable()
    if why:
        raise ImportError(why)
    os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/numba_cache_ldc")
    pv = types.ModuleType("pyvista")
    pv.StructuredGrid = object
    ml = types.ModuleType("mlflow")
    ml.active_run = lambda: None
    ml.log_metrics = lambda *a, **k: None
    sys.modules.setdefault("pyvista", pv)
    sys.modules.setdefault("mlflow", ml)
    sys.modules.setdefault("pyamg", types.ModuleType("pyamg"))
    src = reference_src()
    if src not in sys.path:
        sys.path.insert(0, src)
    import solvers  # noqa: F401
    from solvers.base import LidDrivenCavitySolver
    from solvers.spectral.fsg import FSGSolver
    from solvers.spectral.multigrid import fsg as mg
    from solvers.spectral.operators import transfer_operators as tr
    from solvers.spectral.operators.corner import create_corner_treatment
    from solvers.spectral.sg import SGSolver

    return types.SimpleNamespace(SGSolver=SGSolver, FSGSolver=FSGSolver, LidDrivenCavitySolver=LidDrivenCavitySolver, mg=mg,
                                 tr=tr, create_corner_treatment=create_corner_treatment, src=src)


def make_smoother(ref, N: int, Re: float, corner_smoothing: float = 0.15, beta_squared: float = 5.0, CFL: float = 1.5,
                  lid_velocity: float = 1.0, lean: bool = True):
    """A reference `MultigridSmoother` (`multigrid/fsg.py:744-995`) on a cold-start level of order N.

    lean=True builds the level by calling the reference's own pieces (`basis.nodes`, `basis.diff_matrix`,
    `_build_interpolation_matrix_1d`) and leaves out the n^2 x n^2 Kronecker matrices of `build_spectral_level`
    (`multigrid/fsg.py:410-424`), which the stepping path never reads and which take 11 GB at N = 128."""
    import numpy as np

    from solvers.spectral.basis.spectral import ChebyshevLobattoBasis

    mg = ref.mg
    bx = ChebyshevLobattoBasis(domain=(0.0, 1.0))
    by = ChebyshevLobattoBasis(domain=(0.0, 1.0))
    if not lean:
        lv = mg.build_spectral_level(N, 0, bx, by, 1.0, 1.0)
    else:
        n, m = N + 1, N - 1
        x = bx.nodes(n)
        y = by.nodes(n)
        X, Y = np.meshgrid(x, y, indexing="ij")
        Xi, Yi = np.meshgrid(x[1:-1], y[1:-1], indexing="ij")
        Dx = bx.diff_matrix(x)
        Dy = by.diff_matrix(y)
        z = np.zeros((1, 1))
        zf, zi = (lambda: np.zeros(n * n)), (lambda: np.zeros(m * m))
        lv = mg.SpectralLevel(
            n=N, level_idx=0, x_nodes=x, y_nodes=y, X=X, Y=Y, X_inner=Xi, Y_inner=Yi, shape_full=(n, n), shape_inner=(m, m),
            dx_min=np.min(np.diff(x)), dy_min=np.min(np.diff(y)), Dx_1d=Dx, Dy_1d=Dy, Dxx_1d=Dx @ Dx, Dyy_1d=Dy @ Dy,
            Dx=z, Dy=z, Dxx=z, Dyy=z, Laplacian=z, Dx_inner=z, Dy_inner=z,
            Interp_x=mg._build_interpolation_matrix_1d(x[1:-1], x), Interp_y=mg._build_interpolation_matrix_1d(y[1:-1], y),
            u=zf(), v=zf(), p=zi(), u_prev=zf(), v_prev=zf(), u_stage=zf(), v_stage=zf(), p_stage=zi(), R_u=zf(), R_v=zf(),
            R_p=zi(), du_dx=zf(), du_dy=zf(), dv_dx=zf(), dv_dy=zf(), lap_u=zf(), lap_v=zf(), dp_dx=zf(), dp_dy=zf(),
            dp_dx_inner=zi(), dp_dy_inner=zi())
    sm = mg.MultigridSmoother(lv, Re=Re, beta_squared=beta_squared, lid_velocity=lid_velocity, CFL=CFL,
                              corner_treatment=ref.create_corner_treatment("smoothing", corner_smoothing))
    sm.initialize_lid()
    return sm
