"""Drop-in solver classes for `solver=spectral/sg` and `solver=spectral/fsg`.

Same class names, constructor keywords (= the YAML keys of conf/solver/spectral/{sg,fsg}.yaml), public
attributes and error behaviour as the reference's `solvers.spectral.sg.SGSolver` (src/solvers/spectral/sg.py:29-101)
and `solvers.spectral.fsg.FSGSolver` (src/solvers/spectral/fsg.py:19-129), so `main.py` (run_solver,
main.py:75-120) can instantiate them through Hydra's `_target_` unchanged.  The numerics of `solve()` run in
the CUDA library through the C ABI (one cavity = a batch of one); nothing here falls back to the CPU.

Beyond the reference surface, both classes accept `device=` (CUDA device index, default 0) and `history=`
(default True: `SGSolver.solve` returns the reference's TimeSeries, base.py:243-313 thinned to 1000 points as base.py:137-141
does, forming only the records that survive the thinning -- two passes of the resident DMMA kernel, one history-path
iteration per kept record; "full": a record after every iteration, then thinned; False: the resident kernel runs the
solve once and the TimeSeries holds the final record only) -- both optional, so the reference's YAML files stay valid.
"""
from __future__ import annotations

import logging
import time

import numpy as np

from . import operators as ops
from .batched import BatchedSpectralLDC
from .datastructures import Fields, Metrics, SolverArrays, SpectralParameters, TimeSeries
from .fsg import solve_fsg_batched
from .postprocess import streamfunction_vortex_metrics

log = logging.getLogger(__name__)

_TRANSFER = {"prolongation": ("fft", "polynomial"), "restriction": ("fft", "injection")}


class SGSolver:
    """Single-grid pseudospectral artificial-compressibility solver (reference: spectral/sg.py:29-449)."""

    Parameters = SpectralParameters

    MAX_TIMESERIES_POINTS = 1000  # base.py:100 (_store_results default)

    def __init__(self, params=None, device: int = 0, history=True, **kwargs):
        if isinstance(history, str) and history.lower() not in ("full", "sampled"):
            raise ValueError("history must be True, False, 'sampled' or 'full'")
        self.history = "full" if isinstance(history, str) and history.lower() == "full" else bool(history)
        if params is None:
            params = self.Parameters(**kwargs)  # unknown key -> TypeError, as in the reference (base.py:47-52)
        self.params = params
        self.device = int(device)
        self.metrics = Metrics()
        self.time_series = None
        basis = str(params.basis_type).lower()
        if basis not in ops.BASES:
            raise ValueError(f"Unknown basis_type: {params.basis_type}. Use 'legendre' or 'chebyshev'")
        self.basis_type = basis
        # validates the corner method exactly where the reference does (corner.py:219-223)
        lv = ops.build_level(params.nx, params.ny, params.Lx, params.Ly, basis)
        self.level = lv
        ops.lid_velocity_profile(lv.x, params.corner_treatment, params.corner_smoothing, params.lid_velocity, params.Lx)

        # grids and operators under the reference's attribute names (sg.py:103-119, 181-210, 479-493)
        self.x_full, self.y_full = np.meshgrid(lv.x, lv.y, indexing="ij")
        self.x_inner, self.y_inner = lv.x[1:-1], lv.y[1:-1]
        self.x_reduced, self.y_reduced = np.meshgrid(self.x_inner, self.y_inner, indexing="ij")
        self.dx_min, self.dy_min = lv.dx_min, lv.dy_min
        self.Dx_1d, self.Dy_1d, self.Dxx_1d, self.Dyy_1d = lv.Dx, lv.Dy, lv.Dxx, lv.Dyy
        self.Interp_x, self.Interp_y = lv.Ix, lv.Iy
        self.w_x, self.w_y, self.W_2d = lv.wx, lv.wy, lv.W
        self.shape_full = (params.nx + 1, params.ny + 1)
        self.shape_inner = (params.nx - 1, params.ny - 1)
        self.arrays = SolverArrays.allocate(self.shape_full[0] * self.shape_full[1],
                                            self.shape_inner[0] * self.shape_inner[1])
        self.fields = Fields(u=np.zeros(self.x_full.size), v=np.zeros(self.x_full.size), p=np.zeros(self.x_full.size),
                             x=self.x_full.ravel().copy(), y=self.y_full.ravel().copy())
        self.u_2d = self.arrays.u.reshape(self.shape_full)
        self.v_2d = self.arrays.v.reshape(self.shape_full)
        self.p_2d = self.arrays.p.reshape(self.shape_inner)
        u_lid, v_lid = ops.lid_velocity_profile(lv.x, params.corner_treatment, params.corner_smoothing,
                                                params.lid_velocity, params.Lx)
        self.u_2d[:, -1] = u_lid  # _initialize_lid_velocity (sg.py:250-253)
        self.v_2d[:, -1] = v_lid
        self._gpu = None

    # ---- device handle ------------------------------------------------------------------------------------
    def _batch(self) -> BatchedSpectralLDC:
        if self._gpu is None:
            p = self.params
            self._gpu = BatchedSpectralLDC(p.nx, p.ny, 1, p.Lx, p.Ly, self.basis_type, self.device)
            self._gpu.set_cases(p.Re, p.beta_squared, p.CFL, p.lid_velocity, p.corner_treatment, p.corner_smoothing)
        return self._gpu

    def close(self):
        if self._gpu is not None:
            self._gpu.close()
            self._gpu = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _push_state(self):
        self._batch().set_state(self.u_2d[None], self.v_2d[None], self.p_2d[None])

    def _pull_state(self):
        u, v, p = self._batch().get_state()
        self.arrays.u[:] = u.ravel()
        self.arrays.v[:] = v.ravel()
        self.arrays.p[:] = p.ravel()

    def _pull_residuals(self):
        Ru, Rv, Rp, norms = self._batch().residual()
        self.arrays.R_u[:] = Ru.ravel()
        self.arrays.R_v[:] = Rv.ravel()
        self.arrays.R_p[:] = Rp.ravel()
        return norms[0]

    # ---- reference API --------------------------------------------------------------------------------------
    def step(self):
        """One RK4 pseudo-step with SGSolver semantics (frozen pressure gradient; sg.py:410-449)."""
        self._push_state()
        self._batch().step(1, freeze_pressure=True, want_rel=False)
        self._pull_state()
        return self.arrays.u, self.arrays.v, self.arrays.p

    def _compute_algebraic_residuals(self):
        return {"u_residual": float(np.linalg.norm(self.arrays.R_u)), "v_residual": float(np.linalg.norm(self.arrays.R_v)),
                "continuity_residual": float(np.linalg.norm(self.arrays.R_p))}

    def _extrapolate_to_full_grid(self, inner_2d):
        return ops.extrapolate_pressure_to_full(np.asarray(inner_2d))

    def _finalize_fields(self):
        """fields.p = inner pressure linearly extrapolated to the full grid (sg.py:451-461, 144-179)."""
        self.fields.u[:] = self.arrays.u
        self.fields.v[:] = self.arrays.v
        self.fields.p[:] = self._extrapolate_to_full_grid(self.p_2d).ravel()

    def _compute_energy(self):
        return float(self._diagnostics()[0])

    def _compute_enstrophy(self):
        return float(self._diagnostics()[1])

    def _compute_palinstrophy(self):
        return float(self._diagnostics()[2])

    def _diagnostics(self):
        self._push_state()
        return self._batch().diagnostics()[0]

    def compute_vortex_metrics(self) -> dict:
        """Streamfunction-based vortex metrics (sg.py:556-743) by tensor-product fast diagonalisation, on the device."""
        from .postprocess import vortex_metrics_device

        self._push_state()
        return vortex_metrics_device(self._batch())[0]

    def compute_validation_errors(self, reference_dir: str = "data/validation/fv", save_plots: bool = True) -> dict:
        """L2 errors against the stored FV solutions (base.py:970-1054) -- PyVista-free: the .vts files are parsed by
        `ldc_b200.vts` and the bilinear evaluation + norms run on the device.  As in the reference, `reference_dir` is
        accepted but the two directories data/validation/{fv,fv-regu} (relative to the working directory) are the ones
        read.  The error-distribution plots (matplotlib, base.py:1056-1120) are not produced here."""
        from .validation import compute_validation_errors

        self._push_state()
        return compute_validation_errors(self._batch(), self.params.Re)

    def validation_table(self, reference_csv: str = None):
        """The rows of the vortex-metrics comparison table of base.py:890-964 (computed value, Botella & Peyret value,
        error in per cent; primary, BL and BR vortices), or None when the reference file is missing.  As in the
        reference the file is data/validation/botella/botella_Re<int(Re)>_vortex.csv relative to the working directory,
        read with `comment="#"`, first data row; a column the file does not have gives a "-" row."""
        import os

        import pandas as pd

        if reference_csv is None:
            reference_csv = f"data/validation/botella/botella_Re{int(self.params.Re)}_vortex.csv"
        if not os.path.exists(reference_csv):
            log.warning(f"Reference file not found: {reference_csv}")
            return None
        ref = pd.read_csv(reference_csv, comment="#").iloc[0].to_dict()
        rows = []

        # Row format restated from the reference's own table builder (solvers/base.py:924-958): the column names, the
        # number formats and the `reference and reference != 0` rule ARE the contract of `mlflow.log_table`'s consumers.
        def add_row(vortex, metric, computed, reference, fmt=".6f"):
            if reference and reference != 0:
                error_pct = abs(abs(computed) - abs(reference)) / abs(reference) * 100
                ref_str = f"{reference:{fmt}}" if abs(reference) >= 1e-3 else f"{reference:.4e}"
            else:
                error_pct, ref_str = None, "-"
            comp_str = f"{computed:{fmt}}" if abs(computed) >= 1e-3 else f"{computed:.4e}"
            rows.append({"Vortex": vortex, "Metric": metric, "Computed": comp_str, "Botella": ref_str,
                         "Error (%)": f"{error_pct:.2f}" if error_pct is not None else "-"})

        m = self.metrics
        add_row("Primary", "|\u03c8|", abs(m.psi_min), ref.get("psi_primary"))
        add_row("Primary", "|\u03c9|", abs(m.omega_center), ref.get("omega_primary"))
        add_row("Primary", "x", m.psi_min_x, ref.get("x_primary"))
        add_row("Primary", "y", m.psi_min_y, ref.get("y_primary"))
        for tag in ("BL", "BR"):
            add_row(tag, "|\u03c8|", abs(getattr(m, f"psi_{tag}")), ref.get(f"psi_{tag}"))
            add_row(tag, "|\u03c9|", abs(getattr(m, f"omega_{tag}", 0.0)), ref.get(f"omega_{tag}"))
            add_row(tag, "x", getattr(m, f"psi_{tag}_x"), ref.get(f"x_{tag}"))
            add_row(tag, "y", getattr(m, f"psi_{tag}_y"), ref.get(f"y_{tag}"))
        return rows

    def mlflow_log_validation_table(self, reference_csv: str = None):
        """base.py:890-964: `validation_table` logged as validation_metrics.json into the active MLflow run; a warning
        and nothing else when MLflow is not installed or no run is active."""
        try:
            import mlflow
        except ImportError:
            log.warning("MLflow is not installed - skipping validation table")
            return
        if not mlflow.active_run():
            log.warning("No active MLflow run - skipping validation table")
            return
        rows = self.validation_table(reference_csv)
        if rows is None:
            return
        import pandas as pd

        mlflow.log_table(pd.DataFrame(rows), artifact_file="validation_metrics.json")
        log.info("Logged validation metrics table to MLflow")

    def compute_global_quantities(self) -> dict:
        """base.py:551-567: {'E': kinetic energy, 'Z': enstrophy, 'P': palinstrophy} of the current solution (one device call)."""
        E, Z, Pal = self._diagnostics()
        return {"E": float(E), "Z": float(Z), "P": float(Pal)}

    def save_vtk(self, filepath):
        """base.py:776-790: `to_vtk()` written to `filepath` (.vts), parent directories created."""
        import os

        filepath = os.fspath(filepath)
        os.makedirs(os.path.dirname(filepath) or ".", exist_ok=True)
        self.to_vtk().save(filepath)
        log.info(f"Saved VTS to {filepath}")

    def save(self, filepath):
        """base.py:332-353: params, metrics, time_series and fields as four tables of one HDF5 file (pandas.HDFStore, which
        needs PyTables -- as in the reference)."""
        import os

        import pandas as pd

        filepath = os.fspath(filepath)
        os.makedirs(os.path.dirname(filepath) or ".", exist_ok=True)
        with pd.HDFStore(filepath, mode="w", complevel=5) as store:
            store["params"] = self.params.to_dataframe()
            store["metrics"] = self.metrics.to_dataframe()
            store["time_series"] = self.time_series.to_dataframe()
            store["fields"] = self.fields.to_dataframe()

    def to_vtk(self):
        """base.py:464-522 without PyVista: the solution as a structured grid with u, v, pressure, velocity_magnitude,
        vorticity (the base-class default base.py:524-549, which the spectral solvers do not override: cubic splines built
        as Spline(y, x, F) and evaluated with dx=1 for "dvdx" and dy=1 for "dudy" -- the first spline variable being y,
        the array written as "vorticity" is dv/dy - du/dx; kept, the stored FV files carry exactly that), the velocity vector and the Re / N / solver field data; `to_vtk().save(path)` (main.py:116) writes
        the .vts flavour VTK writes (`ldc_b200.vts.write_vts`).  Point order and dimensions are PyVista's for
        `StructuredGrid(X, Y, Z)` with `X, Y = meshgrid(x, y)`: y runs fastest, dimensions (ny, nx, 1)."""
        from scipy.interpolate import RectBivariateSpline

        from .vts import StructuredGrid

        x_unique = np.sort(np.unique(self.fields.x))
        y_unique = np.sort(np.unique(self.fields.y))
        nx, ny = len(x_unique), len(y_unique)
        order = np.lexsort((self.fields.x, self.fields.y))
        U = self.fields.u[order].reshape(ny, nx)
        V = self.fields.v[order].reshape(ny, nx)
        Pr = self.fields.p[order].reshape(ny, nx)
        X, Y = np.meshgrid(x_unique, y_unique)
        pts = np.column_stack([X.ravel("F"), Y.ravel("F"), np.zeros(nx * ny)])
        dvdx = RectBivariateSpline(y_unique, x_unique, V)(y_unique, x_unique, dx=1)   # sic (base.py:546)
        dudy = RectBivariateSpline(y_unique, x_unique, U)(y_unique, x_unique, dy=1)   # sic (base.py:547)
        grid = StructuredGrid(points=pts, dimensions=(ny, nx, 1))
        grid["u"] = U.ravel("F")
        grid["v"] = V.ravel("F")
        grid["pressure"] = Pr.ravel("F")
        grid["velocity_magnitude"] = np.sqrt(U**2 + V**2).ravel("F")
        grid["vorticity"] = (dvdx - dudy).ravel("F")
        grid["velocity"] = np.column_stack([U.ravel("F"), V.ravel("F"), np.zeros(nx * ny)])
        grid.field_data["Re"] = np.array([self.params.Re])
        grid.field_data["N"] = np.array([self.params.nx])
        grid.field_data["solver"] = str(self.params.name)
        return grid

    def solve(self, tolerance: float = None, max_iter: int = None):
        """base.solve (base.py:202-330) for the spectral single grid: iterate SGSolver.step until
        max(||du||/(||u||+1e-12), same for v) < tolerance after a 10-iteration warm-up."""
        tol = self.params.tolerance if tolerance is None else tolerance
        max_iter = self.params.max_iterations if max_iter is None else max_iter
        g = self._batch()
        t0 = time.time()
        self._push_state()
        hist = None
        if self.history == "full":
            r, hists = g.solve_history(tol, int(max_iter), warmup=10, freeze_pressure=True, nan_exit=True)
            hist = hists[0][10:]  # base.py:262: records are kept for i >= 10 only
        elif self.history:
            # only the records that survive the reference's thinning to MAX_TIMESERIES_POINTS are formed
            r, hists, _ = g.solve_history_sampled(tol, int(max_iter), warmup=10, freeze_pressure=True, nan_exit=True,
                                                  skip=10, max_points=self.MAX_TIMESERIES_POINTS)
            hist = hists[0]
        else:
            r = g.solve(tol, int(max_iter), warmup=10, freeze_pressure=True, nan_exit=True)
        diverged = bool(r.diverged[0])
        # the reference has no NaN exit on this path: it keeps stepping NaNs until max_iter
        iters = int(max_iter) if diverged else int(r.iterations[0])
        self._pull_state()
        wall = time.time() - t0
        norms = self._pull_residuals()
        E, Z, Pal = g.diagnostics()[0]
        self._store_results(float(r.rel_residual[0]), iters, bool(r.converged[0]), wall, norms, (E, Z, Pal), hist)

    @classmethod
    def _downsample(cls, col):
        """base.py:137-141: at most MAX_TIMESERIES_POINTS entries, picked at linspace(0, L-1, ., dtype=int)."""
        col = [float(x) for x in col]
        if len(col) <= cls.MAX_TIMESERIES_POINTS:
            return col
        idx = np.linspace(0, len(col) - 1, cls.MAX_TIMESERIES_POINTS, dtype=int)
        return [col[i] for i in idx]

    def _store_results(self, rel, iters, converged, wall, norms, ezp, hist=None):
        self._finalize_fields()
        if hist is not None:
            # base.py:125-153: the full per-iteration history, down-sampled; Metrics take the FINAL (not sampled) values
            cols = [self._downsample(hist[:, k]) for k in range(7)]
            self.time_series = TimeSeries(*cols)
            if len(hist):
                rel, norms, ezp = float(hist[-1, 0]), hist[-1, 1:4], hist[-1, 4:7]
            else:
                rel, norms, ezp = float("inf"), (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)
        else:
            self.time_series = TimeSeries(rel_iter_residual=[rel], u_residual=[float(norms[0])],
                                          v_residual=[float(norms[1])], continuity_residual=[float(norms[2])],
                                          energy=[float(ezp[0])], enstrophy=[float(ezp[1])], palinstrophy=[float(ezp[2])])
        try:
            vm = self.compute_vortex_metrics()
        except Exception as e:  # base.py:156-160 swallows these too
            log.warning(f"Failed to compute vortex metrics: {e}")
            vm = {}
        self.metrics = Metrics(iterations=iters, converged=converged, final_residual=rel, wall_time_seconds=wall,
                               u_momentum_residual=float(norms[0]), v_momentum_residual=float(norms[1]),
                               continuity_residual=float(norms[2]), final_energy=float(ezp[0]),
                               final_enstrophy=float(ezp[1]), final_palinstrophy=float(ezp[2]),
                               **{k: float(v) for k, v in vm.items() if k in Metrics.__dataclass_fields__})


class FSGSolver(SGSolver):
    """Coarse-to-fine grid sequencing (reference: spectral/fsg.py:19-129, multigrid/fsg.py:1053-1221)."""

    def solve(self, tolerance: float = None, max_iter: int = None):
        p = self.params
        tol = p.tolerance if tolerance is None else tolerance
        max_iter = p.max_iterations if max_iter is None else max_iter
        if p.prolongation_method not in _TRANSFER["prolongation"]:
            raise ValueError(f"Unknown prolongation method: {p.prolongation_method}. Use 'fft' or 'polynomial'.")
        if p.restriction_method not in _TRANSFER["restriction"]:
            raise ValueError(f"Unknown restriction method: {p.restriction_method}. Use 'fft' or 'injection'.")
        t0 = time.time()
        out = solve_fsg_batched(p.nx, p.Re, tol, int(max_iter), n_levels=p.n_levels,
                                coarse_tolerance_factor=p.coarse_tolerance_factor, prolongation=p.prolongation_method,
                                beta_squared=p.beta_squared, CFL=p.CFL, lid_velocity=p.lid_velocity,
                                corner_treatment=p.corner_treatment, corner_smoothing=p.corner_smoothing, Lx=p.Lx, Ly=p.Ly,
                                basis=self.basis_type, device=self.device)
        wall = time.time() - t0
        self.arrays.u[:] = out.u[0].ravel()
        self.arrays.v[:] = out.v[0].ravel()
        self.arrays.p[:] = out.p[0].ravel()
        converged = bool(out.converged[0])
        self._push_state()
        norms = self._pull_residuals()
        ezp = self._batch().diagnostics()[0]
        # fsg.py:107-114: a one-entry history whose rel_iter is tol (converged) or 10 tol (not)
        self._store_results(tol if converged else tol * 10, int(out.iterations[0]), converged, wall, norms, ezp)
