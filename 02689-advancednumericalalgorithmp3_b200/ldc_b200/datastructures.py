"""Config / result containers of the solver interface, mirroring the reference's field names so that the
caller (`main.py`: params.to_mlflow(), metrics.to_mlflow(), time_series.to_mlflow_batch(), fields.*) works
unchanged.  Reference: src/solvers/datastructures.py:29-143 (Parameters, Metrics, TimeSeries, Fields) and
:257-279 (SpectralParameters).  Only names, defaults and conversions are shared; the solver arrays of the
reference (SpectralSolverFields, 21 work buffers) do not exist here: the GPU keeps u, v, p only.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import numpy as np


def _mlflow_value(v):
    return int(v) if isinstance(v, bool) else v


@dataclass
class Parameters:
    """Base solver configuration (datastructures.py:29-51)."""

    name: str = ""
    Re: float = 100
    lid_velocity: float = 1.0
    Lx: float = 1.0
    Ly: float = 1.0
    nx: int = 64
    ny: int = 64
    max_iterations: int = 500
    tolerance: float = 1e-4
    method: str = ""

    def to_mlflow(self) -> dict:
        return {k: _mlflow_value(v) for k, v in self.__dict__.items()}

    def to_dataframe(self):
        import pandas as pd

        return pd.DataFrame([self.to_mlflow()])


@dataclass
class SpectralParameters(Parameters):
    """Spectral configuration: exactly the keys of conf/solver/spectral/{sg,fsg}.yaml
    (datastructures.py:257-279).  An unknown key raises TypeError, as in the reference."""

    basis_type: str = "legendre"
    CFL: float = 0.1
    beta_squared: float = 5.0
    method: str = "Spectral-AC"
    corner_treatment: str = "smoothing"
    corner_smoothing: float = 0.15
    multigrid: str = "none"
    n_levels: int = 3
    coarse_tolerance_factor: float = 10.0
    prolongation_method: str = "fft"
    restriction_method: str = "fft"


@dataclass
class Metrics:
    """Solver outputs (datastructures.py:59-109)."""

    iterations: int = 0
    converged: bool = False
    final_residual: float = float("inf")
    wall_time_seconds: float = 0.0
    u_momentum_residual: float = 0.0
    v_momentum_residual: float = 0.0
    continuity_residual: float = 0.0
    final_energy: float = 0.0
    final_enstrophy: float = 0.0
    final_palinstrophy: float = 0.0
    psi_min: float = 0.0
    psi_min_x: float = 0.0
    psi_min_y: float = 0.0
    omega_center: float = 0.0
    omega_max: float = 0.0
    omega_max_x: float = 0.0
    omega_max_y: float = 0.0
    psi_BR: float = 0.0
    omega_BR: float = 0.0
    psi_BR_x: float = 0.0
    psi_BR_y: float = 0.0
    psi_BL: float = 0.0
    omega_BL: float = 0.0
    psi_BL_x: float = 0.0
    psi_BL_y: float = 0.0
    psi_TL: float = 0.0
    omega_TL: float = 0.0
    psi_TL_x: float = 0.0
    psi_TL_y: float = 0.0

    def to_mlflow(self) -> dict:
        return {k: _mlflow_value(v) for k, v in self.__dict__.items() if v != float("inf")}

    def to_dataframe(self):
        import pandas as pd

        return pd.DataFrame([self.to_mlflow()])


@dataclass
class TimeSeries:
    """Convergence history (datastructures.py:117-143)."""

    rel_iter_residual: List[float] = field(default_factory=list)
    u_residual: List[float] = field(default_factory=list)
    v_residual: List[float] = field(default_factory=list)
    continuity_residual: List[float] = field(default_factory=list)
    energy: List[float] = field(default_factory=list)
    enstrophy: List[float] = field(default_factory=list)
    palinstrophy: List[float] = field(default_factory=list)

    def to_mlflow_batch(self) -> list:
        from mlflow.entities import Metric  # only where MLflow is installed (main.py:106-109)

        return [Metric(key=name, value=value, timestamp=0, step=step)
                for name, values in self.__dict__.items() if values
                for step, value in enumerate(values) if value is not None]

    def to_dataframe(self):
        import pandas as pd

        return pd.DataFrame({k: v for k, v in self.__dict__.items() if v})


@dataclass
class Fields:
    """Flattened solution fields on the full grid (datastructures.py:151-166)."""

    u: np.ndarray
    v: np.ndarray
    p: np.ndarray
    x: np.ndarray
    y: np.ndarray

    def to_dataframe(self):
        import pandas as pd

        return pd.DataFrame({"x": self.x, "y": self.y, "u": self.u, "v": self.v, "p": self.p})


@dataclass
class SolverArrays:
    """The subset of the reference's SpectralSolverFields that its inherited machinery reads after a solve
    (base.py:96-98, 227-280; sg.py:463-473): flattened u, v (full grid), p (inner grid) and the residuals."""

    u: np.ndarray
    v: np.ndarray
    p: np.ndarray
    R_u: np.ndarray
    R_v: np.ndarray
    R_p: np.ndarray

    @classmethod
    def allocate(cls, n_nodes_full: int, n_nodes_inner: int):
        z = np.zeros
        return cls(z(n_nodes_full), z(n_nodes_full), z(n_nodes_inner), z(n_nodes_full), z(n_nodes_full), z(n_nodes_inner))
