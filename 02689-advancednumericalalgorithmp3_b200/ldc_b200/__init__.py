"""ldc_b200: B200-native batched Chebyshev-collocation lid-driven-cavity pseudo-time stepper.

Host side of the C ABI in include/ldc_b200.h.  Importing the package does not need a GPU; creating a
BatchedSpectralLDC does (there is no CPU fallback).
"""
from . import operators
from ._lib import LIB_PATH, LdcError, load
from .batched import BatchedSpectralLDC, SolveResult, fp64_peak_tflops
from .fsg import FSGResult, solve_fsg_batched, solve_sg_batched
from .solvers import FSGSolver, SGSolver

__all__ = ["operators", "BatchedSpectralLDC", "SolveResult", "fp64_peak_tflops", "solve_fsg_batched",
           "solve_sg_batched", "FSGResult", "SGSolver", "FSGSolver", "LdcError", "LIB_PATH", "load"]
__version__ = "0.1.0"
