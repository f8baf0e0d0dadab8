"""Mint the TimeSeries fixtures from the UNMODIFIED reference.

    python tests/golden/make_golden_history.py        # authoring container only (needs /root/reference), ~1 min

`history.npz` holds, for SGSolver Re=100 N=16 cs=0.15 (14 155 iterations):
  short_*   the raw per-iteration series of a 60-iteration run (30 records, not down-sampled): base.py:262-276
  full_*    the 1000-point down-sampled series of the converged run (base.py:137-153) and the final Metrics
"""
from __future__ import annotations

import os

import numpy as np

from make_golden import HERE, import_reference, sg_kwargs

KEYS = ("rel_iter_residual", "u_residual", "v_residual", "continuity_residual", "energy", "enstrophy", "palinstrophy")


def series(ts):
    return np.array([getattr(ts, k) for k in KEYS]).T


def main():
    SGSolver, *_ = import_reference()
    SGSolver.compute_vortex_metrics = lambda self: {}  # not on the path under test
    out = {}
    s = SGSolver(**sg_kwargs(16, 100.0))
    s.solve(max_iter=40)
    out["short_series"] = series(s.time_series)
    out["short_iterations"] = np.array(s.metrics.iterations)
    s = SGSolver(**sg_kwargs(16, 100.0))
    s.solve()
    mt = s.metrics
    out["full_series"] = series(s.time_series)
    out["full_meta"] = np.array([mt.iterations, float(mt.converged), mt.final_residual, mt.u_momentum_residual,
                                 mt.v_momentum_residual, mt.continuity_residual, mt.final_energy, mt.final_enstrophy,
                                 mt.final_palinstrophy])
    print("short", out["short_series"].shape, "full", out["full_series"].shape, "iterations", mt.iterations)
    np.savez_compressed(os.path.join(HERE, "history.npz"), **out)


if __name__ == "__main__":
    main()
