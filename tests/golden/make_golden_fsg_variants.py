"""Mint the FSG-variant fixture from the UNMODIFIED reference (conf/solver/spectral/fsg.yaml:13-17 knobs).

    python tests/golden/make_golden_fsg_variants.py     # authoring container only (needs /root/reference)

`fsg_variants.npz`:
  three   -- Re=100, N=48, n_levels=3 (levels [12, 24, 48]), coarse_tolerance_factor=10, polynomial prolongation;
  capped  -- Re=100, N=24, max_iterations=300: the cap applies per level, the solve ends unconverged (fsg.py:76-114).
"""
from __future__ import annotations

import numpy as np

from make_golden import fsg_kwargs, import_reference, save


def main():
    _, FSGSolver, *_ = import_reference()
    FSGSolver.compute_vortex_metrics = lambda self: {}
    out = {}
    for tag, kw in (("three", fsg_kwargs(48, 100.0, n_levels=3, coarse_tolerance_factor=10.0, prolongation_method="polynomial")),
                    ("capped", fsg_kwargs(24, 100.0, max_iterations=300))):
        f = FSGSolver(**kw)
        f.solve()
        n = kw["nx"] + 1
        ts = f.time_series
        out.update({f"{tag}_u": f.arrays.u.reshape(n, n), f"{tag}_v": f.arrays.v.reshape(n, n),
                    f"{tag}_p": f.arrays.p.reshape(n - 2, n - 2),
                    f"{tag}_meta": np.array([f.metrics.iterations, float(f.metrics.converged), f.metrics.final_residual,
                                             len(ts.rel_iter_residual), ts.rel_iter_residual[-1]])})
        print(tag, f.metrics.iterations, f.metrics.converged, f.metrics.final_residual, ts.rel_iter_residual)
    save("fsg_variants.npz", **out)


if __name__ == "__main__":
    main()
