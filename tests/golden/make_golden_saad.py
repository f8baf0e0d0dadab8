"""Mint the Saad-regularisation fixture from the UNMODIFIED reference (`corner_treatment: saad`,
conf/experiment/validation/saad-regu/spectral.yaml; SaadTreatment, spectral/operators/corner.py).

    python tests/golden/make_golden_saad.py     # authoring container only (needs /root/reference)

`saad.npz`: FSG solves at Re=1000 with the polynomial lid u = 16 x^2 (1-x)^2 -- N=16 (one level) and N=24 (levels [12, 24]).
"""
from __future__ import annotations

import numpy as np

from make_golden import fsg_kwargs, import_reference, save


def main():
    _, FSGSolver, *_ = import_reference()
    FSGSolver.compute_vortex_metrics = lambda self: {}
    out = {}
    for N in (16, 24):
        f = FSGSolver(**fsg_kwargs(N, 1000.0, corner_treatment="saad"))
        f.solve()
        n = N + 1
        out.update({f"u_{N}": f.arrays.u.reshape(n, n), f"v_{N}": f.arrays.v.reshape(n, n), f"p_{N}": f.arrays.p.reshape(n - 2, n - 2),
                    f"meta_{N}": np.array([f.metrics.iterations, float(f.metrics.converged)])})
        print("FSG saad N =", N, f.metrics.iterations, f.metrics.converged)
    save("saad.npz", **out)


if __name__ == "__main__":
    main()
