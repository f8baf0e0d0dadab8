"""Golden vortex metrics: the UNMODIFIED reference's `compute_vortex_metrics` (sparse SuperLU streamfunction,
spectral/sg.py:556-743) applied to the converged fields already stored in solves.npz.  Authoring container only.

    python tests/golden/make_golden_vortex.py
"""
import os

import numpy as np

from make_golden import HERE, fsg_kwargs, import_reference, save, sg_kwargs

SGSolver, FSGSolver, *_ = import_reference()
G = np.load(os.path.join(HERE, "solves.npz"))
out = {}
for tag, N, Re, kw in (("sg_Re100_N16", 16, 100.0, sg_kwargs), ("fsg_Re100_N24", 24, 100.0, fsg_kwargs),
                       ("fsg_Re100_N32", 32, 100.0, fsg_kwargs), ("fsg_Re1000_N30_cs015", 30, 1000.0, fsg_kwargs),
                       ("fsg_Re400_N64", 64, 400.0, fsg_kwargs)):
    s = SGSolver(**sg_kwargs(N, Re))
    s.arrays.u[:] = G[f"{tag}_u"].ravel()
    s.arrays.v[:] = G[f"{tag}_v"].ravel()
    vm = s.compute_vortex_metrics()
    keys = sorted(vm)
    out[f"{tag}_keys"] = np.array(keys)
    out[f"{tag}_vals"] = np.array([vm[k] for k in keys])
    psi, _, _ = s._compute_streamfunction()
    out[f"{tag}_psi"] = psi
    print(tag, vm["psi_min"], vm["psi_min_x"], vm["psi_min_y"], vm["psi_BR"], vm["omega_max"])
save("vortex.npz", **out)
