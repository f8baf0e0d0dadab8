"""Mint the validation-error fixture from the UNMODIFIED reference.

    python tests/golden/make_golden_validation.py     # authoring container only (needs /root/reference)

The reference's `compute_validation_errors` (base.py:970-1054) reads the stored FV solution with `pv.read`; PyVista is
not installed here, so the stub module's `read` is served by this repo's .vts parser (`ldc_b200.vts`, itself checked
against the RangeMin/RangeMax attributes VTK wrote into the files).  Everything after the read -- the bilinear
evaluation with SciPy's RegularGridInterpolator, the interior mask, the norms -- is the reference's own code.
`validation.npz`: the FSG Re=100 N=24 fields, the stored FV Re=100 point cloud (x, y, u, v) and the reference's errors.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

from make_golden import HERE, REF, fsg_kwargs, import_reference

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "02689-advancednumericalalgorithmp3_b200"))


def main():
    from ldc_b200 import vts

    _, FSGSolver, *_ = import_reference()
    import pyvista as pv  # the stub installed by import_reference

    def read(path):
        g = vts.read_vts(path)
        return types.SimpleNamespace(points=g.points, point_data=g.point_data)

    pv.read = read
    FSGSolver.compute_vortex_metrics = lambda self: {}
    os.chdir(REF)  # the reference resolves data/validation/... against the working directory
    s = FSGSolver(**fsg_kwargs(24, 100.0))
    s.solve()
    errs = s.compute_validation_errors(save_plots=False)
    print(errs)
    g = vts.read_vts(os.path.join(REF, "data/validation/fv/Re100/solution.vts"))
    n = s.params.nx + 1
    np.savez_compressed(os.path.join(HERE, "validation.npz"), u=s.arrays.u.reshape(n, n), v=s.arrays.v.reshape(n, n),
                        x=g.points[:, 0], y=g.points[:, 1], ref_u=g.point_data["u"], ref_v=g.point_data["v"],
                        errors=np.array([errs["u_L2_error"], errs["v_L2_error"]]), keys=np.array(sorted(errs)))


if __name__ == "__main__":
    main()
