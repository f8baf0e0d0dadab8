"""Mint the to_vtk fixture from the UNMODIFIED reference.

    python tests/golden/make_golden_vtk.py     # authoring container only (needs /root/reference)

`LidDrivenCavitySolver.to_vtk` (base.py:464-522) builds a `pv.StructuredGrid(X, Y, Z)` and assigns point and field arrays;
PyVista is not installed here, so the stub module's StructuredGrid is a recorder (it keeps X, Y, Z and every array it is
handed).  Everything that decides the numbers -- the lexsort, the Fortran-order ravel, the spline "vorticity", the field
data -- is the reference's own code, run on the reference's own FSG Re=100 N=24 solution.  The point order PyVista gives
`StructuredGrid(X, Y, Z)` (Fortran-order ravel of the 2-D arrays, dimensions (ny, nx, 1)) is checked here against a file
the reference itself wrote through this very method: data/validation/fv/Re100/solution.vts.
`vtk.npz`: the solver arrays (u, v, p inner) and the recorded grid arrays.
"""
from __future__ import annotations

import os
import sys

import numpy as np

from make_golden import HERE, REF, fsg_kwargs, import_reference

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "02689-advancednumericalalgorithmp3_b200"))


class Recorder:
    def __init__(self, X, Y, Z):
        self.X, self.Y, self.Z = np.array(X), np.array(Y), np.array(Z)
        self.arrays, self.field_data = {}, {}

    def __setitem__(self, k, v):
        self.arrays[k] = np.array(v)


def main():
    from ldc_b200 import vts

    _, FSGSolver, *_ = import_reference()
    import pyvista as pv

    pv.StructuredGrid = Recorder
    FSGSolver.compute_vortex_metrics = lambda self: {}
    s = FSGSolver(**fsg_kwargs(24, 100.0))
    s.solve()
    g = s.to_vtk()
    n = s.params.nx + 1
    out = dict(u=s.arrays.u.reshape(n, n), v=s.arrays.v.reshape(n, n), p=s.arrays.p.reshape(n - 2, n - 2), X=g.X, Y=g.Y,
               Re=g.field_data["Re"], N=g.field_data["N"], solver=np.array(str(g.field_data["solver"][0])))
    for k, a in g.arrays.items():
        out["grid_" + k] = a
    np.savez_compressed(os.path.join(HERE, "vtk.npz"), **out)
    print({k: np.shape(a) for k, a in out.items()})

    # PyVista's point order, from a file written by this method: points = Fortran-order ravel of meshgrid(x, y)
    f = vts.read_vts(os.path.join(REF, "data/validation/fv/Re100/solution.vts"))
    x, y = np.unique(f.points[:, 0]), np.unique(f.points[:, 1])
    X, Y = np.meshgrid(x, y)
    assert f.dimensions == (len(y), len(x), 1)
    assert np.array_equal(f.points[:, 0], X.ravel("F")) and np.array_equal(f.points[:, 1], Y.ravel("F"))
    # ... and the stored "vorticity" is what base.py:524-549 computes from the stored u, v (see SGSolver.to_vtk)
    from scipy.interpolate import RectBivariateSpline

    U = f["u"].reshape(len(x), len(y)).T
    V = f["v"].reshape(len(x), len(y)).T
    w = RectBivariateSpline(y, x, V)(y, x, dx=1) - RectBivariateSpline(y, x, U)(y, x, dy=1)
    assert np.array_equal(w.ravel("F"), f["vorticity"])
    print("point order and vorticity convention confirmed against data/validation/fv/Re100/solution.vts")


if __name__ == "__main__":
    main()
