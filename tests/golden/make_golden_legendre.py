"""Mint the Legendre-basis fixture from the UNMODIFIED reference (`basis_type: legendre`, conf/solver/spectral/sg.yaml:19;
LegendreLobattoBasis, spectral/basis/spectral.py:326-408).

    python tests/golden/make_golden_legendre.py     # authoring container only (needs /root/reference)

`legendre.npz`: operators for an even and an odd order, a 10-step SG trajectory, a full SG solve (Re=100, N=16) and a full
FSG solve (Re=100, N=24, levels [12, 24]; the transfer operators keep their Chebyshev formulas whatever the basis, as in
the reference).
"""
from __future__ import annotations

import numpy as np

from make_golden import fsg_kwargs, import_reference, save, sg_kwargs


def main():
    SGSolver, FSGSolver, *_ = import_reference()
    FSGSolver.compute_vortex_metrics = lambda self: {}
    SGSolver.compute_vortex_metrics = lambda self: {}
    out = {}
    for N in (16, 21):
        s = SGSolver(**sg_kwargs(N, 100.0, basis_type="legendre"))
        out.update({f"x_{N}": s.x_full[:, 0], f"Dx_{N}": s.Dx_1d, f"Dxx_{N}": s.Dxx_1d, f"Ix_{N}": s.Interp_x, f"W_{N}": s.W_2d,
                    f"dxmin_{N}": np.array(s.dx_min)})
    s = SGSolver(**sg_kwargs(21, 400.0, basis_type="legendre"))
    snaps = []
    for k in range(10):
        s.step()
        if k in (0, 4, 9):
            snaps.append(np.concatenate([s.arrays.u, s.arrays.v, s.arrays.p]))
    out["traj_21"] = np.array(snaps)
    s = SGSolver(**sg_kwargs(16, 100.0, basis_type="legendre"))
    s.solve()
    out.update(sg_u=s.arrays.u.reshape(17, 17), sg_v=s.arrays.v.reshape(17, 17), sg_p=s.arrays.p.reshape(15, 15),
               sg_meta=np.array([s.metrics.iterations, float(s.metrics.converged), s.metrics.final_residual,
                                 s.metrics.final_energy, s.metrics.final_enstrophy, s.metrics.final_palinstrophy]))
    print("SG legendre N=16:", s.metrics.iterations, s.metrics.converged)
    f = FSGSolver(**fsg_kwargs(24, 100.0, basis_type="legendre"))
    f.solve()
    out.update(fsg_u=f.arrays.u.reshape(25, 25), fsg_v=f.arrays.v.reshape(25, 25), fsg_p=f.arrays.p.reshape(23, 23),
               fsg_meta=np.array([f.metrics.iterations, float(f.metrics.converged)]))
    print("FSG legendre N=24:", f.metrics.iterations, f.metrics.converged)
    save("legendre.npz", **out)


if __name__ == "__main__":
    main()
