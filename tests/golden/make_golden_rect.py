"""Mint the rectangular-cavity fixture from the UNMODIFIED reference (Lx != Ly and nx != ny: `Lx`, `Ly`, `nx`, `ny` of
conf/config.yaml / SpectralParameters).

    python tests/golden/make_golden_rect.py     # authoring container only (needs /root/reference)

`rect.npz`: SG nx=20, ny=14 on [0,2] x [0,0.5] (x and y operators differ in size: the launch-per-stage path), a 10-step
trajectory and the solve to 1e-6; FSG nx=ny=24 on [0,2] x [0,1] (same size, different scaling: the plain resident kernel
with separate x / y operators), solve to 1e-6.
"""
from __future__ import annotations

import numpy as np

from make_golden import fsg_kwargs, import_reference, save, sg_kwargs


def main():
    SGSolver, FSGSolver, *_ = import_reference()
    FSGSolver.compute_vortex_metrics = lambda self: {}
    SGSolver.compute_vortex_metrics = lambda self: {}
    out = {}
    kw = sg_kwargs(20, 100.0, Lx=2.0, Ly=0.5)
    kw["ny"] = 14
    s = SGSolver(**kw)
    snaps = []
    for k in range(10):
        s.step()
        if k in (0, 4, 9):
            snaps.append(np.concatenate([s.arrays.u, s.arrays.v, s.arrays.p]))
    out["sg_traj"] = np.array(snaps)
    s = SGSolver(**kw)
    s.solve()
    out.update(sg_u=s.arrays.u.reshape(21, 15), sg_v=s.arrays.v.reshape(21, 15), sg_p=s.arrays.p.reshape(19, 13),
               sg_meta=np.array([s.metrics.iterations, float(s.metrics.converged), s.metrics.final_energy,
                                 s.metrics.final_enstrophy, s.metrics.final_palinstrophy]))
    print("SG 20x14 on 2 x 0.5:", s.metrics.iterations, s.metrics.converged)
    f = FSGSolver(**fsg_kwargs(24, 100.0, Lx=2.0, Ly=1.0))
    f.solve()
    out.update(fsg_u=f.arrays.u.reshape(25, 25), fsg_v=f.arrays.v.reshape(25, 25), fsg_p=f.arrays.p.reshape(23, 23),
               fsg_meta=np.array([f.metrics.iterations, float(f.metrics.converged)]))
    print("FSG 24x24 on 2 x 1:", f.metrics.iterations, f.metrics.converged)
    save("rect.npz", **out)


if __name__ == "__main__":
    main()
