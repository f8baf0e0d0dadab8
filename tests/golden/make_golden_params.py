"""Mint the parameter-variant fixture from the UNMODIFIED reference: lid_velocity, beta_squared and CFL away from their
defaults (conf/config.yaml, conf/solver/spectral/sg.yaml) -- they enter the lid profile, the continuity residual and the
adaptive pseudo-time step (sg.py:387-408: max(max|u|, lid_velocity)).

    python tests/golden/make_golden_params.py     # authoring container only (needs /root/reference)
"""
from __future__ import annotations

import numpy as np

from make_golden import import_reference, save, sg_kwargs

KW = dict(lid_velocity=2.0, beta_squared=2.5, CFL=1.0, corner_smoothing=0.08)


def main():
    SGSolver, *_ = import_reference()
    SGSolver.compute_vortex_metrics = lambda self: {}
    out = {}
    s = SGSolver(**sg_kwargs(16, 200.0, **KW))
    snaps = []
    for k in range(10):
        s.step()
        if k in (0, 4, 9):
            snaps.append(np.concatenate([s.arrays.u, s.arrays.v, s.arrays.p]))
    out["traj"] = np.array(snaps)
    s = SGSolver(**sg_kwargs(16, 200.0, **KW))
    s.solve()
    out.update(u=s.arrays.u.reshape(17, 17), v=s.arrays.v.reshape(17, 17), p=s.arrays.p.reshape(15, 15),
               meta=np.array([s.metrics.iterations, float(s.metrics.converged), s.metrics.final_energy,
                              s.metrics.final_enstrophy, s.metrics.final_palinstrophy]))
    print(s.metrics.iterations, s.metrics.converged)
    save("params.npz", **out)


if __name__ == "__main__":
    main()
