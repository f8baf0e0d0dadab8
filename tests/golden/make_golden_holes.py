"""Mint the fixtures that close the round-1 parity holes ("Next round" item 7) by running the UNMODIFIED
reference in-process (authoring container only):

    python tests/golden/make_golden_holes.py

  * restriction: `restrict_solution` (`multigrid/fsg.py:622-664`, always injection) and the 2-D `FFTRestriction` /
    `InjectionRestriction` maps (`operators/transfer_operators.py:258-325, 379-428`) applied to a random N=16 state
  * BASELINE config 3: the first three cavities of the jittered-Re batch (Re = 400 (1 + 0.05 U(-1,1)), seed 0) at N = 64,
    200 `MultigridSmoother.step` calls from the cold start
  * BASELINE config 4: the four corners of the corner_smoothing x Re grid (cs in {0.02, 0.35}, Re in {100, 1000}) as
    FSG [16, 32] solves to 1e-6 (N = 32 keeps the minting short)
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import fsg_kwargs, import_reference, save, sg_kwargs  # noqa: E402


def main():
    SGSolver, FSGSolver, mg, tr, create_corner = import_reference()
    out = {}
    rng = np.random.default_rng(77)
    s = SGSolver(**sg_kwargs(16, 100.0))

    # ---- restriction ----
    lf = mg.build_spectral_level(16, 1, s.basis_x, s.basis_y, 1.0, 1.0)
    lc = mg.build_spectral_level(8, 0, s.basis_x, s.basis_y, 1.0, 1.0)
    lf.u[:] = rng.standard_normal(17 * 17)
    lf.v[:] = rng.standard_normal(17 * 17)
    lf.p[:] = rng.standard_normal(15 * 15)
    mg.restrict_solution(lf, lc, tr.create_transfer_operators("fft", "fft"))
    out.update(r_uf=lf.u.reshape(17, 17).copy(), r_vf=lf.v.reshape(17, 17).copy(), r_pf=lf.p.reshape(15, 15).copy(),
               r_uc_solution=lc.u.reshape(9, 9).copy(), r_vc_solution=lc.v.reshape(9, 9).copy(),
               r_pc_solution=lc.p.reshape(7, 7).copy())
    for name, op in (("fft", tr.FFTRestriction()), ("injection", tr.InjectionRestriction())):
        out[f"r_uc_{name}"] = op.restrict_2d(lf.u.reshape(17, 17), (9, 9))
        out[f"r_vc_{name}"] = op.restrict_2d(lf.v.reshape(17, 17), (9, 9))
        out[f"r_pc_{name}"] = op.restrict_2d(lf.p.reshape(15, 15), (7, 7))

    # ---- config 3: jittered Re at N = 64, 200 smoother steps ----
    Re3 = 400.0 * (1.0 + 0.05 * np.random.default_rng(0).uniform(-1.0, 1.0, 256))[:3]
    corner = create_corner("smoothing", 0.15)
    lv = mg.build_spectral_level(64, 0, s.basis_x, s.basis_y, 1.0, 1.0)
    snaps, res = [], []
    for Re in Re3:
        for a in (lv.u, lv.v, lv.p):
            a[:] = 0.0
        sm = mg.MultigridSmoother(lv, Re=float(Re), beta_squared=5.0, lid_velocity=1.0, CFL=1.5, corner_treatment=corner)
        sm.initialize_lid()
        for _ in range(200):
            r = sm.step()
        snaps.append(np.concatenate([lv.u, lv.v, lv.p]))
        res.append(r)
    out.update(c3_Re=Re3, c3_state=np.array(snaps), c3_res=np.array(res))

    # ---- config 4 corners: FSG [16, 32] to tolerance ----
    SGSolver.compute_vortex_metrics = lambda self: {}   # the sparse streamfunction solve is not on the path under test
    k = 0
    for cs in (0.02, 0.35):
        for Re in (100.0, 1000.0):
            t0 = time.time()
            f = FSGSolver(**fsg_kwargs(32, Re, cs=cs, max_iterations=500_000))
            f.solve()
            mt = f.metrics
            out[f"c4_{k}_u"] = f.arrays.u.reshape(33, 33).copy()
            out[f"c4_{k}_v"] = f.arrays.v.reshape(33, 33).copy()
            out[f"c4_{k}_p"] = f.arrays.p.reshape(31, 31).copy()
            out[f"c4_{k}_meta"] = np.array([cs, Re, mt.iterations, float(mt.converged)])
            print("config-4 corner", cs, Re, "iterations", mt.iterations, "converged", mt.converged,
                  round(time.time() - t0, 1), "s", flush=True)
            k += 1
    save("holes.npz", **out)


if __name__ == "__main__":
    main()
