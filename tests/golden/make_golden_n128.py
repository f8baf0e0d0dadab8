"""Mint the BASELINE config-5 parity fixture (N = 128, Re = 5000) by running the UNMODIFIED reference in-process.

Run in the authoring container only (needs /root/reference and ~16 GB of RAM for the reference's Kronecker
matrices at N = 128, `multigrid/fsg.py:410-424`; they are built by the reference's own `build_spectral_level`
and never used on the stepping path):

    python tests/golden/make_golden_n128.py

What it records ("fixed K = 200 steps from an identical (prolongated) state"):
  * Re_b = 5000 (1 + 0.02 U(-1, 1)), seed 1 -- the first NCAV values of the 64-cavity batch; cs = 0.15
  * the N = 64 level state of each cavity after K64 `MultigridSmoother.step` calls from the cold start
    (`multigrid/fsg.py:956-995`) -- the input of the test
  * `prolongate_solution` 64 -> 128 (`multigrid/fsg.py:551-614`, fft method, BC quirk Q3, endpoint weights Q10),
    `initialize_lid`, then 200 smoother steps at N = 128: u, v, p after 1 and 200 steps, their norms after 50,
    and the (u_res, v_res) pairs of all 200 steps.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference, sg_kwargs, save  # noqa: E402

NCAV, K64, K128 = 2, 600, 200


def main():
    SGSolver, FSGSolver, mg, tr, create_corner = import_reference()
    rng = np.random.default_rng(1)
    Re_all = 5000.0 * (1.0 + 0.02 * rng.uniform(-1.0, 1.0, 64))
    s = SGSolver(**sg_kwargs(16, 100.0))   # only for its basis objects
    corner = create_corner("smoothing", 0.15)
    transfer = tr.create_transfer_operators("fft", "fft")
    t0 = time.time()
    out = {"Re": Re_all[:NCAV], "meta": np.array([NCAV, K64, K128])}
    lc = mg.build_spectral_level(64, 0, s.basis_x, s.basis_y, 1.0, 1.0)
    print("level 64 built", round(time.time() - t0, 1), "s")
    lf = mg.build_spectral_level(128, 1, s.basis_x, s.basis_y, 1.0, 1.0)
    print("level 128 built", round(time.time() - t0, 1), "s")
    for b in range(NCAV):
        Re = float(Re_all[b])
        for a in (lc.u, lc.v, lc.p, lf.u, lf.v, lf.p):
            a[:] = 0.0
        sm = mg.MultigridSmoother(lc, Re=Re, beta_squared=5.0, lid_velocity=1.0, CFL=1.5, corner_treatment=corner)
        sm.initialize_lid()
        for _ in range(K64):
            sm.step()
        out[f"c{b}_u64"] = lc.u.reshape(65, 65).copy()
        out[f"c{b}_v64"] = lc.v.reshape(65, 65).copy()
        out[f"c{b}_p64"] = lc.p.reshape(63, 63).copy()
        mg.prolongate_solution(lc, lf, transfer, 1.0)
        out[f"c{b}_u128_0"] = lf.u.reshape(129, 129).copy()   # the prolongated state BEFORE initialize_lid
        out[f"c{b}_v128_0"] = lf.v.reshape(129, 129).copy()
        out[f"c{b}_p128_0"] = lf.p.reshape(127, 127).copy()
        sf = mg.MultigridSmoother(lf, Re=Re, beta_squared=5.0, lid_velocity=1.0, CFL=1.5, corner_treatment=corner)
        sf.initialize_lid()
        res = []
        for k in range(K128):
            res.append(sf.step())
            if k + 1 in (1, K128):
                out[f"c{b}_u128_{k + 1}"] = lf.u.reshape(129, 129).copy()
                out[f"c{b}_v128_{k + 1}"] = lf.v.reshape(129, 129).copy()
                out[f"c{b}_p128_{k + 1}"] = lf.p.reshape(127, 127).copy()
            if k + 1 == 50:
                out[f"c{b}_norms128_50"] = np.array([np.linalg.norm(lf.u), np.linalg.norm(lf.v), np.linalg.norm(lf.p)])
        out[f"c{b}_res128"] = np.array(res)
        print("cavity", b, "Re", Re, "done", round(time.time() - t0, 1), "s; last res", res[-1])
    save("n128.npz", **out)


if __name__ == "__main__":
    main()
