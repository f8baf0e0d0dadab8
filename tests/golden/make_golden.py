"""Mint golden fixtures by running the UNMODIFIED reference in-process.

Run in the authoring container only (needs /root/reference):

    python tests/golden/make_golden.py            # all fixtures (~3 min)
    python tests/golden/make_golden.py --quick    # skip the long solves

The reference imports `pyvista`, `mlflow` and `pyamg` at module top (base.py:10-11,
fv/linear_solvers/scipy_solver.py:4); none is installed here and none is on the spectral hot
path, so empty stub modules are placed in sys.modules.  Outputs are
small .npz files under tests/golden/ that travel with the repo; nothing at test/bench time reads
/root/reference.
"""
from __future__ import annotations

import argparse
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("LDC_REFERENCE", "/root/reference")


def import_reference():
    os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/numba_cache_ldc")
    pv = types.ModuleType("pyvista")
    pv.StructuredGrid = object
    ml = types.ModuleType("mlflow")
    ml.active_run = lambda: None
    ml.log_metrics = lambda *a, **k: None
    sys.modules.setdefault("pyvista", pv)
    sys.modules.setdefault("mlflow", ml)
    sys.modules.setdefault("pyamg", types.ModuleType("pyamg"))
    sys.path.insert(0, os.path.join(REF, "src"))
    import solvers  # noqa: F401
    from solvers.spectral.sg import SGSolver
    from solvers.spectral.fsg import FSGSolver
    from solvers.spectral.multigrid import fsg as mg
    from solvers.spectral.operators import transfer_operators as tr
    from solvers.spectral.operators.corner import create_corner_treatment
    return SGSolver, FSGSolver, mg, tr, create_corner_treatment


def sg_kwargs(N, Re, cs=0.15, **over):
    kw = dict(name="spectral", Re=Re, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=N, ny=N, tolerance=1e-6,
              max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0,
              corner_treatment="smoothing", corner_smoothing=cs, multigrid="none")
    kw.update(over)
    return kw


def fsg_kwargs(N, Re, cs=0.15, **over):
    kw = sg_kwargs(N, Re, cs, name="spectral_fsg", multigrid="fsg", n_levels=2, coarse_tolerance_factor=1.0,
                   prolongation_method="fft", restriction_method="fft")
    kw.update(over)
    return kw


def save(name, **arrays):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"wrote {name}: {os.path.getsize(path) / 1024:.1f} KiB")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    SGSolver, FSGSolver, mg, tr, create_corner = import_reference()

    # ---- operators -------------------------------------------------------
    ops = {}
    for N in (8, 15, 16, 30, 64):
        s = SGSolver(**sg_kwargs(N, 100.0))
        ops[f"x_{N}"] = s.x_full[:, 0]
        ops[f"Dx_{N}"] = s.Dx_1d
        ops[f"Dxx_{N}"] = s.Dxx_1d
        ops[f"Ix_{N}"] = s.Interp_x
        ops[f"W_{N}"] = s.W_2d
        ops[f"dxmin_{N}"] = np.array(s.dx_min)
    # non-unit, non-square domain
    s = SGSolver(**sg_kwargs(12, 100.0, Lx=2.0, Ly=0.5))
    ops.update(x_rect=s.x_full[:, 0], y_rect=s.y_full[0, :], Dx_rect=s.Dx_1d, Dy_rect=s.Dy_1d, Iy_rect=s.Interp_y,
               W_rect=s.W_2d)
    x30 = ops["x_30"]
    for cs in (0.0, 0.05, 0.15, 0.35):
        u_lid, v_lid = create_corner("smoothing", cs).get_lid_velocity(x30, x30 * 0 + 1, lid_velocity=1.0, Lx=1.0, Ly=1.0)
        ops[f"lid_smoothing_{cs}"] = u_lid
    u_lid, _ = create_corner("saad").get_lid_velocity(x30, x30 * 0 + 1, lid_velocity=2.0, Lx=1.0, Ly=1.0)
    ops["lid_saad_U2"] = u_lid
    for (nc, nf) in ((9, 17), (16, 31), (14, 29), (33, 65), (31, 63)):
        ops[f"Pfft_{nc}_{nf}"] = np.column_stack([tr.FFTProlongation().prolongate_1d(e, nf) for e in np.eye(nc)])
        ops[f"Ppoly_{nc}_{nf}"] = np.column_stack([tr.PolynomialProlongation().prolongate_1d(e, nf) for e in np.eye(nc)])
        ops[f"Rfft_{nf}_{nc}"] = np.column_stack([tr.FFTRestriction().restrict_1d(e, nc) for e in np.eye(nf)])
        ops[f"Rinj_{nf}_{nc}"] = np.column_stack([tr.InjectionRestriction().restrict_1d(e, nc) for e in np.eye(nf)])
    s16 = SGSolver(**sg_kwargs(16, 100.0))
    ops["orders"] = np.array([[n, L] + ([lv.n for lv in mg.build_hierarchy(n, L, s16.basis_x, s16.basis_y)] + [0, 0, 0])[:3]
                              for n, L in ((30, 2), (16, 2), (20, 3), (48, 3), (32, 2), (25, 3))])
    save("operators.npz", **ops)

    # ---- one residual evaluation + diagnostics + output pressure ----------
    rng = np.random.default_rng(1234)
    res = {}
    for N in (16, 30):
        n, m = N + 1, N - 1
        u = rng.standard_normal((n, n))
        v = rng.standard_normal((n, n))
        p = rng.standard_normal((m, m))
        s = SGSolver(**sg_kwargs(N, 250.0))
        s.arrays.u[:] = u.ravel()
        s.arrays.v[:] = v.ravel()
        s.arrays.p[:] = p.ravel()
        s._compute_residuals(s.arrays.u, s.arrays.v, s.arrays.p)
        res.update({f"u_{N}": u, f"v_{N}": v, f"p_{N}": p,
                    f"Ru_sg_{N}": s.arrays.R_u.reshape(n, n).copy(), f"Rv_sg_{N}": s.arrays.R_v.reshape(n, n).copy(),
                    f"Rp_sg_{N}": s.arrays.R_p.reshape(m, m).copy(),
                    f"EZP_{N}": np.array([s._compute_energy(), s._compute_enstrophy(), s._compute_palinstrophy()]),
                    f"pfull_{N}": s._extrapolate_to_full_grid(s.p_2d)})
        lv = mg.build_spectral_level(N, 0, s.basis_x, s.basis_y, 1.0, 1.0)
        sm = mg.MultigridSmoother(lv, Re=250.0, beta_squared=5.0, lid_velocity=1.0, CFL=1.5,
                                  corner_treatment=s.corner_treatment)
        sm._compute_residuals(u.ravel().copy(), v.ravel().copy(), p.ravel().copy())
        res.update({f"Ru_mg_{N}": lv.R_u.reshape(n, n).copy(), f"Rv_mg_{N}": lv.R_v.reshape(n, n).copy(),
                    f"Rp_mg_{N}": lv.R_p.reshape(m, m).copy()})
    res["Re"] = np.array(250.0)
    save("residual.npz", **res)

    # ---- K-step trajectories (both pressure semantics) ---------------------
    traj = {}
    for N, Re, K in ((16, 100.0, 25), (30, 1000.0, 10)):
        s = SGSolver(**sg_kwargs(N, Re))
        n, m = N + 1, N - 1
        snaps = []
        for k in range(K):
            s.step()
            if k in (0, 4, K - 1):
                snaps.append(np.concatenate([s.arrays.u, s.arrays.v, s.arrays.p]))
        traj[f"sg_{N}"] = np.array(snaps)
        traj[f"sg_{N}_Rnorm"] = np.array([np.linalg.norm(s.arrays.R_u), np.linalg.norm(s.arrays.R_v),
                                          np.linalg.norm(s.arrays.R_p)])
        lv = mg.build_spectral_level(N, 0, s.basis_x, s.basis_y, 1.0, 1.0)
        sm = mg.MultigridSmoother(lv, Re=Re, beta_squared=5.0, lid_velocity=1.0, CFL=1.5,
                                  corner_treatment=s.corner_treatment)
        sm.initialize_lid()
        snaps, rr = [], []
        for k in range(K):
            rr.append(sm.step())
            if k in (0, 4, K - 1):
                snaps.append(np.concatenate([lv.u, lv.v, lv.p]))
        traj[f"mg_{N}"] = np.array(snaps)
        traj[f"mg_{N}_res"] = np.array(rr)
        traj[f"meta_{N}"] = np.array([Re, K])
    save("steps.npz", **traj)

    # ---- prolongation with the BC quirk, from a random coarse state ---------
    pro = {}
    s = SGSolver(**sg_kwargs(16, 100.0))
    for method in ("fft", "polynomial"):
        lc = mg.build_spectral_level(8, 0, s.basis_x, s.basis_y, 1.0, 1.0)
        lf = mg.build_spectral_level(16, 1, s.basis_x, s.basis_y, 1.0, 1.0)
        lc.u[:] = rng.standard_normal(81)
        lc.v[:] = rng.standard_normal(81)
        lc.p[:] = rng.standard_normal(49)
        mg.prolongate_solution(lc, lf, tr.create_transfer_operators(method, "fft"), lid_velocity=1.0)
        pro.update({f"uc_{method}": lc.u.reshape(9, 9).copy(), f"vc_{method}": lc.v.reshape(9, 9).copy(),
                    f"pc_{method}": lc.p.reshape(7, 7).copy(), f"uf_{method}": lf.u.reshape(17, 17).copy(),
                    f"vf_{method}": lf.v.reshape(17, 17).copy(), f"pf_{method}": lf.p.reshape(15, 15).copy()})
    save("prolongation.npz", **pro)

    # ---- full solves --------------------------------------------------------
    sol = {}

    def record(tag, s):
        n = s.params.nx + 1
        sol[f"{tag}_u"] = s.arrays.u.reshape(n, n).copy()
        sol[f"{tag}_v"] = s.arrays.v.reshape(n, n).copy()
        sol[f"{tag}_p"] = s.arrays.p.reshape(n - 2, n - 2).copy()
        sol[f"{tag}_pfield"] = s.fields.p.reshape(n, n).copy()
        mt = s.metrics
        sol[f"{tag}_meta"] = np.array([mt.iterations, float(mt.converged), mt.final_residual, mt.u_momentum_residual,
                                       mt.v_momentum_residual, mt.continuity_residual, mt.final_energy,
                                       mt.final_enstrophy, mt.final_palinstrophy])
        print(tag, "iterations", mt.iterations, "converged", mt.converged, "wall", round(mt.wall_time_seconds, 1))

    # patch the expensive sparse streamfunction post-processing out: not on the path under test
    SGSolver.compute_vortex_metrics = lambda self: {}

    s = SGSolver(**sg_kwargs(16, 100.0))
    s.solve()
    record("sg_Re100_N16", s)
    s = FSGSolver(**fsg_kwargs(24, 100.0))
    s.solve()
    record("fsg_Re100_N24", s)
    s = FSGSolver(**fsg_kwargs(30, 1000.0, cs=0.05, max_iterations=500_000))
    s.solve()
    record("fsg_Re1000_N30_cs005", s)
    if not args.quick:
        s = FSGSolver(**fsg_kwargs(32, 100.0))
        s.solve()
        record("fsg_Re100_N32", s)
        s = FSGSolver(**fsg_kwargs(30, 1000.0, cs=0.15, max_iterations=500_000))
        s.solve()
        record("fsg_Re1000_N30_cs015", s)
        s = FSGSolver(**fsg_kwargs(64, 400.0))
        s.solve()
        record("fsg_Re400_N64", s)
        s = SGSolver(**sg_kwargs(32, 100.0))
        s.solve()
        record("sg_Re100_N32", s)
    save("solves_quick.npz" if args.quick else "solves.npz", **sol)


if __name__ == "__main__":
    main()
