"""`basis_type: legendre` (conf/solver/spectral/sg.yaml:19; LegendreLobattoBasis spectral/basis/spectral.py:326-408): the
operators on the host and the device path against tests/golden/legendre.npz, minted by make_golden_legendre.py from the
unmodified reference.  Legendre-Gauss-Lobatto nodes are symmetric about the mid-point like the Chebyshev ones, so the
even-odd kernels apply unchanged; these tests are what says so."""
import numpy as np
import pytest


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("N", [16, 21])
def test_legendre_operators_match_reference(golden, N):
    from ldc_b200 import operators as ops

    g = golden.legendre
    lv = ops.build_level(N, N, 1.0, 1.0, "legendre")
    assert np.max(np.abs(lv.x - g[f"x_{N}"])) < 1e-15
    assert rel(lv.Dx, g[f"Dx_{N}"]) < 1e-13 and rel(lv.Dxx, g[f"Dxx_{N}"]) < 1e-12
    assert rel(lv.Ix, g[f"Ix_{N}"]) < 1e-12 and rel(lv.W, g[f"W_{N}"]) < 1e-13
    assert abs(lv.dx_min - float(g[f"dxmin_{N}"])) < 1e-16
    # the symmetry the parity kernels rest on: D is centro-antisymmetric, D^2 and the quadrature centro-symmetric
    # (to rounding, relative to the largest entry -- as for the reference's own matrices)
    assert np.max(np.abs(lv.Dx + lv.Dx[::-1, ::-1])) < 1e-12 * np.max(np.abs(lv.Dx))
    assert np.max(np.abs(lv.Dxx - lv.Dxx[::-1, ::-1])) < 1e-11 * np.max(np.abs(lv.Dxx))


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["auto", "resident", "tiled"])
def test_legendre_trajectory_matches_reference(built_lib, golden, path):
    g = golden.legendre["traj_21"]
    with built_lib.BatchedSpectralLDC(21, batch=1, basis="legendre", path=path) as b:
        b.set_cases(400.0).init_state()
        done = 0
        for row, k in zip(g, (1, 5, 10)):
            b.step(k - done, freeze_pressure=True, want_rel=False)   # SGSolver.step: frozen pressure gradient
            done = k
            u, v, p = b.get_state()
            got = np.concatenate([u.ravel(), v.ravel(), p.ravel()])
            assert rel(got, row) < 1e-12, (path, k)


@pytest.mark.gpu
def test_legendre_solves_match_reference(built_lib, golden):
    g = golden.legendre
    kw = dict(name="spectral", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=16, ny=16, tolerance=1e-6,
              max_iterations=10_000_000, basis_type="legendre", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
              corner_smoothing=0.15, multigrid="none")
    s = built_lib.SGSolver(**kw)
    s.solve()
    meta = g["sg_meta"]
    assert abs(s.metrics.iterations - int(meta[0])) <= 1 and s.metrics.converged == bool(meta[1])
    assert rel(s.arrays.u.reshape(17, 17), g["sg_u"]) < 1e-10 and rel(s.arrays.v.reshape(17, 17), g["sg_v"]) < 1e-10
    assert rel(s.arrays.p.reshape(15, 15), g["sg_p"]) < 1e-10
    got = np.array([s.metrics.final_energy, s.metrics.final_enstrophy, s.metrics.final_palinstrophy])
    assert np.max(np.abs(got - meta[3:6]) / np.abs(meta[3:6])) < 1e-9
    s.close()
    f = built_lib.FSGSolver(**{**kw, "name": "spectral_fsg", "nx": 24, "ny": 24, "multigrid": "fsg", "n_levels": 2,
                               "coarse_tolerance_factor": 1.0, "prolongation_method": "fft", "restriction_method": "fft"})
    f.solve()
    assert abs(f.metrics.iterations - int(g["fsg_meta"][0])) <= 1 and f.metrics.converged
    assert rel(f.arrays.u.reshape(25, 25), g["fsg_u"]) < 1e-10 and rel(f.arrays.v.reshape(25, 25), g["fsg_v"]) < 1e-10
    assert rel(f.arrays.p.reshape(23, 23), g["fsg_p"]) < 1e-10
    f.close()


@pytest.mark.gpu
@pytest.mark.parametrize("N", [16, 24])
def test_saad_regularised_lid_matches_reference(built_lib, golden, N):
    """`corner_treatment: saad` (conf/experiment/validation/saad-regu/spectral.yaml): FSG at Re=1000 with the polynomial lid
    u = 16 x^2 (1-x)^2, against tests/golden/saad.npz (make_golden_saad.py, unmodified reference)."""
    g = golden.saad
    f = built_lib.FSGSolver(name="spectral_fsg", Re=1000.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=N, ny=N, tolerance=1e-6,
                            max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0,
                            corner_treatment="saad", corner_smoothing=0.15, multigrid="fsg", n_levels=2,
                            coarse_tolerance_factor=1.0, prolongation_method="fft", restriction_method="fft")
    f.solve()
    n = N + 1
    assert abs(f.metrics.iterations - int(g[f"meta_{N}"][0])) <= 1 and f.metrics.converged
    assert rel(f.arrays.u.reshape(n, n), g[f"u_{N}"]) < 1e-10 and rel(f.arrays.v.reshape(n, n), g[f"v_{N}"]) < 1e-10
    assert rel(f.arrays.p.reshape(n - 2, n - 2), g[f"p_{N}"]) < 1e-10
    f.close()


RECT_SG = dict(name="spectral", Re=100.0, lid_velocity=1.0, Lx=2.0, Ly=0.5, nx=20, ny=14, tolerance=1e-6,
               max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
               corner_smoothing=0.15, multigrid="none")


@pytest.mark.gpu
def test_rectangular_cavity_trajectory_and_solve_match_reference(built_lib, golden):
    """Lx != Ly and nx != ny (tests/golden/rect.npz, make_golden_rect.py, unmodified reference): SG on a 20 x 14 grid over
    [0,2] x [0,0.5] -- operators of different size, the launch-per-stage path -- step by step and to tolerance."""
    g = golden.rect
    with built_lib.BatchedSpectralLDC(20, 14, batch=1, Lx=2.0, Ly=0.5) as b:
        assert b.kernel_name == "tiled"
        b.set_cases(100.0).init_state()
        done = 0
        for row, k in zip(g["sg_traj"], (1, 5, 10)):
            b.step(k - done, freeze_pressure=True, want_rel=False)
            done = k
            u, v, p = b.get_state()
            assert rel(np.concatenate([u.ravel(), v.ravel(), p.ravel()]), row) < 1e-12, k
    s = built_lib.SGSolver(**RECT_SG)
    s.solve()
    meta = g["sg_meta"]
    assert abs(s.metrics.iterations - int(meta[0])) <= 1 and s.metrics.converged
    assert rel(s.arrays.u.reshape(21, 15), g["sg_u"]) < 1e-10 and rel(s.arrays.v.reshape(21, 15), g["sg_v"]) < 1e-10
    assert rel(s.arrays.p.reshape(19, 13), g["sg_p"]) < 1e-10
    got = np.array([s.metrics.final_energy, s.metrics.final_enstrophy, s.metrics.final_palinstrophy])
    assert np.max(np.abs(got - meta[2:5]) / np.abs(meta[2:5])) < 1e-9
    s.close()


@pytest.mark.gpu
def test_stretched_square_grid_fsg_matches_reference(built_lib, golden):
    """nx == ny on [0,2] x [0,1]: x and y operators of the same size but different scaling -- the plain resident DMMA
    kernel with separate x / y operators (the even-odd kernel needs them identical); FSG [12, 24] to tolerance."""
    g = golden.rect
    f = built_lib.FSGSolver(**{**RECT_SG, "name": "spectral_fsg", "Lx": 2.0, "Ly": 1.0, "nx": 24, "ny": 24, "multigrid": "fsg",
                               "n_levels": 2, "coarse_tolerance_factor": 1.0, "prolongation_method": "fft",
                               "restriction_method": "fft"})
    f.solve()
    assert f._batch().kernel_name == "k_resident_rk4"
    assert abs(f.metrics.iterations - int(g["fsg_meta"][0])) <= 1 and f.metrics.converged
    assert rel(f.arrays.u.reshape(25, 25), g["fsg_u"]) < 1e-10 and rel(f.arrays.v.reshape(25, 25), g["fsg_v"]) < 1e-10
    assert rel(f.arrays.p.reshape(23, 23), g["fsg_p"]) < 1e-10
    f.close()


@pytest.mark.gpu
def test_fsg_knobs_match_reference(built_lib, golden):
    """The knobs of conf/solver/spectral/fsg.yaml beyond the defaults (tests/golden/fsg_variants.npz,
    make_golden_fsg_variants.py): three levels [12, 24, 48] with coarse_tolerance_factor=10 and polynomial prolongation;
    and an iteration cap that every level hits (the cap is per level, the result unconverged, the one-entry TimeSeries
    carries 10 x tolerance -- fsg.py:76-114)."""
    g = golden.fsg_variants
    base = dict(name="spectral_fsg", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, tolerance=1e-6, max_iterations=10_000_000,
                basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing", corner_smoothing=0.15,
                multigrid="fsg", n_levels=2, coarse_tolerance_factor=1.0, prolongation_method="fft", restriction_method="fft")
    for tag, over in (("three", dict(nx=48, ny=48, n_levels=3, coarse_tolerance_factor=10.0, prolongation_method="polynomial")),
                      ("capped", dict(nx=24, ny=24, max_iterations=300))):
        f = built_lib.FSGSolver(**{**base, **over})
        f.solve()
        meta, n = g[f"{tag}_meta"], over["nx"] + 1
        assert abs(f.metrics.iterations - int(meta[0])) <= (1 if tag == "three" else 0)
        assert f.metrics.converged == bool(meta[1])
        assert f.metrics.final_residual == pytest.approx(meta[2], rel=1e-12)
        assert len(f.time_series.rel_iter_residual) == int(meta[3])
        assert f.time_series.rel_iter_residual[-1] == pytest.approx(meta[4], rel=1e-12)
        assert rel(f.arrays.u.reshape(n, n), g[f"{tag}_u"]) < 1e-10 and rel(f.arrays.v.reshape(n, n), g[f"{tag}_v"]) < 1e-10
        assert rel(f.arrays.p.reshape(n - 2, n - 2), g[f"{tag}_p"]) < 1e-10
        f.close()


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["auto", "resident_eo", "resident", "tiled"])
def test_non_default_lid_velocity_beta_cfl_match_reference(built_lib, golden, path):
    """lid_velocity=2, beta_squared=2.5, CFL=1, corner_smoothing=0.08 at Re=200 (tests/golden/params.npz,
    make_golden_params.py): they enter the lid profile, the continuity residual and the adaptive step."""
    g = golden.params
    kw = dict(lid_velocity=2.0, beta_squared=2.5, CFL=1.0, corner_smoothing=0.08)
    with built_lib.BatchedSpectralLDC(16, batch=1, path=path) as b:
        b.set_cases(200.0, **kw).init_state()
        done = 0
        for row, k in zip(g["traj"], (1, 5, 10)):
            b.step(k - done, freeze_pressure=True, want_rel=False)
            done = k
            u, v, p = b.get_state()
            assert rel(np.concatenate([u.ravel(), v.ravel(), p.ravel()]), row) < 1e-12, (path, k)
    if path != "auto":
        return
    s = built_lib.SGSolver(name="spectral", Re=200.0, Lx=1.0, Ly=1.0, nx=16, ny=16, tolerance=1e-6, max_iterations=10_000_000,
                           basis_type="chebyshev", corner_treatment="smoothing", multigrid="none", **kw)
    s.solve()
    meta = g["meta"]
    assert abs(s.metrics.iterations - int(meta[0])) <= 1 and s.metrics.converged == bool(meta[1])
    assert rel(s.arrays.u.reshape(17, 17), g["u"]) < 1e-10 and rel(s.arrays.v.reshape(17, 17), g["v"]) < 1e-10
    assert rel(s.arrays.p.reshape(15, 15), g["p"]) < 1e-10
    got = np.array([s.metrics.final_energy, s.metrics.final_enstrophy, s.metrics.final_palinstrophy])
    assert np.max(np.abs(got - meta[2:5]) / np.abs(meta[2:5])) < 1e-9
    s.close()
