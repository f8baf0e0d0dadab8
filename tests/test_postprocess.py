"""Streamfunction / vortex metrics (fast diagonalisation) against the reference's sparse SuperLU solve
(fixtures minted by tests/golden/make_golden_vortex.py) and the solver-interface containers."""
import numpy as np
import pytest

from ldc_b200 import operators as ops
from ldc_b200 import postprocess as pp
from ldc_b200.datastructures import Metrics, SpectralParameters, TimeSeries

CASES = [("sg_Re100_N16", 16), ("fsg_Re100_N24", 24), ("fsg_Re100_N32", 32), ("fsg_Re1000_N30_cs015", 30),
         ("fsg_Re400_N64", 64)]


@pytest.mark.parametrize("tag,N", CASES)
def test_vortex_metrics_match_reference(golden, tag, N):
    g, v = golden.solves, golden.vortex
    lv = ops.build_level(N, N)
    psi, om = pp.streamfunction(lv, g[f"{tag}_u"], g[f"{tag}_v"])
    ref_psi = v[f"{tag}_psi"]
    assert np.linalg.norm(psi - ref_psi) / np.linalg.norm(ref_psi) < 1e-9
    got = pp.streamfunction_vortex_metrics(lv, g[f"{tag}_u"], g[f"{tag}_v"])
    for k, ref in zip(v[f"{tag}_keys"], v[f"{tag}_vals"]):
        assert got[str(k)] == pytest.approx(ref, rel=1e-8, abs=1e-12), k


def test_streamfunction_is_batched(golden):
    g = golden.solves
    lv = ops.build_level(16, 16)
    u = np.stack([g["sg_Re100_N16_u"], 2 * g["sg_Re100_N16_u"]])
    v = np.stack([g["sg_Re100_N16_v"], 2 * g["sg_Re100_N16_v"]])
    psi, _ = pp.streamfunction(lv, u, v)
    assert np.allclose(psi[1], 2 * psi[0], rtol=1e-12, atol=1e-15)
    assert not psi[:, 0].any() and not psi[:, :, -1].any()


def test_parameter_surface_matches_reference_yaml_keys():
    # conf/solver/spectral/fsg.yaml + sg.yaml + config.yaml keys
    kw = dict(name="spectral_fsg", Re=1000, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=30, ny=30, max_iterations=500000,
              tolerance=1e-6, basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
              corner_smoothing=0.15, multigrid="fsg", n_levels=2, coarse_tolerance_factor=1.0,
              prolongation_method="fft", restriction_method="fft")
    p = SpectralParameters(**kw)
    assert p.method == "Spectral-AC" and p.to_mlflow()["nx"] == 30
    with pytest.raises(TypeError):
        SpectralParameters(bogus_key=1)
    m = Metrics()
    assert "final_residual" not in m.to_mlflow()  # inf is skipped (datastructures.py:101-107)
    assert Metrics(converged=True).to_mlflow()["converged"] == 1
    assert TimeSeries(energy=[1.0, 2.0]).to_dataframe().shape == (2, 1)


def test_solver_construction_and_errors_without_gpu():
    from ldc_b200.solvers import FSGSolver, SGSolver

    s = SGSolver(name="spectral", Re=100, nx=16, ny=16, basis_type="chebyshev", CFL=1.5, corner_smoothing=0.15)
    assert s.shape_full == (17, 17) and s.shape_inner == (15, 15)
    assert s.arrays.u.size == 289 and s.arrays.p.size == 225
    assert s.u_2d[8, -1] == 1.0 and s.u_2d[0, -1] == 0.0  # smoothed lid is initialised on the host state
    assert s.W_2d.shape == (17, 17) and s.Dxx_1d.shape == (17, 17) and s.fields.x.size == 289
    with pytest.raises(ValueError, match="basis_type"):
        SGSolver(nx=8, ny=8, basis_type="fourier")
    with pytest.raises(ValueError, match="corner treatment"):
        SGSolver(nx=8, ny=8, basis_type="chebyshev", corner_treatment="bogus")
    with pytest.raises(TypeError):
        FSGSolver(nx=8, ny=8, not_a_key=3)
    f = FSGSolver(nx=16, ny=16, basis_type="chebyshev", prolongation_method="cubic")
    with pytest.raises(ValueError, match="prolongation"):
        f.solve()
