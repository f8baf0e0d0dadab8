"""Host-side operator setup of the product (ldc_b200.operators) against the reference-minted fixtures. CPU only."""
import numpy as np
import pytest

from ldc_b200 import operators as ops


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("N", [8, 15, 16, 30, 64])
def test_level_operators(golden, N):
    g = golden.operators
    lv = ops.build_level(N, N, 1.0, 1.0, "chebyshev")
    assert np.array_equal(lv.x, g[f"x_{N}"])
    assert rel(lv.Dx, g[f"Dx_{N}"]) < 1e-15
    assert rel(lv.Dxx, g[f"Dxx_{N}"]) < 1e-14
    assert rel(lv.Ix, g[f"Ix_{N}"]) < 1e-13
    assert rel(lv.W, g[f"W_{N}"]) < 1e-14
    assert lv.dx_min == pytest.approx(float(g[f"dxmin_{N}"]), rel=1e-15)
    # the resident kernel treats the interior rows of the interpolation as the identity
    assert np.abs(lv.Ix[1:-1] - np.eye(N - 1)).max() < 1e-12


def test_rectangular_level(golden):
    g = golden.operators
    lv = ops.build_level(12, 12, 2.0, 0.5, "chebyshev")
    assert rel(lv.x, g["x_rect"]) < 1e-15 and rel(lv.y, g["y_rect"]) < 1e-15
    assert rel(lv.Dx, g["Dx_rect"]) < 1e-15 and rel(lv.Dy, g["Dy_rect"]) < 1e-15
    assert rel(lv.Iy, g["Iy_rect"]) < 1e-13 and rel(lv.W, g["W_rect"]) < 1e-14


def test_lid_profiles(golden):
    g = golden.operators
    x = g["x_30"]
    for cs in (0.0, 0.05, 0.15, 0.35):
        u, v = ops.lid_velocity_profile(x, "smoothing", cs, 1.0, 1.0)
        assert np.array_equal(u, g[f"lid_smoothing_{cs}"]) and not v.any()
    u, _ = ops.lid_velocity_profile(x, "saad", 0.0, 2.0, 1.0)
    assert np.array_equal(u, g["lid_saad_U2"])
    u2, _ = ops.lid_velocity_profile(x, "polynomial", 0.0, 2.0, 1.0)
    assert np.array_equal(u, u2)
    with pytest.raises(ValueError, match="Unknown corner treatment"):
        ops.lid_velocity_profile(x, "subtraction", 0.1, 1.0, 1.0)


@pytest.mark.parametrize("nc,nf", [(9, 17), (16, 31), (14, 29), (33, 65), (31, 63)])
def test_transfer_matrices(golden, nc, nf):
    g = golden.operators
    assert rel(ops.prolongation_matrix(nc, nf, "fft"), g[f"Pfft_{nc}_{nf}"]) < 1e-13
    assert rel(ops.prolongation_matrix(nc, nf, "polynomial"), g[f"Ppoly_{nc}_{nf}"]) < 1e-9
    assert rel(ops.restriction_matrix(nf, nc, "fft"), g[f"Rfft_{nf}_{nc}"]) < 1e-13
    if f"Rinj_{nf}_{nc}" in g:
        assert np.array_equal(ops.restriction_matrix(nf, nc, "injection"), g[f"Rinj_{nf}_{nc}"])
    with pytest.raises(ValueError):
        ops.prolongation_matrix(nf, nc, "fft")
    with pytest.raises(ValueError):
        ops.prolongation_matrix(nc, nf, "spline")


def test_hierarchy_orders(golden):
    for n, L, *lv in golden.operators["orders"]:
        assert ops.hierarchy_orders(int(n), int(L)) == [int(a) for a in lv if a > 0]


def test_pressure_extrapolation(golden):
    g = golden.residual
    for N in (16, 30):
        assert rel(ops.extrapolate_pressure_to_full(g[f"p_{N}"]), g[f"pfull_{N}"]) < 1e-15


def test_legendre_basis_is_a_differentiation_matrix():
    lv = ops.build_level(12, 12, 1.0, 1.0, "legendre")
    x = lv.x
    assert x[0] == 0.0 and x[-1] == 1.0 and np.all(np.diff(x) > 0)
    f = x**5 - 2 * x**2
    assert np.allclose(lv.Dx @ f, 5 * x**4 - 4 * x, atol=1e-11)
    assert np.isclose(lv.W.sum(), 1.0, rtol=1e-13)
    with pytest.raises(ValueError, match="Unknown basis_type"):
        ops.build_level(8, 8, 1.0, 1.0, "fourier")
