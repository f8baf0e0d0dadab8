"""Pin the CPU oracle (oracle/ldc_oracle.py) against fixtures minted by running the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest

from oracle import ldc_oracle as orc


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("N", [8, 15, 16, 30, 64])
def test_operators_match_reference(golden, N):
    g = golden.operators
    lv = orc.build_level(N)
    assert np.array_equal(lv.x, g[f"x_{N}"])
    assert rel(lv.Dx, g[f"Dx_{N}"]) < 1e-15
    assert rel(lv.Dxx, g[f"Dxx_{N}"]) < 1e-14
    assert rel(lv.Ix, g[f"Ix_{N}"]) < 1e-13
    assert rel(lv.W, g[f"W_{N}"]) < 1e-14
    assert lv.dx_min == pytest.approx(float(g[f"dxmin_{N}"]), rel=1e-15)


def test_rectangular_domain(golden):
    g = golden.operators
    lx = orc.build_level(12, 2.0, 0.5)
    assert rel(lx.x, g["x_rect"]) < 1e-15 and rel(lx.y, g["y_rect"]) < 1e-15
    assert rel(lx.Dx, g["Dx_rect"]) < 1e-15 and rel(lx.Dy, g["Dy_rect"]) < 1e-15
    assert rel(lx.Iy, g["Iy_rect"]) < 1e-13 and rel(lx.W, g["W_rect"]) < 1e-14


def test_lid_profiles(golden):
    g = golden.operators
    x = g["x_30"]
    for cs in (0.0, 0.05, 0.15, 0.35):
        u, v = orc.lid_profile(x, "smoothing", cs, 1.0, 1.0)
        assert np.array_equal(u, g[f"lid_smoothing_{cs}"]) and not v.any()
    u, _ = orc.lid_profile(x, "saad", 0.0, 2.0, 1.0)
    assert np.array_equal(u, g["lid_saad_U2"])
    with pytest.raises(ValueError):
        orc.lid_profile(x, "nope", 0.1, 1.0, 1.0)


@pytest.mark.parametrize("nc,nf", [(9, 17), (16, 31), (14, 29), (33, 65), (31, 63)])
def test_transfer_matrices(golden, nc, nf):
    g = golden.operators
    assert rel(orc.transfer_matrix(orc.prolong_1d_fft, nc, nf), g[f"Pfft_{nc}_{nf}"]) < 1e-13
    assert rel(orc.transfer_matrix(orc.prolong_1d_poly, nc, nf), g[f"Ppoly_{nc}_{nf}"]) < 1e-9
    assert rel(orc.transfer_matrix(orc.restrict_1d_fft, nf, nc), g[f"Rfft_{nf}_{nc}"]) < 1e-13
    if f"Rinj_{nf}_{nc}" in g:
        assert np.array_equal(orc.transfer_matrix(orc.restrict_1d_injection, nf, nc), g[f"Rinj_{nf}_{nc}"])


def test_hierarchy_orders(golden):
    for n, L, *lv in golden.operators["orders"]:
        assert orc.hierarchy_orders(int(n), int(L)) == [int(a) for a in lv if a > 0]


@pytest.mark.parametrize("N", [16, 30])
def test_residual_one_evaluation(golden, N):
    g = golden.residual
    lv = orc.build_level(N)
    Ru, Rv, Rp = orc.residual(lv, g[f"u_{N}"], g[f"v_{N}"], g[f"p_{N}"], float(g["Re"]), 5.0)
    for tag in ("sg", "mg"):
        assert rel(Ru, g[f"Ru_{tag}_{N}"]) < 1e-13
        assert rel(Rv, g[f"Rv_{tag}_{N}"]) < 1e-13
        assert rel(Rp, g[f"Rp_{tag}_{N}"]) < 1e-13
    assert np.allclose(orc.diagnostics(lv, g[f"u_{N}"], g[f"v_{N}"]), g[f"EZP_{N}"], rtol=1e-12)
    assert rel(orc.extrapolate_pressure(g[f"p_{N}"]), g[f"pfull_{N}"]) < 1e-15


@pytest.mark.parametrize("N", [16, 30])
def test_step_trajectories(golden, N):
    g = golden.steps
    Re, K = g[f"meta_{N}"]
    K = int(K)
    lv = orc.build_level(N)
    case = orc.Case(Re=float(Re))
    ul, vl = orc.lid_profile(lv.x, "smoothing", 0.15, 1.0, 1.0)
    n2, m2 = (N + 1) ** 2, (N - 1) ** 2
    for tag, freeze in (("sg", True), ("mg", False)):
        u, v, p = orc.initial_state(lv, ul, vl)
        snaps, rr = [], []
        for k in range(K):
            ur, vr, Ru, Rv, Rp = orc.rk4_step(lv, u, v, p, case, ul, vl, freeze)
            rr.append((ur, vr))
            if k in (0, 4, K - 1):
                snaps.append(np.concatenate([u.ravel(), v.ravel(), p.ravel()]))
        ref = g[f"{tag}_{N}"]
        for s, r in zip(snaps, ref):
            assert rel(s[:n2], r[:n2]) < 1e-12 and rel(s[n2:2 * n2], r[n2:2 * n2]) < 1e-12
            assert rel(s[2 * n2:], r[2 * n2:]) < 1e-12
        if tag == "mg":
            assert np.allclose(rr, g[f"mg_{N}_res"], rtol=1e-10)
        else:
            got = [np.linalg.norm(Ru), np.linalg.norm(Rv), np.linalg.norm(Rp)]
            assert np.allclose(got, g[f"sg_{N}_Rnorm"], rtol=1e-10)


@pytest.mark.parametrize("method", ["fft", "polynomial"])
def test_prolongation_with_bc_quirk(golden, method):
    g = golden.prolongation
    lc, lf = orc.build_level(8), orc.build_level(16)
    uf, vf, pf = orc.prolongate_state(lc, lf, g[f"uc_{method}"], g[f"vc_{method}"], g[f"pc_{method}"], method, 1.0)
    tol = 1e-13 if method == "fft" else 1e-10
    assert rel(uf, g[f"uf_{method}"]) < tol and rel(vf, g[f"vf_{method}"]) < tol and rel(pf, g[f"pf_{method}"]) < tol


def test_sg_solve_matches_reference(golden):
    g = golden.solves
    out = orc.solve_sg(16, orc.Case(Re=100.0), 1e-6, 10_000_000)
    meta = g["sg_Re100_N16_meta"]
    assert abs(out["iterations"] - int(meta[0])) <= 1 and out["converged"]
    for f in ("u", "v", "p"):
        assert rel(out[f], g[f"sg_Re100_N16_{f}"]) < 1e-10
    assert np.allclose(out["res_norms"], meta[3:6], rtol=1e-8)


def test_fsg_solve_matches_reference(golden):
    g = golden.solves
    out = orc.solve_fsg(24, orc.Case(Re=100.0), 1e-6, 10_000_000)
    meta = g["fsg_Re100_N24_meta"]
    assert out["orders"] == [12, 24]
    assert abs(out["iterations"] - int(meta[0])) <= 1 and out["converged"]
    for f in ("u", "v", "p"):
        assert rel(out[f], g[f"fsg_Re100_N24_{f}"]) < 1e-10


def test_fsg_divergence_semantics(golden):
    """Re=1000 N=30 cs=0.05 diverges on the coarse level: zero fields, converged False.  The NaN
    iteration index itself grows from round-off, so only its order of magnitude is compared."""
    g = golden.solves
    out = orc.solve_fsg(30, orc.Case(Re=1000.0, corner_smoothing=0.05), 1e-6, 500_000)
    meta = g["fsg_Re1000_N30_cs005_meta"]
    assert not out["converged"] and out["diverged"] and meta[1] == 0.0
    assert not out["u"].any() and not g["fsg_Re1000_N30_cs005_u"].any()
    assert 0.5 * meta[0] < out["iterations"] < 2.0 * meta[0]


def test_config5_n128_prolongated_trajectory(golden):
    """BASELINE config 5 (tests/golden/make_golden_n128.py): the oracle's prolongation 64 -> 128 and 200 steps at N = 128,
    Re = 5000 against the reference's own run -- the same fixture the streamed CUDA kernels are held to."""
    g = golden.n128
    lc, lf = orc.build_level(64), orc.build_level(128)
    b = 0
    Re = float(g["Re"][b])
    uf, vf, pf = orc.prolongate_state(lc, lf, g[f"c{b}_u64"], g[f"c{b}_v64"], g[f"c{b}_p64"], "fft", 1.0)
    assert rel(uf, g[f"c{b}_u128_0"]) < 1e-13 and rel(vf, g[f"c{b}_v128_0"]) < 1e-13 and rel(pf, g[f"c{b}_p128_0"]) < 1e-13
    ul, vl = orc.lid_profile(lf.x, "smoothing", 0.15, 1.0, 1.0)
    uf[:, -1] = ul
    vf[:, -1] = vl
    case = orc.Case(Re=Re)
    for k in range(1, 201):
        orc.rk4_step(lf, uf, vf, pf, case, ul, vl, freeze_pressure=False)
        if k == 1:
            assert rel(uf, g[f"c{b}_u128_1"]) < 1e-12 and rel(pf, g[f"c{b}_p128_1"]) < 1e-12
        if k == 50:
            assert np.allclose([np.linalg.norm(uf), np.linalg.norm(vf), np.linalg.norm(pf)], g[f"c{b}_norms128_50"], rtol=1e-11)
    assert rel(uf, g[f"c{b}_u128_200"]) < 1e-10 and rel(vf, g[f"c{b}_v128_200"]) < 1e-10 and rel(pf, g[f"c{b}_p128_200"]) < 1e-10


def test_restriction_maps(golden):
    g = golden.holes
    for method in ("fft", "injection"):
        fn = orc.RESTRICT_1D[method]
        assert rel(orc.transfer_2d(fn, g["r_uf"], (9, 9)), g[f"r_uc_{method}"]) < 1e-13
        assert rel(orc.transfer_2d(fn, g["r_pf"], (7, 7)), g[f"r_pc_{method}"]) < 1e-13
    assert np.array_equal(g["r_uc_solution"], g["r_uc_injection"])   # restrict_solution always injects


def test_config3_jittered_cavity_n64(golden):
    g = golden.holes
    lv = orc.build_level(64)
    Re = float(g["c3_Re"][1])
    ul, vl = orc.lid_profile(lv.x, "smoothing", 0.15, 1.0, 1.0)
    u, v, p = orc.initial_state(lv, ul, vl)
    case = orc.Case(Re=Re)
    for _ in range(200):
        orc.rk4_step(lv, u, v, p, case, ul, vl, freeze_pressure=False)
    n2 = 65 * 65
    ref = g["c3_state"][1]
    assert rel(u.ravel(), ref[:n2]) < 1e-11 and rel(v.ravel(), ref[n2:2 * n2]) < 1e-11 and rel(p.ravel(), ref[2 * n2:]) < 1e-11
