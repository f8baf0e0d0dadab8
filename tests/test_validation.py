"""L2 error against a stored FV solution without PyVista (base.py:970-1054, 1122-1160).
CPU: the .vts reader, the interpolation plan and the oracle against the reference-minted fixture.
GPU: ldc_validation_errors against the fixture (1e-12 relative) and against the oracle on awkward point sets."""
import os

import numpy as np
import pytest


def plan_eval(plan, F):
    """NumPy evaluation of a BilinearPlan (test-side check of the plan the device kernel consumes)."""
    n1y = F.shape[1]
    f = F.ravel()
    o = plan.cell.astype(np.int64)
    tx, ty = plan.tx, plan.ty
    val = f[o] * ((1 - ty) * (1 - tx)) + f[o + n1y] * ((1 - ty) * tx) + f[o + 1] * (ty * (1 - tx)) + f[o + n1y + 1] * (ty * tx)
    return np.where(plan.mask == 1, val, np.nan)


def test_vts_roundtrip_and_reader(tmp_path):
    from ldc_b200 import vts

    rng = np.random.default_rng(3)
    nx, ny = 37, 21  # 37*21*3*8 bytes of points > one 32 KiB block: exercises the multi-block header
    pts = rng.standard_normal((nx * ny, 3))
    data = {"u": rng.standard_normal(nx * ny), "v": rng.standard_normal(nx * ny), "velocity": rng.standard_normal((nx * ny, 3)),
            "id": np.arange(nx * ny, dtype=np.int64)}
    path = os.path.join(tmp_path, "solution.vts")
    vts.write_vts(path, pts, (nx, ny, 1), data)
    g = vts.read_vts(path)
    assert g.dimensions == (nx, ny, 1) and np.array_equal(g.points, pts)
    for k, a in data.items():
        assert np.array_equal(g.point_data[k], a) and g.point_data[k].dtype == a.dtype
    with open(path) as f:
        text = f.read()
    with open(path, "w") as f:
        f.write(text.replace('type="StructuredGrid"', 'type="PolyData"', 1))
    with pytest.raises(ValueError):
        vts.read_vts(path)


def test_plan_matches_scipy_interpolator():
    from scipy.interpolate import RegularGridInterpolator

    from ldc_b200 import operators as ops
    from ldc_b200.validation import bilinear_plan

    lv = ops.build_level(12, 12)
    rng = np.random.default_rng(0)
    F = rng.standard_normal((13, 13))
    xq = np.concatenate([rng.uniform(-0.1, 1.1, 500), lv.x, [0.0, 1.0, 0.5, 1e-11]])
    yq = np.concatenate([rng.uniform(-0.1, 1.1, 500), lv.y[::-1], [0.3, 1.0, 0.0, 0.5]])
    plan = bilinear_plan(lv.x, lv.y, xq, yq, 1.0, 1.0)
    ref = RegularGridInterpolator((lv.y, lv.x), F.T.copy(), method="linear", bounds_error=False, fill_value=np.nan)(
        np.column_stack([yq, xq]))
    interior = (xq > 1e-10) & (xq < 1 - 1e-10) & (yq > 1e-10) & (yq < 1 - 1e-10)
    want_mask = interior & ~np.isnan(ref)
    assert np.array_equal(plan.mask == 1, want_mask)
    got = plan_eval(plan, F)
    assert np.allclose(got[want_mask], ref[want_mask], rtol=1e-13, atol=1e-15)


def test_oracle_matches_reference(golden):
    from oracle import ldc_oracle as orc

    g = golden.validation
    lv = orc.build_level(24)
    eu, ev, n = orc.validation_errors(lv, g["u"], g["v"], g["x"], g["y"], g["ref_u"], g["ref_v"])
    assert abs(eu - g["errors"][0]) < 1e-13 and abs(ev - g["errors"][1]) < 1e-13 and n == 128 * 128


@pytest.mark.gpu
def test_device_matches_reference_and_oracle(built_lib, golden):
    from ldc_b200.validation import StoredSolution, batched_validation_errors, bilinear_plan
    from oracle import ldc_oracle as orc

    g = golden.validation
    u, v = g["u"], g["v"]
    B = 3
    U = np.stack([u, 2 * u, u.copy()])
    V = np.stack([v, -v, v.copy()])
    U[2, 5, 7] = np.nan  # a cavity with a NaN: the points touching it are dropped (base.py:1028)
    with built_lib.BatchedSpectralLDC(24, batch=B) as s:
        s.set_cases(100.0)
        s.set_state(U, V, np.zeros(s.shape_inner))
        st = StoredSolution(g["x"], g["y"], g["ref_u"], g["ref_v"])
        e = batched_validation_errors(s, st)
        assert abs(e[0, 0] - g["errors"][0]) < 1e-12 and abs(e[0, 1] - g["errors"][1]) < 1e-12 and e[0, 2] == 128 * 128
        lv = orc.build_level(24)
        for b in range(B):
            eu, ev, n = orc.validation_errors(lv, U[b], V[b], st.x, st.y, st.u, st.v)
            assert abs(e[b, 0] - eu) < 1e-12 * max(1, eu) and abs(e[b, 1] - ev) < 1e-12 * max(1, ev) and e[b, 2] == n
        assert e[2, 2] < 128 * 128
        # scattered points incl. outside the domain and on the walls
        rng = np.random.default_rng(1)
        xq = np.concatenate([rng.uniform(-0.2, 1.2, 3001), [0.0, 1.0, 0.5]])
        yq = np.concatenate([rng.uniform(-0.2, 1.2, 3001), [0.5, 0.5, 1.0]])
        ru, rv = rng.standard_normal(xq.size), rng.standard_normal(xq.size)
        e = s.validation_errors(bilinear_plan(lv.x, lv.y, xq, yq, 1.0, 1.0), ru, rv)
        for b in range(B):
            eu, ev, n = orc.validation_errors(lv, U[b], V[b], xq, yq, ru, rv)
            assert abs(e[b, 0] - eu) < 1e-12 * max(1, eu) and abs(e[b, 1] - ev) < 1e-12 * max(1, ev) and e[b, 2] == n


@pytest.mark.gpu
def test_solver_compute_validation_errors(built_lib, golden, tmp_path, monkeypatch):
    """Drop-in method: reads data/validation/fv/Re100/solution.vts relative to the working directory."""
    from ldc_b200 import vts

    g = golden.validation
    d = tmp_path / "data" / "validation" / "fv" / "Re100"
    d.mkdir(parents=True)
    pts = np.column_stack([g["x"], g["y"], np.zeros(g["x"].size)])
    vts.write_vts(str(d / "solution.vts"), pts, (128, 128, 1), {"u": g["ref_u"], "v": g["ref_v"]})
    monkeypatch.chdir(tmp_path)
    s = built_lib.FSGSolver(name="spectral_fsg", Re=100.0, nx=24, ny=24, basis_type="chebyshev", max_iterations=10, n_levels=2,
                            multigrid="fsg")
    s.arrays.u[:] = g["u"].ravel()
    s.arrays.v[:] = g["v"].ravel()
    errs = s.compute_validation_errors()
    assert sorted(errs) == ["u_L2_error", "v_L2_error"]
    assert abs(errs["u_L2_error"] - g["errors"][0]) < 1e-12 and abs(errs["v_L2_error"] - g["errors"][1]) < 1e-12
    s.close()
