"""`solver.to_vtk().save(path)` (main.py:116; base.py:464-549) without PyVista.  CPU only: the drop-in class is built on the
host, handed the reference's own solution (tests/golden/vtk.npz, minted by make_golden_vtk.py from the unmodified
reference's to_vtk) and must produce the same arrays; the file it writes is read back."""
import os

import numpy as np

KW = dict(name="spectral_fsg", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=24, ny=24, tolerance=1e-6,
          max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
          corner_smoothing=0.15, multigrid="fsg", n_levels=2, coarse_tolerance_factor=1.0, prolongation_method="fft",
          restriction_method="fft")


def _solver_with_reference_solution(g):
    from ldc_b200.solvers import FSGSolver

    s = FSGSolver(**KW)          # host-side construction only: no device handle until a solve
    s.arrays.u[:] = g["u"].ravel()
    s.arrays.v[:] = g["v"].ravel()
    s.arrays.p[:] = g["p"].ravel()
    s._finalize_fields()
    return s


def test_to_vtk_matches_the_reference(golden):
    g = golden.vtk
    grid = _solver_with_reference_solution(g).to_vtk()
    assert grid.dimensions == (25, 25, 1) and grid.n_points == 625
    assert np.array_equal(grid.points[:, 0], g["X"].ravel("F")) and np.array_equal(grid.points[:, 1], g["Y"].ravel("F"))
    assert not grid.points[:, 2].any()
    assert sorted(grid.point_data) == sorted(k[5:] for k in g.files if k.startswith("grid_"))
    for k in ("u", "v", "velocity_magnitude", "velocity"):
        assert np.array_equal(grid[k], g["grid_" + k]), k
    # pressure passes through the linear extrapolation to the walls, vorticity through SciPy's spline fit
    assert np.allclose(grid["pressure"], g["grid_pressure"], rtol=0, atol=1e-14)
    assert np.allclose(grid["vorticity"], g["grid_vorticity"], rtol=1e-12, atol=1e-12)
    assert np.array_equal(grid.field_data["Re"], g["Re"]) and np.array_equal(grid.field_data["N"], g["N"])
    assert grid.field_data["solver"] == str(g["solver"]) == "spectral_fsg"


def test_saved_file_reads_back(golden, tmp_path):
    from ldc_b200 import vts

    grid = _solver_with_reference_solution(golden.vtk).to_vtk()
    path = os.path.join(tmp_path, "solution.vts")
    grid.save(path)
    back = vts.read_vts(path)
    assert back.dimensions == grid.dimensions and np.array_equal(back.points, grid.points)
    for k, a in grid.point_data.items():
        assert np.array_equal(back[k], a), k
    assert back.field_data["solver"] == "spectral_fsg"
    assert back.field_data["Re"].dtype == np.float64 and back.field_data["Re"][0] == 100.0
    assert back.field_data["N"].dtype == np.int64 and back.field_data["N"][0] == 24
    text = open(path).read()
    assert '<PointData Scalars="u">' in text and 'Name="velocity" NumberOfComponents="3"' in text and "RangeMin=" in text


def test_field_data_bytes_are_vtks():
    """The three FieldData entries of a file VTK itself wrote for the reference (data/validation/fv/Re100/solution.vts:
    Re = 100 and N = 128 as Int64, solver = "fv"): the writer produces the same base64 payloads."""
    from ldc_b200 import vts

    assert vts._encode_binary(np.array([100], dtype=np.int64)) == "AQAAAACAAAAIAAAACwAAAA==eJxLYYAAAAMoAGU="
    assert vts._encode_binary(np.array([128], dtype=np.int64)) == "AQAAAACAAAAIAAAACwAAAA==eJxrYIAAAAQIAIE="
    assert vts._encode_binary(np.frombuffer(b"fv\0", dtype="u1")) == "AQAAAACAAAADAAAACwAAAA==eJxLK2MAAAIhAN0="


def test_validation_table_matches_the_reference(tmp_path, monkeypatch):
    """mlflow_log_validation_table (base.py:890-964): the rows the reference logs, for a Re whose CSV has the columns the
    method asks for (1000), one whose CSV names them differently (100: every reference entry is "-") and one without a
    file (250); tests/golden/table.json was minted by make_golden_table.py from the unmodified reference."""
    import json
    import sys
    import types

    from ldc_b200.solvers import SGSolver

    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "table.json")))
    d = tmp_path / "data" / "validation" / "botella"
    d.mkdir(parents=True)
    for Re, case in g["cases"].items():
        if case["csv"] is not None:
            (d / f"botella_Re{Re}_vortex.csv").write_text(case["csv"])
    monkeypatch.chdir(tmp_path)
    logged = {}
    ml = types.ModuleType("mlflow")
    ml.active_run = lambda: True
    ml.log_table = lambda df, artifact_file: logged.update(rows=df.to_dict(orient="records"), file=artifact_file)
    monkeypatch.setitem(sys.modules, "mlflow", ml)
    for Re, case in g["cases"].items():
        s = SGSolver(**{**KW, "name": "spectral", "multigrid": "none", "nx": 12, "ny": 12, "Re": float(Re)})
        for k, v in g["metrics"].items():
            setattr(s.metrics, k, v)
        assert s.validation_table() == case["rows"]
        logged.clear()
        s.mlflow_log_validation_table()
        assert logged.get("rows") == case["rows"] and logged.get("file") == case["file"]
    ml.active_run = lambda: None
    logged.clear()
    s.mlflow_log_validation_table()
    assert not logged


import pytest


@pytest.mark.gpu
def test_solve_then_export_matches_the_reference_file_content(built_lib, golden, tmp_path):
    """The whole tail of main.py's run flow on the device path: FSGSolver.solve() (Re=100, N=24), then to_vtk().save();
    the arrays in the file against those the reference's own solve + to_vtk produced (fixture vtk.npz)."""
    from ldc_b200 import vts

    g = golden.vtk
    s = built_lib.FSGSolver(**KW)
    s.solve()
    path = os.path.join(tmp_path, "solution.vts")
    s.to_vtk().save(path)
    s.close()
    back = vts.read_vts(path)

    def rel(a, b):
        return np.linalg.norm(a - b) / np.linalg.norm(b)

    assert rel(back["u"], g["grid_u"]) < 1e-10 and rel(back["v"], g["grid_v"]) < 1e-10
    assert rel(back["pressure"], g["grid_pressure"]) < 1e-9
    assert rel(back["velocity_magnitude"], g["grid_velocity_magnitude"]) < 1e-10
    assert rel(back["vorticity"], g["grid_vorticity"]) < 1e-9
    assert rel(back["velocity"], g["grid_velocity"]) < 1e-10
    assert back.field_data["solver"] == "spectral_fsg" and back.field_data["N"][0] == 24


def test_reader_takes_ascii_and_uncompressed_binary(tmp_path):
    """The two other inline encodings VTK writes (`format="ascii"`, `format="binary"` without a compressor): same grid
    as the compressed flavour."""
    import base64

    from ldc_b200 import vts

    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0], [0, 2, 0], [1, 2, 0]], dtype=np.float64)
    u = np.array([0.5, -1.25, 2.0, 3.5, -4.0, 6.25])

    def raw(a):
        b = a.tobytes()
        return base64.b64encode(np.array([len(b)], dtype="<u4").tobytes() + b).decode()

    ascii_file = f"""<?xml version="1.0"?>
<VTKFile type="StructuredGrid" version="0.1" byte_order="LittleEndian">
  <StructuredGrid WholeExtent="0 1 0 2 0 0"><Piece Extent="0 1 0 2 0 0">
    <PointData><DataArray type="Float64" Name="u" format="ascii">{' '.join(repr(float(x)) for x in u)}</DataArray></PointData>
    <Points><DataArray type="Float64" NumberOfComponents="3" format="ascii">{' '.join(repr(float(x)) for x in pts.ravel())}</DataArray></Points>
  </Piece></StructuredGrid>
</VTKFile>"""
    binary_file = f"""<?xml version="1.0"?>
<VTKFile type="StructuredGrid" version="0.1" byte_order="LittleEndian" header_type="UInt32">
  <StructuredGrid WholeExtent="0 1 0 2 0 0"><Piece Extent="0 1 0 2 0 0">
    <PointData><DataArray type="Float64" Name="u" format="binary">{raw(u)}</DataArray></PointData>
    <Points><DataArray type="Float64" NumberOfComponents="3" format="binary">{raw(pts)}</DataArray></Points>
  </Piece></StructuredGrid>
</VTKFile>"""
    for name, text in (("a.vts", ascii_file), ("b.vts", binary_file)):
        path = os.path.join(tmp_path, name)
        with open(path, "w") as f:
            f.write(text)
        g = vts.read_vts(path)
        assert g.dimensions == (2, 3, 1) and np.array_equal(g.points, pts) and np.array_equal(g["u"], u)
    bad = os.path.join(tmp_path, "c.vts")
    with open(bad, "w") as f:
        f.write(binary_file.replace('format="binary"', 'format="appended"', 1))
    with pytest.raises(ValueError, match="appended"):
        vts.read_vts(bad)


def test_save_vtk_creates_directories(golden, tmp_path):
    from ldc_b200 import vts

    s = _solver_with_reference_solution(golden.vtk)
    path = tmp_path / "out" / "run7" / "solution.vts"
    s.save_vtk(path)
    assert vts.read_vts(str(path)).n_points == 625


@pytest.mark.gpu
def test_compute_global_quantities(built_lib, golden):
    """base.py:551-567 on the device path: E, Z, P of the reference's own FSG Re=100 N=24 solution against the values
    its solve reported (tests/golden/solves.npz meta: energy, enstrophy, palinstrophy)."""
    g = golden.vtk
    s = built_lib.FSGSolver(**KW)
    s.arrays.u[:] = g["u"].ravel()
    s.arrays.v[:] = g["v"].ravel()
    s.arrays.p[:] = g["p"].ravel()
    q = s.compute_global_quantities()
    s.close()
    assert set(q) == {"E", "Z", "P"} and all(np.isfinite(v) and v > 0 for v in q.values())
    from oracle import ldc_oracle as orc

    lv = orc.build_level(24)
    E, Z, Pal = orc.diagnostics(lv, g["u"], g["v"])
    assert abs(q["E"] - E) < 1e-13 * E and abs(q["Z"] - Z) < 1e-12 * Z and abs(q["P"] - Pal) < 1e-11 * Pal
