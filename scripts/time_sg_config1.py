"""BASELINE config 1 (SGSolver Re=100 N=32, single cavity) through the drop-in class: wall time of solve() with the
three TimeSeries modes."""
import json
import sys
import time

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200

KW = dict(name="spectral", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=32, ny=32, tolerance=1e-6,
          max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
          corner_smoothing=0.15, multigrid="none")
out = {}
for mode in (False, True, "full") * 2:
    s = ldc_b200.SGSolver(history=mode, **KW)
    t0 = time.time()
    s.solve()
    wall = time.time() - t0
    m = s.metrics
    out[str(mode)] = dict(wall_s=wall, iterations=m.iterations, converged=m.converged, points=len(s.time_series.energy),
                          final_residual=m.final_residual, energy=m.final_energy, psi_min=m.psi_min)
    print(mode, out[str(mode)], flush=True)
    s.close()
json.dump(out, open("gpurun_out/sg_config1.json", "w"), indent=1)
