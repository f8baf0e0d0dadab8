"""One SGSolver solve with the sampled TimeSeries (replay route) at a small size, as an ncu launch-list target:
    ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file out.csv python scripts/sg_history_target.py
"""
import sys

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import ldc_b200

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
s = ldc_b200.SGSolver(name="spectral", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=N, ny=N, tolerance=1e-6,
                      max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0,
                      corner_treatment="smoothing", corner_smoothing=0.15, multigrid="none")
s.solve()
print(s.metrics.iterations, s.metrics.converged, len(s.time_series.energy), s.metrics.wall_time_seconds)
s.close()
