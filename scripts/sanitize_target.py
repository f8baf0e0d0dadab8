"""Small runs of every stepping kernel and of the post-processing kernels, for compute-sanitizer (memcheck / racecheck):

    compute-sanitizer --tool racecheck python scripts/sanitize_target.py

(compute-sanitizer is closed on the round-1 GPU pool; run plainly it is a smoke of every path, and the run-to-run /
batch-composition bitwise tests in tests/test_gpu_parity.py stand in for racecheck.)
"""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200
from ldc_b200.validation import bilinear_plan

which = sys.argv[1:] or ["resident_eo", "resident_eo_split", "resident", "stream_eo", "tiled", "history", "validation"]
rng = np.random.default_rng(0)
for w in which:
    if w in ("resident_eo", "resident_eo_split", "resident", "tiled"):
        for N in ((15, 30) if w != "resident_eo" else (15, 30, 64)):
            with ldc_b200.BatchedSpectralLDC(N, batch=3, path=w) as s:
                s.set_cases([100.0, 400.0, 1000.0]).init_state()
                s.step(3)
                s.step(2, freeze_pressure=True)
                assert np.isfinite(s.get_state()[0]).all()
    elif w == "stream_eo":
        for N, B in ((66, 2), (128, 1), (96, 12)):   # double-buffered split-K shape and the two-CTA-per-SM shape
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path=w) as s:
                s.set_cases(400.0).init_state()
                s.step(2)
                assert np.isfinite(s.get_state()[0]).all()
    elif w == "history":
        with ldc_b200.BatchedSpectralLDC(12, batch=2) as s:
            s.set_cases([100.0, 200.0]).init_state()
            r, h = s.solve_history(1e-3, 40, chunk=16)
            assert h[0].shape[1] == 7
    elif w == "validation":
        with ldc_b200.BatchedSpectralLDC(12, batch=2) as s:
            s.set_cases(100.0).init_state()
            xq, yq = rng.uniform(-0.1, 1.1, 500), rng.uniform(-0.1, 1.1, 500)
            e = s.validation_errors(bilinear_plan(s.level.x, s.level.y, xq, yq, 1.0, 1.0), rng.standard_normal(500),
                                    rng.standard_normal(500))
            assert np.isfinite(e).all()
    print("sanitize target OK:", w, flush=True)
