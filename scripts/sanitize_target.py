"""Small runs of every stepping kernel and of the post-processing kernels, for compute-sanitizer (memcheck / racecheck):

    compute-sanitizer --tool racecheck python scripts/sanitize_target.py

Small shapes throughout (the tools slow kernels down by two orders of magnitude); LDC_STREAM_TEAMS=2 makes the persistent
streamed kernel hand its five cavities round between two teams, so the queue, the team barriers and the mailboxes all run.
Summaries are kept under profiles/r02_sanitizer_*.txt.
"""
import os
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200
from ldc_b200.validation import bilinear_plan

which = sys.argv[1:] or ["resident_eo", "resident_eo_split", "resident", "stream_eo", "stream_tma", "tiled", "history", "validation",
                         "vortex"]
rng = np.random.default_rng(0)
for w in which:
    if w in ("resident_eo", "resident_eo_split", "resident", "tiled"):
        for N in ((15, 30) if w != "resident_eo" else (15, 30, 64)):
            # (the even-odd kernel also with more cavities than CTA slots would need > 148 cavities: its queue is the one
            # the streamed kernel exercises below with two teams)
            with ldc_b200.BatchedSpectralLDC(N, batch=3, path=w) as s:
                s.set_cases([100.0, 400.0, 1000.0]).init_state()
                s.step(3)
                s.step(2, freeze_pressure=True)
                assert np.isfinite(s.get_state()[0]).all()
    elif w == "stream_eo":
        for N, B in ((66, 2), (96, 12)):   # double-buffered split-K shape and the two-CTA-per-SM shape
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path=w) as s:
                s.set_cases(400.0).init_state()
                s.step(2)
                assert np.isfinite(s.get_state()[0]).all()
    elif w == "stream_tma":
        os.environ["LDC_STREAM_TEAMS"] = "2"
        for N, B, K in ((34, 5, 9), (66, 2, 2)):   # 4-CTA teams with cavities migrating between them; 9-CTA teams
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path=w) as s:
                s.set_cases(np.linspace(100.0, 400.0, B)).init_state()
                s.step(K)
                s.solve(1e-1, 6, warmup=0)
                assert np.isfinite(s.get_state()[0]).all()
        del os.environ["LDC_STREAM_TEAMS"]
    elif w == "vortex":
        from ldc_b200 import postprocess as pp

        with ldc_b200.BatchedSpectralLDC(12, batch=2) as s:
            s.set_cases(100.0).init_state()
            s.step(5)
            assert len(pp.vortex_metrics_device(s)) == 2
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
