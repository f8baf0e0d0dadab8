"""Sampled history of BASELINE config 1 (SG Re=100 N=32): wall time of the replay route against the snapshot spacing."""
import json
import sys
import time

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200
from ldc_b200 import BatchedSpectralLDC as B

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
Re = float(sys.argv[2]) if len(sys.argv) > 2 else 100.0
out = []
with B(N, batch=1) as b:
    b.set_cases(Re)
    for method, seg in (("plain", 0), ("two_pass", 0), ("replay", 64), ("replay", 128), ("replay", 256), ("replay", 512),
                        ("replay", 1024), ("replay", 256)):
        b.init_state()
        B.REPLAY_SEGMENT = seg or 256
        t0 = time.time()
        if method == "plain":
            r = b.solve(1e-6, 10_000_000, warmup=10, freeze_pressure=True)
        else:
            r, h, idx = b.solve_history_sampled(1e-6, 10_000_000, nan_exit=True, method=method)
        rec = dict(method=method, segment=seg, wall_s=time.time() - t0, kernel_ms=r.kernel_ms, iterations=int(r.iterations[0]))
        out.append(rec)
        print(rec, flush=True)
json.dump(out, open("gpurun_out/replay_timing.json", "w"), indent=1)
