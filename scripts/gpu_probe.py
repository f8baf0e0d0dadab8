"""GPU probe: FP64 ceilings + a first timing of the resident kernel at the bench shape."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200

out = {}
for w in ("dmma", "dfma", "mixed"):
    out[f"fp64_{w}_tflops"] = ldc_b200.fp64_peak_tflops(0, w)
print(out, flush=True)


def fstep(N):
    n, m = N + 1, N - 1
    return 4 * (20 * n**3 + 2 * n * m * m + 2 * n * n * m)


CASES = [(64, 256, 400), (64, 148, 400), (64, 1024, 200), (48, 512, 400), (32, 1024, 1000), (30, 1, 2000), (16, 2048, 2000)]
BIG = [(128, 8, 100), (128, 64, 50), (96, 64, 100)]
RUNS = [(p, c) for c in CASES for p in ("resident", "resident_eo")] + [(p, c) for c in BIG for p in ("tiled", "stream_eo")]
SMALL = [(30, 1, 4000), (30, 16, 2000), (30, 148, 2000), (30, 296, 2000), (30, 1024, 1000), (15, 1, 4000), (48, 1, 2000), (48, 148, 1000),
         (16, 2048, 2000), (32, 1024, 1000)]
if len(sys.argv) > 1 and sys.argv[1] == "small":
    RUNS = [(p, c) for c in SMALL for p in ("resident_eo", "resident_eo_split")]
MID = [(64, 1, 1000), (64, 8, 1000), (64, 16, 1000), (64, 37, 1000), (64, 74, 500), (64, 148, 500), (48, 8, 1000), (48, 37, 1000),
       (40, 8, 1000), (34, 8, 1000)]
if len(sys.argv) > 1 and sys.argv[1] == "mid":
    RUNS = [(p, c) for c in MID for p in ("resident_eo", "stream_eo")]
if len(sys.argv) > 1 and sys.argv[1] == "big":
    RUNS = [(p, c) for c in BIG for p in ("tiled", "stream_eo")]
for path, (N, B, K) in RUNS:
    Re = 400.0 * (1 + 0.05 * np.random.default_rng(0).uniform(-1, 1, B))
    with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
        s.set_cases(Re).init_state()
        s.step(50 if N <= 64 else 5, want_rel=False)
        t0 = time.time()
        s.step(K, want_rel=False)
        wall = time.time() - t0
        ms = s.last_kernel_ms
    rate = B * K / (ms * 1e-3)
    out[f"{path}_N{N}_B{B}"] = dict(steps_per_s=rate, kernel_ms=ms, wall_ms=wall * 1e3, tflops=rate * fstep(N) / 1e12)
    print(path, N, B, out[f"{path}_N{N}_B{B}"], flush=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
