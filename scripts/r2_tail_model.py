"""Per-level iteration counts of one rank's share of the config-4 sweep (128 of the 1024 cavities, round-robin) and what they
say about the phases of the solve: python scripts/r2_tail_model.py"""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import ldc_b200

B = 128
cs, Re = np.meshgrid(np.linspace(0.02, 0.35, 32), np.linspace(100.0, 1000.0, 32), indexing="ij")
pick = np.arange(0, 1024, 1024 // B)[:B]
cs, Re = cs.ravel()[pick], Re.ravel()[pick]
r = ldc_b200.solve_fsg_batched(64, Re, 1e-6, 2_000_000, corner_smoothing=cs, n_levels=2, coarse_tolerance_factor=1.0)
lv = [np.asarray(x) for x in r.per_level_iterations]
out = dict(kernel_s=r.kernel_ms * 1e-3, coarse=np.sort(lv[0]).tolist(), fine=np.sort(lv[1]).tolist())
json.dump(out, open("gpurun_out/tail_model_B128.json", "w"))
for name, it in (("N=32", lv[0]), ("N=64", lv[1])):
    s = np.sort(it)
    print(name, "iterations: min", s[0], "median", s[B // 2], "max", s[-1], "| active count falls below 74 / 31 / 15 at iteration",
          s[B - 74], s[B - 31], s[B - 15])
print("kernel s", out["kernel_s"])
