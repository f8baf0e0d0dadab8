"""Clock samples of the persistent streamed kernel (LDC_S2_DEBUG): python scripts/r2_s2_debug.py N B K out.txt"""
import os
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
N, B, K, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
import ldc_b200

Re = 5000.0 * (1 + 0.02 * np.random.default_rng(1).uniform(-1, 1, B))
with ldc_b200.BatchedSpectralLDC(N, batch=B, path="stream_tma") as s:
    s.set_cases(Re).init_state()
    s.step(10, want_rel=False)
    os.environ["LDC_S2_DEBUG"] = out
    s.step(K, want_rel=False)
    del os.environ["LDC_S2_DEBUG"]
    print("N", N, "B", B, "kernel ms", s.last_kernel_ms, "us/step", 1e3 * s.last_kernel_ms / K)
d = np.loadtxt(out)
d = d[8:8 + 4 * 40]   # skip the first two steps
names = ["g0 first-chunk wait", "g0 K loop", "g0 waits in loop", "g1 first-chunk wait", "g1 K loop", "g1 waits in loop",
         "exchange+epilogue", "arrive", "producer barrier wait", "producer state issue", "epi: wait other group",
         "epi: accumulator swap", "epi: assembly+stores"]
cols = [1, 2, 3, 5, 6, 7, 8, 9, 11, 12, 13, 14, 15]
for nm, c in zip(names, cols):
    print(f"{nm:24s} mean {d[:, c].mean():9.0f} clk  min {d[:, c].min():9.0f}  max {d[:, c].max():9.0f}")
print("stage period (g0 start to start): mean", np.diff(d[:, 0]).mean(), "clk")
