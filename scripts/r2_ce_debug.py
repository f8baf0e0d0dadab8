"""Clock samples of the cluster-resident kernel (LDC_S2_DEBUG): python scripts/r2_ce_debug.py N B K out.txt"""
import os
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
N, B, K, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
import ldc_b200

os.environ["LDC_STREAM_CW"] = "3"
Re = 5000.0 * (1 + 0.02 * np.random.default_rng(1).uniform(-1, 1, B))
with ldc_b200.BatchedSpectralLDC(N, batch=B, path="stream_tma") as s:
    s.set_cases(Re).init_state()
    s.step(10, want_rel=False)
    os.environ["LDC_S2_DEBUG"] = out
    s.step(K, want_rel=False)
    del os.environ["LDC_S2_DEBUG"]
    print("N", N, "B", B, s.kernel_name, "kernel ms", s.last_kernel_ms, "us/step", 1e3 * s.last_kernel_ms / K)
d = np.loadtxt(out)
d = d[8:8 + 4 * 40]   # skip the first two steps
order = [0, 1, 2, 3, 4, 6, 5, 10, 8]
names = ["K loop, remote part", "own elements + arrive", "accumulator swap", "assembly + own u, v panels",
         "sync + wait: panels read everywhere", "u, v sends, pressure, own p panels, sync", "K loop, local part + p sends",
         "wait: the others' output has arrived"]
for k, nm in enumerate(names):
    c = d[:, order[k + 1]] - d[:, order[k]]
    print(f"{nm:42s} mean {c.mean():8.0f} clk  min {c.min():8.0f}  max {c.max():8.0f}")
r3 = d[3::4]
print("reduction on warp 0 (last stage): wait for the sums + their addition", (r3[:, 12] - r3[:, 11]).mean(), "| residuals", (r3[:, 14] - r3[:, 12]).mean(), "| status", (r3[:, 15] - r3[:, 14]).mean(), "| start after the stage's [5]:", (r3[:, 11] - r3[:, 5]).mean())
print("stage period: mean", np.diff(d[:, 0]).mean(), "clk;  last stage of a step to the next:", np.diff(d[:, 0])[3::4].mean())
