"""Tail hand-over of a to-tolerance solve (resident even-odd kernel -> streamed kernel once 2 x active <= #SM): the same
FSG [32, 64] solve of B cavities with and without it -- iteration counts, fields, kernel time.

    python scripts/r2_tail_switch.py [B]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import ldc_b200

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
cs, Re = np.meshgrid(np.linspace(0.02, 0.35, 32), np.linspace(100.0, 1000.0, 32), indexing="ij")
pick = np.arange(0, 1024, 1024 // B)[:B]          # what one of 1024 / B ranks gets from the config-4 sweep (round-robin)
cs, Re = cs.ravel()[pick], Re.ravel()[pick]
out = {}
for mode in ("0", "1"):
    os.environ["LDC_TAIL_SWITCH"] = mode
    t0 = time.time()
    r = ldc_b200.solve_fsg_batched(64, Re, 1e-6, 2_000_000, corner_smoothing=cs, n_levels=2, coarse_tolerance_factor=1.0)
    out[mode] = dict(wall=time.time() - t0, kernel_s=r.kernel_ms * 1e-3, it=r.iterations.copy(), conv=int(r.converged.sum()), u=r.u, p=r.p)
    print("tail switch", mode, "wall", out[mode]["wall"], "kernel s", out[mode]["kernel_s"], "iterations max", int(r.iterations.max()),
          "converged", out[mode]["conv"], flush=True)
a, b = out["0"], out["1"]
dit = int(np.abs(a["it"] - b["it"]).max())
du = float(max(np.linalg.norm(a["u"][k] - b["u"][k]) / np.linalg.norm(a["u"][k]) for k in range(B)))
print("max |iteration difference|", dit, "max rel-L2 difference of u", du)
json.dump(dict(B=B, kernel_s_off=a["kernel_s"], kernel_s_on=b["kernel_s"], iterations_max=int(a["it"].max()), max_iteration_diff=dit,
               max_rel_l2_u=du, converged=[a["conv"], b["conv"]]), open(f"gpurun_out/tail_switch_B{B}.json", "w"), indent=1)
assert dit <= 1 and du < 1e-10
