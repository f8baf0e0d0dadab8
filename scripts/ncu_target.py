"""Short program for `ncu --set full`: two launches of the resident RK4 kernel at the bench shape."""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200

N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
P = int(sys.argv[3]) if len(sys.argv) > 3 else 8
PATH = sys.argv[4] if len(sys.argv) > 4 else "auto"
Re = 400.0 * (1 + 0.05 * np.random.default_rng(0).uniform(-1, 1, B))
with ldc_b200.BatchedSpectralLDC(N, batch=B, path=PATH) as s:
    s.set_cases(Re).init_state()
    print("kernel:", s.kernel_name)
    for _ in range(3):
        s.step(P, want_rel=False)
    u, v, p = s.get_state()
    assert np.isfinite(u).all()
print("ncu target OK", N, B, P, PATH)
