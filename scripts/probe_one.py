"""One timing of one stepping kernel: python scripts/probe_one.py N B K path [repeat] -> steps/s, us per step and a hash of
the final state (to compare variants bit for bit)."""
import hashlib
import json
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import ldc_b200

N, B, K, path = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
rep = int(sys.argv[5]) if len(sys.argv) > 5 else 3
Re = 1000.0 * (1 + 0.05 * np.random.default_rng(0).uniform(-1, 1, B))
with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
    s.set_cases(Re).init_state()
    s.step(200, want_rel=False)
    ms = []
    for _ in range(rep):
        s.step(K, want_rel=False)
        ms.append(s.last_kernel_ms)
    h = hashlib.sha1(b"".join(np.ascontiguousarray(a).tobytes() for a in s.get_state())).hexdigest()[:12]
    best = min(ms)
    print(json.dumps(dict(kernel=s.kernel_name, N=N, B=B, K=K, steps_per_s=B * K / (best * 1e-3), us_per_step=best * 1e3 / K,
                          ms=ms, state_sha1=h)), flush=True)
