"""Round-2 check of the persistent TMA-fed streamed path (`stream_tma`): parity against the oracle and against the
launch-per-stage streamed kernels (`stream_eo`), a to-tolerance solve, then timings of both at the config-5 shapes.

    python scripts/r2_stream_check.py [parity|time|all]
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")
import __graft_entry__ as ge

ge.build()
import ldc_b200
from oracle import ldc_oracle as orc

what = sys.argv[1] if len(sys.argv) > 1 else "all"
out = {}
CW = os.environ.get("LDC_STREAM_CW", "auto")   # 1: 16 x 16 tiles (k_stream_tma), 2: 16 x 32 (k_stream_tma_wide)
print("tile shape LDC_STREAM_CW =", CW, flush=True)


def rel(a, o):
    return float(np.linalg.norm(a - o) / max(np.linalg.norm(o), 1e-300))


def fstep(N):
    n, m = N + 1, N - 1
    return 4 * (20 * n**3 + 2 * n * m * m + 2 * n * n * m)


if what in ("parity", "all"):
    for (N, B, K, freeze) in ((34, 3, 5, False), (40, 2, 6, True), (64, 2, 9, False), (100, 2, 4, False), (128, 2, 3, False),
                              (127, 1, 3, True)):
        Re = np.array([100.0, 250.0, 400.0])[:B]
        cs = np.array([0.15, 0.1, 0.2])[:B]
        res = {}
        for path in ("stream_tma", "stream_eo"):
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
                s.set_cases(Re, corner_smoothing=cs).init_state()
                s.step(K, freeze_pressure=freeze)
                res[path] = s.get_state()
                kn = s.kernel_name
        lv = orc.build_level(N)
        worst = 0.0
        for b in range(B):
            case = orc.Case(Re=Re[b], corner_smoothing=cs[b])
            ul, vl = orc.lid_profile(lv.x, "smoothing", cs[b], 1.0, 1.0)
            uo, vo, po = orc.initial_state(lv, ul, vl)
            for _ in range(K):
                orc.rk4_step(lv, uo, vo, po, case, ul, vl, freeze_pressure=freeze)
            for a, o in zip((res["stream_tma"][0][b], res["stream_tma"][1][b], res["stream_tma"][2][b]), (uo, vo, po)):
                worst = max(worst, rel(a, o))
        vs_old = max(rel(a, o) for a, o in zip(res["stream_tma"], res["stream_eo"]))
        assert vs_old < 1e-11
        out[f"parity_N{N}_B{B}_K{K}_fz{int(freeze)}"] = dict(vs_oracle=worst, vs_stream_eo=vs_old, kernel=kn)
        print(N, B, K, freeze, "vs oracle", worst, "vs stream_eo", vs_old, kn, flush=True)
        assert worst < 1e-10, "parity"
    # general start state (boundary lines not the boundary conditions), several launches, two runs bitwise equal
    N, B = 72, 5
    rng = np.random.default_rng(3)
    n, m = N + 1, N - 1
    u0, v0, p0 = 0.1 * rng.standard_normal((B, n, n)), 0.1 * rng.standard_normal((B, n, n)), 0.1 * rng.standard_normal((B, m, m))
    Re = np.linspace(100, 900, B)
    fin = {}
    for path in ("stream_tma", "stream_eo", "stream_tma"):
        with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
            s.set_cases(Re).set_state(u0, v0, p0)
            s.step(7)
            s.step(5)
            st = s.get_state()
        if path in fin:
            assert all(np.array_equal(a, c) for a, c in zip(fin[path], st)), "stream_tma is not bitwise reproducible"
        fin[path] = st
    d = max(rel(a, o) for a, o in zip(fin["stream_tma"], fin["stream_eo"]))
    out["general_state_N72"] = d
    print("general state N=72 B=5, 7+5 steps: vs stream_eo", d, flush=True)
    assert d < 1e-10
    # to tolerance: iteration counts of both kernels (many cavities per team: exercises the queue)
    N, B = 40, 24
    Re = np.linspace(60, 200, B)
    its = {}
    for path in ("stream_tma", "stream_eo"):
        with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
            s.set_cases(Re).init_state()
            t0 = time.time()
            r = s.solve(1e-5, 200000, warmup=0)
            its[path] = (r.iterations.copy(), r.converged.copy(), time.time() - t0, s.get_state())
    print("to tolerance N=40 B=24:", its["stream_tma"][0][:6], its["stream_eo"][0][:6], "wall", its["stream_tma"][2], its["stream_eo"][2],
          flush=True)
    out["solve_N40"] = dict(iters_tma=its["stream_tma"][0].tolist(), iters_eo=its["stream_eo"][0].tolist(),
                            wall_tma=its["stream_tma"][2], wall_eo=its["stream_eo"][2])
    assert np.all(np.abs(its["stream_tma"][0] - its["stream_eo"][0]) <= 1)
    assert its["stream_tma"][1].all()
    assert max(rel(a, o) for a, o in zip(its["stream_tma"][3], its["stream_eo"][3])) < 1e-9

if what in ("time", "all"):
    peak = ldc_b200.fp64_peak_tflops(0, "dmma")
    out["fp64_dmma_tflops"] = peak
    for (N, B, K) in ((128, 8, 200), (128, 64, 100), (128, 16, 200), (128, 1, 200), (96, 64, 100), (64, 8, 500), (64, 37, 300)):
        for path in ("stream_eo", "stream_tma"):
            Re = 5000.0 * (1 + 0.02 * np.random.default_rng(1).uniform(-1, 1, B))
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
                s.set_cases(Re).init_state()
                s.step(10, want_rel=False)
                ms = []
                for _ in range(3):
                    s.step(K, want_rel=False)
                    ms.append(s.last_kernel_ms)
            best = min(ms)
            rate = B * K / (best * 1e-3)
            issued = 4 * 40 * 2 * (4 * ((((N - 1) + 1) // 2 + 3) // 4)) ** 3   # DMMA flops issued per step (K4^3 blocks)
            out[f"{path}_N{N}_B{B}"] = dict(steps_per_s=rate, us_per_step=best * 1e3 / K, alg_tflops=rate * fstep(N) / 1e12,
                                            pipe_frac=rate * issued / 1e12 / peak, ms=ms)
            print(path, N, B, out[f"{path}_N{N}_B{B}"], flush=True)
json.dump(out, open(f"gpurun_out/r2_stream_check_cw{CW}.json", "w"), indent=1)
print("OK")
