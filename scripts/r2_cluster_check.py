"""Round-2 check of the cluster-resident kernel `k_cluster_eo` (the third shape of the `stream_tma` path, 50 <= N <= 65):
parity against the oracle and against the launch-per-stage streamed kernels (`stream_eo`), a queue run with more cavities
than clusters, a to-tolerance solve, then timings beside the 16 x 16 TMA shape.

    python scripts/r2_cluster_check.py [parity|time|all]
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


def rel(a, o):
    return float(np.linalg.norm(a - o) / max(np.linalg.norm(o), 1e-300))


def run(path, N, B, K, freeze, cw, Re, cs, state=None, steps2=0):
    if cw is None:
        os.environ.pop("LDC_STREAM_CW", None)
    else:
        os.environ["LDC_STREAM_CW"] = str(cw)
    with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
        s.set_cases(Re, corner_smoothing=cs)
        if state is None:
            s.init_state()
        else:
            s.set_state(*state)
        s.step(K, freeze_pressure=freeze)
        if steps2:
            s.step(steps2, freeze_pressure=freeze)
        return s.get_state(), s.kernel_name


if what in ("parity", "all"):
    for (N, B, K, freeze) in ((64, 2, 9, False), (64, 3, 6, True), (50, 2, 5, False), (57, 1, 4, True), (65, 2, 3, False)):
        Re = np.array([100.0, 250.0, 400.0])[:B]
        cs = np.array([0.15, 0.1, 0.2])[:B]
        got, kn = run("stream_tma", N, B, K, freeze, 3, Re, cs)
        old, _ = run("stream_eo", N, B, K, freeze, None, Re, cs)
        assert kn == "k_cluster_eo", kn
        lv = orc.build_level(N)
        worst = 0.0
        for b in range(B):
            case = orc.Case(Re=Re[b], corner_smoothing=cs[b])
            ul, vl = orc.lid_profile(lv.x, "smoothing", cs[b], 1.0, 1.0)
            uo, vo, po = orc.initial_state(lv, ul, vl)
            for _ in range(K):
                orc.rk4_step(lv, uo, vo, po, case, ul, vl, freeze_pressure=freeze)
            for a, o in zip((got[0][b], got[1][b], got[2][b]), (uo, vo, po)):
                worst = max(worst, rel(a, o))
        vs_old = max(rel(a, o) for a, o in zip(got, old))
        out[f"parity_N{N}_B{B}_K{K}_fz{int(freeze)}"] = dict(vs_oracle=worst, vs_stream_eo=vs_old, kernel=kn)
        print(N, B, K, freeze, "vs oracle", worst, "vs stream_eo", vs_old, kn, flush=True)
        assert vs_old < 1e-11 and worst < 1e-10, "parity"
    # general start state (boundary lines not the boundary conditions), two launches, more cavities than clusters (queue),
    # two runs bitwise equal
    N, B = 60, 21
    rng = np.random.default_rng(3)
    n, m = N + 1, N - 1
    st0 = (0.1 * rng.standard_normal((B, n, n)), 0.1 * rng.standard_normal((B, n, n)), 0.1 * rng.standard_normal((B, m, m)))
    Re = np.linspace(100, 900, B)
    cs = np.full(B, 0.15)
    a1, _ = run("stream_tma", N, B, 7, False, 3, Re, cs, st0, 5)
    a2, _ = run("stream_tma", N, B, 7, False, 3, Re, cs, st0, 5)
    ref, _ = run("stream_eo", N, B, 7, False, None, Re, cs, st0, 5)
    assert all(np.array_equal(x, y) for x, y in zip(a1, a2)), "k_cluster_eo is not bitwise reproducible"
    d = max(rel(a, o) for a, o in zip(a1, ref))
    out["general_state_N60_B21"] = d
    print("general state N=60 B=21 (queue), 7+5 steps: vs stream_eo", d, flush=True)
    assert d < 1e-10
    # to tolerance: iteration counts against the launch-per-stage kernels
    N, B = 52, 6
    Re = np.linspace(60, 200, B)
    its = {}
    for path, cw in (("stream_tma", 3), ("stream_eo", None)):
        if cw is None:
            os.environ.pop("LDC_STREAM_CW", None)
        else:
            os.environ["LDC_STREAM_CW"] = str(cw)
        with ldc_b200.BatchedSpectralLDC(N, batch=B, path=path) as s:
            s.set_cases(Re).init_state()
            t0 = time.time()
            r = s.solve(1e-5, 200000, warmup=0)
            its[path] = (r.iterations.copy(), r.converged.copy(), time.time() - t0, s.get_state(), s.last_kernel_ms)
    print("to tolerance N=52 B=6:", its["stream_tma"][0], its["stream_eo"][0], "kernel ms", its["stream_tma"][4], its["stream_eo"][4],
          flush=True)
    out["solve_N52"] = dict(iters_cluster=its["stream_tma"][0].tolist(), iters_eo=its["stream_eo"][0].tolist(),
                            ms_cluster=its["stream_tma"][4], ms_eo=its["stream_eo"][4])
    assert np.all(np.abs(its["stream_tma"][0] - its["stream_eo"][0]) <= 1)
    assert its["stream_tma"][1].all()
    assert max(rel(a, o) for a, o in zip(its["stream_tma"][3], its["stream_eo"][3])) < 1e-9

if what in ("time", "all"):
    peak = ldc_b200.fp64_peak_tflops(0, "dmma")
    out["fp64_dmma_tflops"] = peak
    for (N, B, K) in ((64, 1, 1000), (64, 8, 1000), (64, 15, 500), (64, 30, 300), (50, 8, 1000)):
        for cw in (1, 3):
            os.environ["LDC_STREAM_CW"] = str(cw)
            Re = 5000.0 * (1 + 0.02 * np.random.default_rng(1).uniform(-1, 1, B))
            with ldc_b200.BatchedSpectralLDC(N, batch=B, path="stream_tma") as s:
                s.set_cases(Re).init_state()
                s.step(10, want_rel=False)
                ms = []
                for _ in range(3):
                    s.step(K, want_rel=False)
                    ms.append(s.last_kernel_ms)
                kn = s.kernel_name
            best = min(ms)
            rate = B * K / (best * 1e-3)
            issued = 4 * 40 * 2 * 32 ** 3
            out[f"cw{cw}_N{N}_B{B}"] = dict(kernel=kn, steps_per_s=rate, us_per_step=best * 1e3 / K,
                                           pipe_frac=rate * issued / 1e12 / peak, ms=ms)
            print(cw, N, B, out[f"cw{cw}_N{N}_B{B}"], flush=True)
os.environ.pop("LDC_STREAM_CW", None)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/r2_cluster_check.json", "w"), indent=1)
print("OK")
