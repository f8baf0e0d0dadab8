"""Per-iteration TimeSeries of `solver=spectral/sg` (base.py:243-313, 125-153).
CPU: the oracle against the reference-minted fixture and the down-sampling rule.  GPU: ldc_solve_history through the
drop-in SGSolver against the same fixture (tolerance 1e-9 relative on every record, iteration count exact)."""
import numpy as np
import pytest

KW = dict(name="spectral", Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, nx=16, ny=16, tolerance=1e-6,
          max_iterations=10_000_000, basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
          corner_smoothing=0.15, multigrid="none")


def relcols(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300), axis=0)


def test_oracle_history_matches_reference(golden):
    from oracle import ldc_oracle as orc

    g = golden.history
    out = orc.solve_sg(16, orc.Case(Re=100.0), 1e-6, 40, history=True)
    h = out["history"][10:]
    assert h.shape == g["short_series"].shape == (30, 7)
    assert relcols(h, g["short_series"]).max() < 1e-11


def test_downsample_rule():
    import sys

    from ldc_b200.solvers import SGSolver

    col = np.arange(14145, dtype=float)
    out = SGSolver._downsample(col)
    assert len(out) == 1000 and out[0] == 0.0 and out[-1] == 14144.0
    assert out == [float(i) for i in np.linspace(0, 14144, 1000, dtype=int)]
    assert SGSolver._downsample(col[:1000]) == list(col[:1000])


@pytest.mark.gpu
def test_short_history_matches_reference(built_lib, golden):
    g = golden.history
    s = built_lib.SGSolver(**KW)
    s.solve(max_iter=40)
    ts = s.time_series
    got = np.array([ts.rel_iter_residual, ts.u_residual, ts.v_residual, ts.continuity_residual, ts.energy, ts.enstrophy,
                    ts.palinstrophy]).T
    assert got.shape == (30, 7)
    assert relcols(got, g["short_series"]).max() < 1e-10
    assert s.metrics.iterations == int(g["short_iterations"]) and not s.metrics.converged
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [True, "full"])   # True: only the records that survive the thinning are formed
def test_full_history_and_metrics_match_reference(built_lib, golden, mode):
    g = golden.history
    s = built_lib.SGSolver(history=mode, **KW)
    s.solve()
    meta = g["full_meta"]
    assert s.metrics.iterations == int(meta[0]) and s.metrics.converged
    ts = s.time_series
    got = np.array([ts.rel_iter_residual, ts.u_residual, ts.v_residual, ts.continuity_residual, ts.energy, ts.enstrophy,
                    ts.palinstrophy]).T
    assert got.shape == g["full_series"].shape == (1000, 7)
    # the trajectory is contractive: records agree to ~1e-12 early and stay within 1e-8 relative of the tiny late residuals
    assert relcols(got, g["full_series"]).max() < 1e-8
    m = s.metrics
    mine = np.array([m.final_residual, m.u_momentum_residual, m.v_momentum_residual, m.continuity_residual, m.final_energy,
                     m.final_enstrophy, m.final_palinstrophy])
    assert np.max(np.abs(mine - meta[2:]) / np.abs(meta[2:])) < 1e-8
    # chunked streaming (several ldc_solve_history calls) gives the same records as one call
    with built_lib.BatchedSpectralLDC(16, batch=2) as b:
        b.set_cases([100.0, 150.0]).init_state()
        r1, h1 = b.solve_history(5e-4, 3000, chunk=4096)
        b.init_state()
        r2, h2 = b.solve_history(5e-4, 3000, chunk=250)
        assert np.array_equal(r1.iterations, r2.iterations) and r1.iterations[0] != r1.iterations[1]
        for a, c, it in zip(h1, h2, r1.iterations):
            assert a.shape == (it, 7) and np.array_equal(a, c)
    s.close()


@pytest.mark.gpu
def test_history_off_uses_the_resident_kernel(built_lib, golden):
    s = built_lib.SGSolver(history=False, **KW)
    s.solve()
    assert s.metrics.iterations == int(golden.history["full_meta"][0])
    assert len(s.time_series.rel_iter_residual) == 1
    s.close()


@pytest.mark.gpu
def test_fused_history_kernel_matches_the_launch_per_stage_path(built_lib):
    """k_history_fused (one launch per chunk, everything in shared memory) against the launch-per-stage history path
    (forced with path='tiled'): same per-node arithmetic, so records agree to rounding of the reductions."""
    out = {}
    for path in ("auto", "tiled"):
        with built_lib.BatchedSpectralLDC(20, 14, batch=3, path=path) as b:   # non-square grid, three cavities
            b.set_cases([100.0, 400.0, 50.0], corner_smoothing=[0.15, 0.1, 0.2]).init_state()
            launches0 = b.launches
            r, h = b.solve_history(2e-3, 600, warmup=10, freeze_pressure=(path == "auto") or True, chunk=128)
            out[path] = (r, h, b.launches - launches0, b.get_state())
    ra, ha, la, sa = out["auto"]
    rt, ht, lt, st = out["tiled"]
    assert la < lt / 50                      # a handful of launches against ~22 per iteration
    assert np.array_equal(ra.iterations, rt.iterations) and np.array_equal(ra.converged, rt.converged)
    assert len(set(ra.iterations)) > 1       # the cavities stop at different iterations inside a chunk
    for a, c in zip(ha, ht):
        assert a.shape == c.shape and relcols(a, c).max() < 1e-11
    for a, c in zip(sa, st):
        assert np.max(np.abs(a - c)) < 1e-13


def test_sample_iterations_rule():
    from ldc_b200.batched import BatchedSpectralLDC as B

    assert B.sample_iterations(7).size == 0 and B.sample_iterations(10).size == 0
    assert B.sample_iterations(13).tolist() == [10, 11, 12]
    assert B.sample_iterations(1010).tolist() == list(range(10, 1010))
    idx = B.sample_iterations(14155)
    assert idx.size == 1000 and idx[0] == 10 and idx[-1] == 14154
    assert idx.tolist() == [10 + int(i) for i in np.linspace(0, 14144, 1000, dtype=int)]


@pytest.mark.gpu
@pytest.mark.parametrize("n,path", [(16, "auto"), (16, "tiled"), (40, "auto"), (72, "auto")])
def test_sampled_history_is_the_full_history_at_the_sampled_iterations(built_lib, n, path):
    """ldc_solve_sampled: same iteration counts and final state as a plain solve (bitwise: the record steps run on a copy),
    records equal to the every-iteration history at the kept iterations (two trajectories that differ by rounding)."""
    Re, tol, cap, pts = [100.0, 150.0, 70.0], 2e-3 if n < 72 else 2e-2, 4000, 37
    with built_lib.BatchedSpectralLDC(n, batch=3, path=path) as b:
        b.set_cases(Re).init_state()
        plain = b.solve(tol, cap, warmup=10, freeze_pressure=True)
        s_plain = b.get_state()
        b.init_state()
        rs, hs, idx = b.solve_history_sampled(tol, cap, max_points=pts)
        s_samp = b.get_state()
        b.init_state()
        rf, hf = b.solve_history(tol, cap)
    assert np.array_equal(rs.iterations, plain.iterations) and np.array_equal(rs.converged, plain.converged)
    assert plain.converged.all() and len(set(plain.iterations.tolist())) == 3
    for a, c in zip(s_plain, s_samp):
        assert np.array_equal(a, c)
    assert np.all(np.abs(rf.iterations - rs.iterations) <= 1)
    for b_ in range(3):
        want = built_lib.BatchedSpectralLDC.sample_iterations(plain.iterations[b_], 10, pts)
        assert np.array_equal(idx[b_], want) and hs[b_].shape == (pts, 7)
        ok = idx[b_] < rf.iterations[b_]
        assert ok.sum() >= pts - 1
        assert relcols(hs[b_][ok], hf[b_][idx[b_][ok]]).max() < 1e-8
        assert np.all(hs[b_][:, 4] > 0)      # every kept record was formed (no zero rows)


@pytest.mark.gpu
def test_sampled_history_edge_cases(built_lib):
    """ldc_solve_sampled at the ends of its domain: a record of iteration 0 and of the last iteration of a solve that
    stops at max_iter, a cavity that goes non-finite (its later rows stay zero), argument checks."""
    from ldc_b200 import _lib

    with built_lib.BatchedSpectralLDC(16, batch=2) as b:
        b.set_cases([100.0, 1000.0], corner_smoothing=[0.15, 0.01], CFL=[1.5, 50.0]).init_state()   # second one blows up
        full, hf = b.solve_history(1e-12, 300, nan_exit=True)
        assert full.diverged[1] and not full.diverged[0] and full.iterations[0] == 300 and full.iterations[1] < 300
        b.init_state()
        tol = np.full(2, 1e-12)
        rec = np.array([0, 1, 7, int(full.iterations[1]) - 1, int(full.iterations[1]) + 5, 299], dtype=np.int64)
        buf = np.zeros((2, rec.size, 7))
        iters, conv, divg, rel = np.zeros(2, np.int64), np.zeros(2, np.int32), np.zeros(2, np.int32), np.zeros(2)
        W = np.ascontiguousarray(b.level.W)

        def call(r, max_iter=300):
            return b.lib.ldc_solve_sampled(b._h, _lib.ptr(tol), max_iter, 10, 1, 1, _lib.ptr(W), r.size, _lib.ptr(r), _lib.ptr(buf),
                                           _lib.ptr(iters), _lib.ptr(conv), _lib.ptr(divg), _lib.ptr(rel))

        assert call(rec) == 0
        assert np.array_equal(iters, full.iterations) and divg[1] == 1 and divg[0] == 0
        for k, it in enumerate(rec):
            assert relcols(buf[0, k][None], hf[0][it][None]).max() < 1e-9
            if it < full.iterations[1] - 1:
                assert relcols(buf[1, k][None], hf[1][it][None]).max() < 1e-7      # a trajectory about to blow up
            elif it >= full.iterations[1]:
                assert not buf[1, k].any()                                          # iteration not taken: zero row
        for bad in (np.array([5, 5], dtype=np.int64), np.array([7, 3], dtype=np.int64), np.array([300], dtype=np.int64),
                    np.array([-1], dtype=np.int64)):
            assert call(bad) != 0 and "rec_iter" in b.lib.ldc_last_error().decode()


@pytest.mark.gpu
@pytest.mark.parametrize("n,path,batch", [(16, "auto", 1), (16, "auto", 3), (16, "tiled", 2), (40, "auto", 200), (72, "auto", 2),
                                          (56, "stream_tma", 3)])   # the last one: the cluster-resident kernel
def test_solve_cut_into_segments_takes_the_steps_of_the_uncut_solve(built_lib, n, path, batch):
    """ldc_solve_segment: iteration counts, flags, residuals and the final state of a solve cut into uneven segments (with
    the state read back in between) are bitwise those of ldc_solve."""
    Re = np.linspace(70.0, 150.0, batch)
    tol, cap = (2e-3 if n < 56 else 2e-2), 4000
    with built_lib.BatchedSpectralLDC(n, batch=batch, path=path) as b:
        b.set_cases(Re).init_state()
        whole = b.solve(tol, cap, warmup=10, freeze_pressure=True)
        s_whole = b.get_state()
        b.init_state()
        done, cuts = 0, [1, 7, 64, 65, 300, 1000, cap]
        for stop in cuts:
            part = b.solve_segment(tol, done, stop, cap, warmup=10, freeze_pressure=True)
            b.get_state()
            done = stop
        s_cut = b.get_state()
    assert whole.converged.all() and whole.iterations.max() > 65
    assert np.array_equal(part.iterations, whole.iterations) and np.array_equal(part.converged, whole.converged)
    assert np.array_equal(part.rel_residual, whole.rel_residual)
    for a, c in zip(s_whole, s_cut):
        assert np.array_equal(a, c)


@pytest.mark.gpu
def test_solve_segment_argument_checks(built_lib):
    from ldc_b200 import _lib

    with built_lib.BatchedSpectralLDC(16, batch=1) as b:
        b.set_cases(100.0).init_state()
        tol = np.full(1, 1e-3)
        for start, stop, cap in ((-1, 5, 10), (5, 3, 10), (0, 11, 10)):
            assert b.lib.ldc_solve_segment(b._h, _lib.ptr(tol), start, stop, cap, 0, 0, 1, None, None, None, None) != 0
            assert "ldc_solve_segment" in b.lib.ldc_last_error().decode()
        assert b.lib.ldc_solve_segment(b._h, None, 0, 5, 10, 0, 0, 1, None, None, None, None) != 0
        r = b.solve_segment(1e-3, 0, 0, 10)            # an empty first segment resets the status
        assert r.iterations[0] == 0 and not r.converged[0]
        r = b.solve_segment(1e-3, 0, 10, 10)
        assert r.iterations[0] == 10


@pytest.mark.gpu
@pytest.mark.parametrize("n,batch,seg,max_seg", [(16, 1, 64, 512), (16, 3, 64, 12), (16, 2, 32, 7), (40, 1, 64, 512),
                                                 (72, 1, 32, 512)])
def test_sampled_history_by_replay_equals_the_two_pass_route(built_lib, monkeypatch, n, batch, seg, max_seg):
    """The replay route (snapshots at segment starts, ONE batch of segments for the records) against the two-pass route:
    same SolveResult and final state (bitwise), same kept iterations, records equal to rounding; also with a segment
    limit small enough that snapshots are thinned on the way (even and odd limits)."""
    cls = built_lib.BatchedSpectralLDC
    monkeypatch.setattr(cls, "REPLAY_SEGMENT", seg)
    monkeypatch.setattr(cls, "REPLAY_MAX_SEGMENTS", max_seg)
    Re = np.linspace(80.0, 140.0, batch)
    tol, cap, pts = (2e-3 if n < 72 else 2e-2), 4000, 61
    with cls(n, batch=batch) as b:
        b.set_cases(Re).init_state()
        r2, h2, i2 = b.solve_history_sampled(tol, cap, max_points=pts, method="two_pass")
        s2 = b.get_state()
        b.init_state()
        r1, h1, i1 = b.solve_history_sampled(tol, cap, max_points=pts, method="replay")
        s1 = b.get_state()
    assert r2.converged.all() and r2.iterations.min() > 2 * seg      # at least three segments each
    assert np.array_equal(r1.iterations, r2.iterations) and np.array_equal(r1.converged, r2.converged)
    assert np.array_equal(r1.rel_residual, r2.rel_residual)
    for a, c in zip(s1, s2):
        assert np.array_equal(a, c)
    for b_ in range(batch):
        assert np.array_equal(i1[b_], i2[b_]) and h1[b_].shape == h2[b_].shape == (pts, 7)
        assert relcols(h1[b_], h2[b_]).max() < 1e-9
        assert np.all(h1[b_][:, 4] > 0)


@pytest.mark.gpu
def test_sampled_history_by_replay_edge_cases(built_lib, monkeypatch):
    """Replay route: a solve shorter than the record warm-up (no records), one that stops at max_iter inside a segment,
    a cavity that goes non-finite next to one that does not, and the byte limit that sends it back to two passes."""
    cls = built_lib.BatchedSpectralLDC
    monkeypatch.setattr(cls, "REPLAY_SEGMENT", 32)
    with cls(16, batch=2) as b:
        b.set_cases([100.0, 1000.0], corner_smoothing=[0.15, 0.01], CFL=[1.5, 50.0]).init_state()   # second one blows up
        r, h, idx = b.solve_history_sampled(1e-12, 7, nan_exit=True, method="replay")
        assert r.iterations.tolist() == [7, 7] and all(x.shape == (0, 7) for x in h)
        b.init_state()
        full, hf = b.solve_history(1e-12, 150, nan_exit=True)
        assert full.diverged[1] and full.iterations[0] == 150 and 20 < full.iterations[1] < 150
        for limit in (cls.REPLAY_MAX_BYTES, 1):
            monkeypatch.setattr(cls, "REPLAY_MAX_BYTES", limit)
            b.init_state()
            r, h, idx = b.solve_history_sampled(1e-12, 150, nan_exit=True, max_points=25, method="replay")
            assert np.array_equal(r.iterations, full.iterations) and np.array_equal(r.diverged, full.diverged)
            assert idx[0][-1] == 149 and idx[1][-1] == full.iterations[1] - 1
            assert relcols(h[0], hf[0][idx[0]]).max() < 1e-9
            early = idx[1] < full.iterations[1] // 2
            assert early.any() and relcols(h[1][early], hf[1][idx[1][early]]).max() < 1e-6   # a trajectory that blows up
    with pytest.raises(ValueError):
        b.solve_history_sampled(1e-3, 10, method="nope")
