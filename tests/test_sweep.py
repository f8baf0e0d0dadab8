"""Batch-aware sweep front-end: Hydra-style override expansion, grouping into batched solves,
objectives (main.py:142-225) and the fan-out of per-trial records.  CPU tests use a stand-in solver; the GPU test runs a
small real sweep and checks it against the drop-in solver class trial by trial."""
import math
import os

import numpy as np
import pytest

from ldc_b200 import sweep


def test_expand_overrides_is_the_hydra_product():
    t = sweep.expand_overrides(["solver=spectral/fsg", "N=30,40", "Re=100,400,1000", "+experiment/optimization=corner_smoothing"])
    assert len(t) == 6 and [x["N"] for x in t] == [30, 30, 30, 40, 40, 40] and [x["Re"] for x in t][:3] == [100, 400, 1000]
    t = sweep.expand_overrides(["N=range(16,33,8)", "solver.corner_smoothing=interval(0.02,0.35)"], n_trials=4)
    assert len(t) == 12 and t[0]["solver.corner_smoothing"] == 0.02 and t[3]["solver.corner_smoothing"] == 0.35
    a = sweep.expand_overrides(["solver.corner_smoothing=interval(0.02,0.35)"], n_trials=5, sampler="random", seed=3)
    b = sweep.expand_overrides(["solver.corner_smoothing=interval(0.02,0.35)"], n_trials=5, sampler="random", seed=3)
    assert a == b and all(0.02 <= x["solver.corner_smoothing"] <= 0.35 for x in a)
    with pytest.raises(ValueError):
        sweep.expand_overrides(["solver.corner_smoothing=interval(0.02,0.35)"])


def test_solver_kwargs_resolve_like_the_yaml():
    kw = sweep.solver_kwargs({"solver": "spectral/fsg", "N": 30, "Re": 1000, "solver.corner_smoothing": 0.07,
                              "max_iterations": 500000})
    assert kw["nx"] == kw["ny"] == 30 and kw["Re"] == 1000.0 and kw["corner_smoothing"] == 0.07 and kw["n_levels"] == 2
    assert kw["max_iterations"] == 500000 and kw["name"] == "spectral_fsg" and kw["CFL"] == 1.5
    # the kwargs are exactly what the drop-in class accepts
    from ldc_b200.datastructures import SpectralParameters

    SpectralParameters(**{k: v for k, v in kw.items() if not k.startswith("_")})
    assert sweep.solver_kwargs({"solver": "spectral/sg"})["multigrid"] == "none"
    with pytest.raises(ValueError):
        sweep.solver_kwargs({"solver": "fv"})


def test_grouping_shares_one_solve_per_shape():
    trials = sweep.expand_overrides(["N=16,24", "Re=100,400", "solver.corner_smoothing=0.1,0.2"])
    groups = sweep.group_trials([sweep.solver_kwargs(t) for t in trials])
    assert [len(g) for g in groups] == [4, 4] and groups[0] == [0, 1, 2, 3]


def test_objectives(tmp_path):
    assert sweep.fv_l2_objective({"u_L2_error": 3.0, "v_L2_error": 4.0}) == 5.0
    assert sweep.fv_l2_objective({}) == float("inf")
    d = tmp_path / "data" / "validation" / "botella"
    d.mkdir(parents=True)
    (d / "botella_Re400_vortex.csv").write_text("# comment\nRe,psi_min,psi_min_x,psi_min_y\n400,-0.1,0.5,0.6\n")
    (d / "botella_Re1000_vortex.csv").write_text("Re,psi_primary,x_primary,y_primary\n1000,0.1189,0.4692,0.5652\n")
    m = sweep.Metrics(psi_min=-0.11, psi_min_x=0.53, psi_min_y=0.56)
    want = math.sqrt((0.1**2 + 0.03**2 + 0.04**2) / 3)
    assert abs(sweep.botella_vortex_objective(m, 400, str(tmp_path)) - want) < 1e-15
    assert sweep.botella_vortex_objective(m, 1000, str(tmp_path)) == float("inf")   # the reference's column-name quirk
    assert sweep.botella_vortex_objective(m, 100, str(tmp_path)) == float("inf")    # no table
    with pytest.raises(ValueError):
        sweep.objective_value("multi", {}, m, 400)


def _fake_sweep(N, Re, cs, tol, max_iter, *, kind, fields, device, **kw):
    from ldc_b200.sharded import SweepResult

    B, n = Re.size, N + 1
    u = np.zeros((B, n, n)) + Re[:, None, None]
    return SweepResult((Re * 10).astype(np.int64), Re < 900, Re >= 900, cs * 1e-6, u, u.copy(), np.zeros((B, N - 1, N - 1)),
                       1.0, [(0, B)])


def test_run_trials_fans_results_out_in_trial_order(tmp_path):
    trials = sweep.expand_overrides(["solver=spectral/fsg", "N=8,12", "Re=100,1000", "solver.corner_smoothing=0.1,0.2"])
    res = sweep.run_trials(trials, "fv_l2_error", root=str(tmp_path), vortex_metrics=False, solve=_fake_sweep)
    assert [r.index for r in res] == list(range(8)) and {r.group for r in res} == {0, 1}
    for r, t in zip(res, trials):
        assert r.params["nx"] == t["N"] and r.params["Re"] == t["Re"] and r.metrics["iterations"] == t["Re"] * 10
        assert r.metrics["converged"] == (1 if t["Re"] < 900 else 0) and r.extras["diverged"] == (t["Re"] >= 900)
        assert r.validation_errors == {} and r.objective == float("inf")
        assert abs(r.metrics["final_residual"] - t["solver.corner_smoothing"] * 1e-6) < 1e-20


@pytest.mark.gpu
def test_small_real_sweep_matches_the_solver_class(built_lib, golden, tmp_path):
    from ldc_b200 import vts

    g = golden.validation
    d = tmp_path / "data" / "validation" / "fv" / "Re100"
    d.mkdir(parents=True)
    vts.write_vts(str(d / "solution.vts"), np.column_stack([g["x"], g["y"], np.zeros(g["x"].size)]), (128, 128, 1),
                  {"u": g["ref_u"], "v": g["ref_v"]})
    trials = sweep.expand_overrides(["solver=spectral/fsg", "N=24", "Re=100,200", "solver.corner_smoothing=0.1,0.15"])
    res = sweep.run_trials(trials, "fv_l2_error", root=str(tmp_path))
    assert len(res) == 4 and all(r.metrics["converged"] == 1 for r in res)
    # cs = 0.15, Re = 100 is the fixture's solve: the reference's errors and iteration count
    r = res[1]
    meta = golden.solves["fsg_Re100_N24_meta"]
    assert r.metrics["iterations"] == int(meta[0])
    assert abs(r.validation_errors["u_L2_error"] - g["errors"][0]) < 1e-9
    assert abs(r.objective - math.hypot(g["errors"][0], g["errors"][1])) < 1e-9
    assert res[2].validation_errors == {} and res[2].objective == float("inf")   # no stored solution for Re=200
    assert r.metrics["psi_min"] < -0.09
    # trial by trial against the drop-in class
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        for t, rr in zip(trials, res):
            kw = {k: v for k, v in sweep.solver_kwargs(t).items() if not k.startswith("_")}
            s = built_lib.FSGSolver(**kw)
            s.solve()
            assert s.metrics.iterations == rr.metrics["iterations"]
            assert abs(s.metrics.psi_min - rr.metrics["psi_min"]) < 1e-12
            assert s.compute_validation_errors() == pytest.approx(rr.validation_errors, rel=1e-12)
            s.close()
    finally:
        os.chdir(cwd)


def _fake_sweep_sharded(N, Re, cs, tol, max_iter, *, kind, fields, device, **kw):
    import torch.distributed as dist

    from ldc_b200.sharded import shard_bounds

    r = _fake_sweep(N, Re, cs, tol, max_iter, kind=kind, fields=fields, device=device, **kw)
    r.bounds = shard_bounds(Re.size, dist.get_world_size())
    return r


def _sweep_worker(rank, world, port, root, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        trials = sweep.expand_overrides(["N=8", "Re=100,400,1000", "solver.corner_smoothing=0.1,0.2,0.3"])
        res = sweep.run_trials(trials, "fv_l2_error", root=root, vortex_metrics=False, solve=_fake_sweep_sharded)
        q.put((rank, [(r.index, r.metrics["iterations"], r.objective, r.validation_errors) for r in res]))
    finally:
        dist.destroy_process_group()


def test_post_processing_is_sharded_and_gathered_over_two_gloo_ranks(tmp_path):
    """Each rank post-processes its own block of the batch; the per-trial records are all-gathered in trial order."""
    import socket

    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sweep_worker, args=(r, 2, port, str(tmp_path), q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=120) for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got[0][1] == got[1][1] and [x[0] for x in got[0][1]] == list(range(9))
    assert [x[1] for x in got[0][1]] == [1000, 1000, 1000, 4000, 4000, 4000, 10000, 10000, 10000]
