"""Host logic of the multi-GPU sweep (ldc_b200/sharded.py) on CPU: world_size-2 gloo process group with a
deterministic stand-in for the per-rank GPU solver.  The GPU numerics are covered by the -m gpu tests; here the
subject is the sharding, the padded all-gather and the result ordering."""
import os
import socket

import numpy as np
import pytest

from ldc_b200.sharded import shard_bounds


def test_shard_bounds_cover_the_batch():
    assert shard_bounds(1024, 8) == [(i * 128, (i + 1) * 128) for i in range(8)]
    assert shard_bounds(5, 2) == [(0, 3), (3, 5)]
    assert shard_bounds(1, 4) == [(0, 1), (1, 1), (1, 1), (1, 1)]
    for B, W in ((7, 3), (64, 8), (3, 8)):
        b = shard_bounds(B, W)
        assert b[0][0] == 0 and b[-1][1] == B and all(b[i][1] == b[i + 1][0] for i in range(W - 1))


def _fake_solver(N, Re, tol, max_iter, *, corner_smoothing, device, **kw):
    """Stand-in with the signature / result type of solve_fsg_batched: every output is a function of the case."""
    from ldc_b200.fsg import FSGResult

    Re = np.asarray(Re)
    B, n = Re.size, N + 1
    base = np.arange(n * n, dtype=np.float64).reshape(n, n)
    u = Re[:, None, None] * base
    v = corner_smoothing[:, None, None] * base
    p = (Re + corner_smoothing)[:, None, None] * np.ones((N - 1, N - 1))
    div = Re > 900
    if "CFL" in kw:  # per-cavity keyword inputs must arrive sliced like Re
        assert np.shape(kw["CFL"]) == Re.shape and len(kw["corner_treatment"]) == B
        p = p + np.asarray(kw["CFL"])[:, None, None]
    return FSGResult(u, v, p, (10 * Re).astype(np.int64), ~div, div, np.where(div, np.nan, 1e-7 * Re), kernel_ms=float(B))


def _worker(rank, world, port, B, q, balance="block"):
    import torch.distributed as dist

    from ldc_b200.sharded import solve_sweep

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        Re = np.linspace(100, 1000, B)
        cs = np.linspace(0.02, 0.35, B)
        kw = dict(CFL=np.linspace(1.0, 2.0, B), corner_treatment=["smoothing"] * B, n_levels=2) if B == 5 else {}
        r = solve_sweep(6, Re, cs, 1e-6, 100, fields="rank0", local_solver=_fake_solver, balance=balance, **kw)
        q.put((rank, r.iterations, r.converged, r.diverged, r.rel_residual, None if r.u is None else r.u.copy(),
               None if r.p is None else r.p.copy(), r.kernel_ms, r.bounds))
    finally:
        dist.destroy_process_group()


def test_interleave_permutation():
    from ldc_b200.sharded import interleave_permutation

    assert interleave_permutation(7, 3).tolist() == [0, 3, 6, 1, 4, 2, 5]
    assert interleave_permutation(4, 1).tolist() == [0, 1, 2, 3] and interleave_permutation(0, 2).size == 0
    for B, W in ((1024, 8), (5, 2), (1, 4), (13, 5)):
        perm, lo = interleave_permutation(B, W), 0
        assert sorted(perm.tolist()) == list(range(B))
        for r, (a, b) in enumerate(shard_bounds(B, W)):   # the block sizes are those of the contiguous split
            assert perm[a:b].tolist() == list(range(r, B, W))


@pytest.mark.parametrize("B,balance", [(5, "block"), (8, "block"), (1, "block"), (5, "interleave"), (8, "interleave")])
def test_sweep_sharded_over_two_gloo_ranks_equals_unsharded(B, balance):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q, balance)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=120) for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    Re = np.linspace(100, 1000, B)
    cs = np.linspace(0.02, 0.35, B)
    kw = dict(CFL=np.linspace(1.0, 2.0, B), corner_treatment=["smoothing"] * B, n_levels=2) if B == 5 else {}
    ref = _fake_solver(6, Re, None, 100, corner_smoothing=cs, device=0, **kw)
    for rank, it, conv, div, rel, u, p, kms, bounds in got:
        assert np.array_equal(it, ref.iterations) and np.array_equal(conv, ref.converged) and np.array_equal(div, ref.diverged)
        assert np.array_equal(np.isnan(rel), np.isnan(ref.rel_residual))
        assert np.allclose(rel[~np.isnan(rel)], ref.rel_residual[~np.isnan(rel)])
        assert bounds == shard_bounds(B, 2)
        assert kms == max(hi - lo for lo, hi in bounds)
        if rank == 0:
            assert np.array_equal(u, ref.u) and np.array_equal(p, ref.p)
        else:
            assert u is None and p is None


def test_sweep_without_process_group_runs_locally():
    from ldc_b200.sharded import solve_sweep

    Re = np.array([100.0, 400.0, 950.0])
    r = solve_sweep(6, Re, 0.15, 1e-6, 100, local_solver=_fake_solver)
    assert r.bounds == [(0, 3)] and list(r.iterations) == [1000, 4000, 9500] and list(r.diverged) == [False, False, True]
    assert r.u.shape == (3, 7, 7)


def _gather_worker(rank, world, port, q):
    """The device-side gather route of solve_sweep (`_gather_fields`, used with NCCL) on gloo / CPU tensors: padded blocks of
    different sizes to rank 0 (`dist.gather`) and to every rank, the point-to-point warm-up, per-rank kernel times."""
    import torch
    import torch.distributed as dist

    from ldc_b200.sharded import _gather_fields, solve_sweep, warm_up_gather

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        warm_up_gather()
        counts = [3, 2]
        pad = max(counts)
        t = torch.zeros((pad, 4, 4), dtype=torch.float64)
        t[:counts[rank]] = torch.arange(counts[rank] * 16, dtype=torch.float64).reshape(counts[rank], 4, 4) + 100 * rank
        to0 = _gather_fields([t], counts, dst_all=False)[0]
        toall = _gather_fields([t], counts, dst_all=True)[0]
        r = solve_sweep(6, np.linspace(100, 800, 5), 0.1, 1e-6, 100, fields="none", local_solver=_fake_solver,
                        per_rank_kernel_ms=True)
        q.put((rank, None if to0 is None else to0.copy(), toall.copy(), r.per_rank_kernel_ms.copy(), r.u))
    finally:
        dist.destroy_process_group()


def test_gather_route_and_per_rank_kernel_times_over_gloo():
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=120) for _ in procs), key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = np.concatenate([np.arange(48.0).reshape(3, 4, 4), np.arange(32.0).reshape(2, 4, 4) + 100])
    for rank, to0, toall, kms, u in got:
        assert np.array_equal(toall, want)
        assert (to0 is None) == (rank != 0) and (rank != 0 or np.array_equal(to0, want))
        assert kms.tolist() == [3.0, 2.0] and u is None   # the stand-in reports its block size as kernel time
