"""Sweeps of independent cavities sharded over the GPUs of one node.

Cavities never talk to each other, so each rank solves a contiguous block of the batch on its own GPU with
no collective on the stepping path.  `torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU
tests) is used once, at the end: an all-gather of the per-cavity results (iteration counts, flags, residuals)
and a gather of the fields.  One process per GPU, launched with torchrun.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .fsg import FSGResult, solve_fsg_batched, solve_sg_batched


def shard_bounds(n_cases: int, world: int):
    """Contiguous blocks [lo, hi) per rank; the first n_cases % world ranks get one extra cavity."""
    base, extra = divmod(int(n_cases), int(world))
    bounds, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        bounds.append((lo, hi))
        lo = hi
    return bounds


@dataclass
class SweepResult:
    iterations: np.ndarray   # int64 [B]
    converged: np.ndarray    # bool [B]
    diverged: np.ndarray     # bool [B]
    rel_residual: np.ndarray  # float64 [B]
    u: np.ndarray | None     # [B, n, n] on the ranks that asked for fields, else None
    v: np.ndarray | None
    p: np.ndarray | None
    kernel_ms: float         # max over ranks
    bounds: list
    per_rank_kernel_ms: np.ndarray | None = None   # [world], when asked for


def _dist():
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return None
    return dist


def _comm_device():
    import torch

    dist = _dist()
    if dist is not None and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def _allgather_blocks(local: np.ndarray, counts, dtype):
    """All-gather of per-rank blocks with different leading sizes (padded to the largest block)."""
    import torch

    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return np.ascontiguousarray(local)
    dev = _comm_device()
    world, pad = dist.get_world_size(), max(counts)
    tail = tuple(local.shape[1:])
    buf = torch.zeros((pad,) + tail, dtype=dtype, device=dev)
    if local.shape[0]:
        buf[:local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local)).to(dev)
    out = torch.empty((world * pad,) + tail, dtype=dtype, device=dev)
    dist.all_gather_into_tensor(out, buf)
    out = out.cpu().numpy().reshape((world, pad) + tail)
    return np.concatenate([out[r, :counts[r]] for r in range(world)], axis=0)


def _gather_fields(tensors, counts, dst_all: bool):
    """Gather of the per-rank field blocks (torch tensors on the communication device, padded to the largest block) to
    rank 0 -- `dist.gather`, SURVEY 8(e) -- or to every rank (`all_gather_into_tensor`).  Returns NumPy arrays where the
    fields are wanted, else None; one device -> host copy per field on the receiving ranks only."""
    import torch

    dist = _dist()
    world, rank, pad = dist.get_world_size(), dist.get_rank(), max(counts)
    out = []
    for t in tensors:
        if dst_all:
            full = torch.empty((world * pad,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(full, t)
            parts = full.reshape((world, pad) + tuple(t.shape[1:]))
            out.append(np.concatenate([parts[r, :counts[r]].cpu().numpy() for r in range(world)], axis=0))
        else:
            recv = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
            dist.gather(t, recv, dst=0)
            out.append(np.concatenate([recv[r][:counts[r]].cpu().numpy() for r in range(world)], axis=0) if rank == 0 else None)
    return out


def warm_up_gather():
    """Bring up NCCL's point-to-point connections to rank 0 (they are made lazily, ~0.1-0.2 s per peer, on the first
    `dist.gather`): call once before a timed sweep so that the sweep's clock does not include communicator set-up."""
    import torch

    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return
    t = torch.zeros(8, dtype=torch.float64, device=_comm_device())
    recv = [torch.empty_like(t) for _ in range(dist.get_world_size())] if dist.get_rank() == 0 else None
    dist.gather(t, recv, dst=0)
    if t.is_cuda:
        torch.cuda.synchronize()


def interleave_permutation(n_cases: int, world: int) -> np.ndarray:
    """Order in which the cavities are laid out before the contiguous split so that rank r gets cavities r, r + world,
    r + 2 world, ...: a sweep ordered by Re or corner_smoothing then hands every rank the same mix of short and long solves
    (the block sizes are those of `shard_bounds`)."""
    return np.concatenate([np.arange(r, n_cases, world) for r in range(world)]) if n_cases else np.zeros(0, dtype=np.int64)


def solve_sweep(N, Re, corner_smoothing, tolerance, max_iter, *, kind: str = "fsg", fields: str = "rank0",
                local_solver=None, device=None, balance: str = "block", per_rank_kernel_ms: bool = False,
                **solver_kw) -> SweepResult:
    """Solve B = len(Re) cavities (per-cavity Re and corner_smoothing; everything in solver_kw is shared),
    sharded over the ranks of the initialised process group (or on this process alone if there is none).

    kind   : "fsg" (solve_fsg_batched) or "sg" (solve_sg_batched)
    fields : "rank0" gathers u, v, p on rank 0 only, "all" on every rank, "none" skips the field gather
    local_solver : override of the per-rank solver (tests); signature of solve_fsg_batched
    balance : "block" -- rank r solves the contiguous block `shard_bounds` gives it; "interleave" -- rank r solves
              cavities r, r + world, ... (`interleave_permutation`); results come back in the caller's order either way
    """
    import torch

    Re = np.atleast_1d(np.asarray(Re, dtype=np.float64))
    cs = np.broadcast_to(np.asarray(corner_smoothing, dtype=np.float64), Re.shape).copy()
    tol = np.broadcast_to(np.asarray(tolerance, dtype=np.float64), Re.shape).copy()
    dist = _dist()
    rank = dist.get_rank() if dist else 0
    world = dist.get_world_size() if dist else 1
    if balance not in ("block", "interleave"):
        raise ValueError("balance must be 'block' or 'interleave'")
    n_all = Re.size
    perm = interleave_permutation(n_all, world) if balance == "interleave" and world > 1 else None
    if perm is not None:
        Re, cs, tol = Re[perm], cs[perm], tol[perm]
        solver_kw = {k: (([val[i] for i in perm] if isinstance(val, (list, tuple)) else np.asarray(val)[perm])
                         if not isinstance(val, str) and np.ndim(val) == 1 and len(val) == n_all else val)
                     for k, val in solver_kw.items()}
    bounds = shard_bounds(Re.size, world)
    lo, hi = bounds[rank]
    counts = [b - a for a, b in bounds]
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    fn = local_solver or (solve_fsg_batched if kind == "fsg" else solve_sg_batched)
    # per-cavity keyword inputs (beta_squared, CFL, lid_velocity arrays; a list of corner treatments) follow the shard
    local_kw = {}
    for k, val in solver_kw.items():
        per_case = not isinstance(val, str) and np.ndim(val) == 1 and len(val) == Re.size
        local_kw[k] = (list(val[lo:hi]) if isinstance(val, (list, tuple)) else np.asarray(val)[lo:hi]) if per_case else val
    n1 = N + 1
    # with NCCL the fields stay on the device from the solver's state to the gather (no host round trip on the senders)
    dev_route = (dist is not None and world > 1 and dist.get_backend() == "nccl" and fields != "none"
                 and local_solver is None)
    if dev_route:
        local_kw.update(keep_handle=True, host_fields=False)
    if hi > lo:
        r: FSGResult = fn(N, Re[lo:hi], tol[lo:hi], max_iter, corner_smoothing=cs[lo:hi], device=device, **local_kw)
    else:  # more ranks than cavities
        z = np.zeros
        r = FSGResult(z((0, n1, n1)), z((0, n1, n1)), z((0, N - 1, N - 1)), z(0, dtype=np.int64), z(0, dtype=bool),
                      z(0, dtype=bool), z(0))
    scal = np.stack([r.iterations.astype(np.float64), r.converged.astype(np.float64), r.diverged.astype(np.float64),
                     r.rel_residual], axis=1) if hi > lo else np.zeros((0, 4))
    scal = _allgather_blocks(scal, counts, torch.float64)
    kms = np.array([[r.kernel_ms]])
    kms_all = _allgather_blocks(kms, [1] * world, torch.float64).ravel()
    kms = kms_all.max()
    u = v = p = None
    if dev_route:
        dev = _comm_device()
        pad, cnt = max(counts), hi - lo
        bufs = [torch.zeros((pad, n1, n1), dtype=torch.float64, device=dev), torch.zeros((pad, n1, n1), dtype=torch.float64, device=dev),
                torch.zeros((pad, N - 1, N - 1), dtype=torch.float64, device=dev)]
        if cnt:
            r.finest.get_state(tuple(b[:cnt] for b in bufs))
            if r.zero_mask is not None and r.zero_mask.any():
                zm = torch.from_numpy(np.ascontiguousarray(r.zero_mask)).to(dev)
                for b in bufs:
                    b[:cnt][zm] = 0.0
            r.finest.close()
        u, v, p = _gather_fields(bufs, counts, dst_all=(fields == "all"))
    elif fields != "none":
        # NCCL has no variable-size gather: all-gather the padded blocks, keep them where asked
        gu = _allgather_blocks(r.u, counts, torch.float64)
        gv = _allgather_blocks(r.v, counts, torch.float64)
        gp = _allgather_blocks(r.p, counts, torch.float64)
        if fields == "all" or rank == 0:
            u, v, p = gu, gv, gp
    if perm is not None:   # back to the caller's order
        inv = np.argsort(perm)
        scal = scal[inv]
        u, v, p = (None if a is None else a[inv] for a in (u, v, p))
    return SweepResult(np.rint(scal[:, 0]).astype(np.int64), scal[:, 1] > 0.5, scal[:, 2] > 0.5, scal[:, 3], u, v, p,
                       float(kms), bounds, kms_all if per_rank_kernel_ms else None)
