"""BASELINE config 4 in miniature or in full: a corner_smoothing x Re sweep of FSG solves sharded over the ranks
of a torchrun job (NCCL), checked against an unsharded solve of a few cavities on rank 0.

    torchrun --nproc-per-node 2 scripts/sweep_demo.py --N 32 --grid 8          # 64 cavities, N=32
    torchrun --nproc-per-node 8 scripts/sweep_demo.py --N 64 --grid 32         # 1024 cavities, N=64 (config 4)
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, ".")
sys.path.insert(0, "02689-advancednumericalalgorithmp3_b200")

ap = argparse.ArgumentParser()
ap.add_argument("--N", type=int, default=32)
ap.add_argument("--grid", type=int, default=8)
ap.add_argument("--max-iter", type=int, default=2_000_000)
ap.add_argument("--tol", type=float, default=1e-6)
ap.add_argument("--balance", default="block", choices=["block", "interleave"])
ap.add_argument("--out", default="gpurun_out/sweep_demo.json")
args = ap.parse_args()

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if "RANK" in os.environ:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank = dist.get_rank() if dist.is_initialized() else 0
world = dist.get_world_size() if dist.is_initialized() else 1
if rank == 0:
    import __graft_entry__ as ge
    ge.build()
if world > 1:
    dist.barrier()
import ldc_b200
from ldc_b200.sharded import solve_sweep

cs, Re = np.meshgrid(np.linspace(0.02, 0.35, args.grid), np.linspace(100, 1000, args.grid), indexing="ij")
cs, Re = cs.ravel(), Re.ravel()
torch.cuda.synchronize()
t0 = time.perf_counter()
r = solve_sweep(args.N, Re, cs, args.tol, args.max_iter, kind="fsg", n_levels=2, coarse_tolerance_factor=1.0, balance=args.balance)
torch.cuda.synchronize()
wall = time.perf_counter() - t0
if rank == 0:
    # spot check: three cavities solved alone (AUTO may pick another kernel for a batch of one: same iteration count,
    # fields equal to rounding) must reproduce the sharded result
    pick = [0, Re.size // 2, Re.size - 1]
    ok, worst = True, 0.0
    for k in pick:
        one = ldc_b200.solve_fsg_batched(args.N, Re[k:k + 1], args.tol, args.max_iter, corner_smoothing=cs[k:k + 1],
                                         n_levels=2, coarse_tolerance_factor=1.0, device=local)
        ok &= int(one.iterations[0]) == int(r.iterations[k])
        worst = max(worst, float(np.linalg.norm(one.u[0] - r.u[k]) / np.linalg.norm(r.u[k])))
    ok &= worst < 1e-10
    out = dict(N=args.N, cavities=int(Re.size), world=world, balance=args.balance, wall_s=wall, kernel_ms_max=r.kernel_ms,
               converged=int(r.converged.sum()), diverged=int(r.diverged.sum()),
               total_iterations=int(r.iterations.sum()), iters_min=int(r.iterations.min()), iters_max=int(r.iterations.max()),
               cavity_steps_per_s=float(r.iterations.sum() / wall), spot_check=bool(ok), spot_check_rel_l2=worst, bounds=r.bounds)
    print(json.dumps(out))
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
