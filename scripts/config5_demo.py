"""BASELINE config 5: Re = 5000, N = 128, coarse-to-fine spectral multigrid (FSG 32 -> 64 -> 128), a batch of cavities with
Re jitter, sharded over the ranks (8 cavities per GPU on a full box):

    torchrun --nproc-per-node 8 scripts/config5_demo.py --batch 64
    python scripts/config5_demo.py --batch 8                               # one GPU's share

Reports wall time, per-level iteration counts and convergence; the fields are checked for the physics the reference's
plots show at this Re (primary vortex strength, secondary eddies) through the streamfunction metrics.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "02689-advancednumericalalgorithmp3_b200"))

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--N", type=int, default=128)
ap.add_argument("--Re", type=float, default=5000.0)
ap.add_argument("--levels", type=int, default=3)
ap.add_argument("--tol", type=float, default=1e-6)
ap.add_argument("--max-iter", type=int, default=400000)
ap.add_argument("--out", default="gpurun_out/config5_demo.json")
args = ap.parse_args()

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
if rank == 0:
    import __graft_entry__ as ge

    ge.build()
import torch

if world > 1:
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    torch.distributed.init_process_group("nccl")
from ldc_b200 import operators as ops
from ldc_b200 import postprocess
from ldc_b200.sharded import solve_sweep

if world > 1:
    torch.distributed.barrier()   # NCCL's communicator comes up here, not inside the timed solve
Re = args.Re * (1.0 + 0.02 * np.random.default_rng(1).uniform(-1, 1, args.batch))   # seed 1, 2 % jitter
torch.cuda.synchronize()
t0 = time.time()
r = solve_sweep(args.N, Re, 0.15, args.tol, args.max_iter, kind="fsg", fields="rank0", n_levels=args.levels)
torch.cuda.synchronize()
wall = time.time() - t0
if rank == 0:
    lvl = ops.build_level(args.N, args.N)
    psi = [postprocess.streamfunction_vortex_metrics(lvl, r.u[b], r.v[b]) for b in range(min(2, args.batch))]
    out = dict(config=dict(N=args.N, Re=args.Re, batch=args.batch, levels=ops.hierarchy_orders(args.N, args.levels),
                           tolerance=args.tol, max_iter_per_level=args.max_iter, n_gpus=world),
               wall_s=wall, kernel_ms_max_over_ranks=r.kernel_ms, iterations=r.iterations.tolist(),
               converged=int(r.converged.sum()), diverged=int(r.diverged.sum()), rel_residual_max=float(r.rel_residual.max()),
               cavity_steps_per_s_all_levels=float(r.iterations.sum() / wall),
               vortex_metrics_first=[{k: float(v) for k, v in m.items()} for m in psi])
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
    print(json.dumps(out))
if world > 1:
    torch.distributed.destroy_process_group()
