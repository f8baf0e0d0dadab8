"""Batch-aware sweep front-end.

The reference runs a parameter sweep as one Hydra MULTIRUN job per trial (`main.py -m Re=100,400 ...`, joblib / LSF
launchers, `conf/machine/local.yaml:9`) or as an Optuna study driven by hydra-optuna-sweeper
(`conf/hydra/sweeper/optuna_corner.yaml`, `conf/experiment/optimization/corner_smoothing.yaml`): every trial builds its own
solver, solves one cavity, computes its validation errors and returns `compute_optuna_objective` (main.py:206-252).

Here the trials of a sweep are collected first, grouped by what a batch must share (solver kind, N, level hierarchy,
transfer method, iteration cap), and each group is ONE batched solve -- sharded over the ranks of the process group when
there is one (`sharded.solve_sweep`) -- followed by the batched validation-error kernel and the per-cavity vortex
metrics.  Results fan back out as one record per trial carrying what `run_solver` logs per run (params, Metrics,
validation errors, objective), optionally to MLflow child runs when MLflow is installed.

    python -m ldc_b200.sweep solver=spectral/fsg N=32 Re=100,400 solver.corner_smoothing=0.05,0.1,0.15 --out sweep.json
    python -m ldc_b200.sweep solver=spectral/fsg N=30 Re=1000 'solver.corner_smoothing=interval(0.02,0.35)' \
        --n-trials 64 optuna.objective=botella_vortex

Hydra, Optuna and MLflow are not needed (none is installed on the GPU box); `BatchObjective` is the hook for an Optuna
study where it is: ask a batch of trials, evaluate them in one call, tell the values.
"""
from __future__ import annotations

import csv
import itertools
import json
import math
import os
import re
import time
from dataclasses import asdict, dataclass, field

import numpy as np

from .datastructures import Metrics

# conf/config.yaml + conf/problem/ldc.yaml + conf/solver/spectral/{sg,fsg}.yaml
GLOBAL_DEFAULTS = dict(N=32, Re=100.0, lid_velocity=1.0, Lx=1.0, Ly=1.0, tolerance=1.0e-6, max_iterations=10_000_000)
SOLVER_DEFAULTS = {
    "spectral/sg": dict(name="spectral", basis_type="chebyshev", CFL=1.5, beta_squared=5.0, corner_treatment="smoothing",
                        corner_smoothing=0.15, multigrid="none"),
    "spectral/fsg": dict(name="spectral_fsg", basis_type="chebyshev", CFL=1.5, beta_squared=5.0,
                         corner_treatment="smoothing", corner_smoothing=0.15, multigrid="fsg", n_levels=2,
                         coarse_tolerance_factor=1.0, prolongation_method="fft", restriction_method="fft"),
}


# ------------------------------------------------------------------------------------------------------------------
# Hydra-style override parsing
# ------------------------------------------------------------------------------------------------------------------
def _scalar(tok: str):
    tok = tok.strip().strip("'\"")
    for cast in (int, float):
        try:
            return cast(tok)
        except ValueError:
            pass
    return {"true": True, "false": False, "null": None}.get(tok.lower(), tok)


@dataclass
class Interval:
    """`interval(lo, hi)` of the Optuna sweeper: a continuous range that a sampler draws from."""

    lo: float
    hi: float


def parse_value(text: str):
    """One override value -> scalar, list of scalars (a sweep axis) or Interval."""
    text = text.strip()
    m = re.fullmatch(r"interval\(\s*([^,]+),\s*([^)]+)\)", text)
    if m:
        return Interval(float(m.group(1)), float(m.group(2)))
    m = re.fullmatch(r"range\(\s*([^)]*)\)", text)
    if m:
        a = [int(x) for x in m.group(1).split(",")]
        return list(range(*a))
    m = re.fullmatch(r"choice\(\s*([^)]*)\)", text)
    if m:
        return [_scalar(x) for x in m.group(1).split(",")]
    if "," in text:
        return [_scalar(x) for x in text.split(",")]
    return _scalar(text)


def expand_overrides(overrides, n_trials: int = 0, sampler: str = "grid", seed: int = 0):
    """Hydra MULTIRUN semantics: the cartesian product of every comma list, in override order (last axis fastest).
    Intervals are sampled `n_trials` times: sampler="grid" -> linspace over the interval, "random" -> uniform draws
    (what a TPE study starts from); each sample is combined with the full product of the list axes."""
    fixed, axes, intervals = {}, [], []
    for ov in overrides:
        if ov.startswith("+") or "=" not in ov:   # "+experiment/...": config-group additions carry no value here
            continue
        key, val = ov.split("=", 1)
        v = parse_value(val)
        if isinstance(v, Interval):
            intervals.append((key, v))
        elif isinstance(v, list):
            axes.append((key, v))
        else:
            fixed[key] = v
    if intervals and n_trials < 1:
        raise ValueError("interval(...) overrides need n_trials >= 1")
    rng = np.random.default_rng(seed)
    samples = [dict()]
    if intervals:
        cols = []
        for key, iv in intervals:
            if sampler == "grid":
                cols.append(np.linspace(iv.lo, iv.hi, n_trials))
            elif sampler == "random":
                cols.append(rng.uniform(iv.lo, iv.hi, n_trials))
            else:
                raise ValueError(f"unknown sampler '{sampler}' (grid or random)")
        samples = [{k: float(c[i]) for (k, _), c in zip(intervals, cols)} for i in range(n_trials)]
    trials = []
    for combo in itertools.product(*[vals for _, vals in axes]):
        for smp in samples:
            t = dict(fixed)
            t.update({k: v for (k, _), v in zip(axes, combo)})
            t.update(smp)
            trials.append(t)
    return trials


def solver_kwargs(trial: dict) -> dict:
    """The kwargs Hydra would hand to SGSolver / FSGSolver for this trial (the YAML interpolations resolved)."""
    kind = str(trial.get("solver", "spectral/fsg"))
    if kind not in SOLVER_DEFAULTS:
        raise ValueError(f"solver={kind}: this front-end drives spectral/sg and spectral/fsg")
    g = dict(GLOBAL_DEFAULTS)
    g.update({k: v for k, v in trial.items() if k in GLOBAL_DEFAULTS})
    kw = dict(SOLVER_DEFAULTS[kind])
    kw.update(Re=float(g["Re"]), lid_velocity=g["lid_velocity"], Lx=g["Lx"], Ly=g["Ly"], nx=int(g["N"]), ny=int(g["N"]),
              tolerance=float(g["tolerance"]), max_iterations=int(g["max_iterations"]))
    for k, v in trial.items():
        if k.startswith("solver."):
            kw[k[len("solver."):]] = v
    kw["_kind"] = kind
    return kw


# ------------------------------------------------------------------------------------------------------------------
# Objectives (main.py:142-225)
# ------------------------------------------------------------------------------------------------------------------
def fv_l2_objective(validation_errors: dict) -> float:
    """sqrt(u_L2_error^2 + v_L2_error^2) against the non-regularised FV solution; inf when it is missing (main.py:142-154)."""
    u = validation_errors.get("u_L2_error", float("inf"))
    v = validation_errors.get("v_L2_error", float("inf"))
    return math.sqrt(u**2 + v**2)


def botella_vortex_objective(metrics: Metrics, Re: int, root: str = ".") -> float:
    """RMS of |psi_min rel. error|, |dx|, |dy| of the primary vortex against data/validation/botella (main.py:157-203).
    As in the reference, only the columns psi_min, psi_min_x, psi_min_y count and falsy entries are skipped -- the
    Re=1000 table names its columns psi_primary / x_primary / y_primary, so that objective is inf there too."""
    path = os.path.join(root, "data/validation/botella", f"botella_Re{int(Re)}_vortex.csv")
    if not os.path.exists(path):
        return float("inf")
    with open(path) as f:
        rows = list(csv.DictReader(line for line in f if not line.lstrip().startswith("#")))
    if not rows:
        return float("inf")
    ref = {k: float(v) for k, v in rows[0].items() if v not in (None, "")}
    errs = []
    if ref.get("psi_min"):
        errs.append(abs(metrics.psi_min - ref["psi_min"]) / abs(ref["psi_min"]))
    if ref.get("psi_min_x"):
        errs.append(abs(metrics.psi_min_x - ref["psi_min_x"]))
    if ref.get("psi_min_y"):
        errs.append(abs(metrics.psi_min_y - ref["psi_min_y"]))
    return math.sqrt(sum(e * e for e in errs) / len(errs)) if errs else float("inf")


def objective_value(kind: str, validation_errors: dict, metrics: Metrics, Re, root: str = ".") -> float:
    if kind == "multi":
        raise ValueError("Multi-objective optimization is not supported by hydra-optuna-sweeper 1.x. "
                         "Use objective=fv_l2_error or objective=botella_vortex instead.")
    if kind == "botella_vortex":
        return botella_vortex_objective(metrics, int(Re), root)
    return fv_l2_objective(validation_errors)


# ------------------------------------------------------------------------------------------------------------------
# Batched execution
# ------------------------------------------------------------------------------------------------------------------
@dataclass
class TrialResult:
    index: int
    overrides: dict
    params: dict
    metrics: dict
    validation_errors: dict
    objective: float
    group: int = 0
    extras: dict = field(default_factory=dict)


_SHARED = ("_kind", "nx", "ny", "Lx", "Ly", "basis_type", "max_iterations", "n_levels", "coarse_tolerance_factor",
           "prolongation_method", "restriction_method", "multigrid", "name")


def group_trials(kwargs_list):
    """Indices of trials that can share one batched solve: everything but Re, corner smoothing / treatment, tolerance,
    beta^2, CFL and lid velocity (the per-cavity inputs of `ldc_set_cases`) must agree."""
    groups = {}
    for i, kw in enumerate(kwargs_list):
        if kw["nx"] != kw["ny"]:
            raise ValueError("the sweep front-end drives square grids (nx == ny == N), as conf/solver/spectral/sg.yaml does")
        key = tuple((k, kw.get(k)) for k in _SHARED)
        groups.setdefault(key, []).append(i)
    return list(groups.values())


def run_trials(trials, objective: str = "fv_l2_error", root: str = ".", device=None, vortex_metrics: bool = True,
               solve=None):
    """Solve every trial (a dict of overrides, see `expand_overrides`) in as few batched solves as possible.

    Returns one TrialResult per trial, in trial order, on every rank (the sharded solve all-gathers its results).
    `root` is the directory holding data/validation/...; `solve` overrides `sharded.solve_sweep` (tests)."""
    from .sharded import solve_sweep

    kws = [solver_kwargs(t) for t in trials]
    results = [None] * len(trials)
    solve = solve or solve_sweep
    for gi, idx in enumerate(group_trials(kws)):
        k0 = kws[idx[0]]
        kind = "fsg" if k0["_kind"] == "spectral/fsg" else "sg"
        N = k0["nx"]
        per = {name: np.array([kws[i][name] for i in idx], dtype=np.float64)
               for name in ("Re", "corner_smoothing", "tolerance", "beta_squared", "CFL", "lid_velocity")}
        treat = [kws[i]["corner_treatment"] for i in idx]
        shared = dict(beta_squared=per["beta_squared"], CFL=per["CFL"], lid_velocity=per["lid_velocity"],
                      corner_treatment=treat if len(set(treat)) > 1 else treat[0], Lx=k0["Lx"], Ly=k0["Ly"],
                      basis=str(k0["basis_type"]).lower())
        if kind == "fsg":
            shared.update(n_levels=k0["n_levels"], coarse_tolerance_factor=k0["coarse_tolerance_factor"],
                          prolongation=k0["prolongation_method"])
        t0 = time.time()
        r = solve(N, per["Re"], per["corner_smoothing"], per["tolerance"], int(k0["max_iterations"]), kind=kind,
                  fields="all", device=device, balance="interleave", **shared)   # sweeps arrive ordered by Re / smoothing
        wall = time.time() - t0
        # post-processing of this rank's shard only (validation-error kernel, streamfunction solves), then one
        # all-gather of the small per-trial records
        lo, hi = _my_block(r.bounds, len(idx))
        verr_loc = _group_validation_errors(N, k0, {k: v[lo:hi] for k, v in per.items()}, r.u[lo:hi], r.v[lo:hi], r.p[lo:hi],
                                            root, device) if hi > lo else []
        # streamfunction / vortex metrics of the whole block in one batched device call (`ldc_vortex_metrics`)
        vm_loc = [{} for _ in range(lo, hi)]
        if vortex_metrics and hi > lo:
            try:
                vm_all = _block_vortex_metrics(N, k0, r.u[lo:hi], r.v[lo:hi], device)
                vm_loc = [{} if r.diverged[j] else vm_all[j - lo] for j in range(lo, hi)]
            except Exception:   # base.py:156-160 swallows these too
                pass
        verrs, vms = _gather_records(verr_loc, vm_loc, r.bounds, len(idx))
        for j, i in enumerate(idx):
            vm = vms[j]
            m = Metrics(iterations=int(r.iterations[j]), converged=bool(r.converged[j]),
                        final_residual=float(r.rel_residual[j]), wall_time_seconds=wall / len(idx),
                        **{k: float(v) for k, v in vm.items() if k in Metrics.__dataclass_fields__})
            params = {k: v for k, v in kws[i].items() if not k.startswith("_")}
            # a diverged trial has no solution: its objective is +inf (an optimiser must not mistake the error of an
            # all-zero or non-finite field for a poor but valid one)
            obj = float("inf") if r.diverged[j] else objective_value(objective, verrs[j], m, params["Re"], root)
            results[i] = TrialResult(i, dict(trials[i]), params, m.to_mlflow(), verrs[j], obj, gi,
                                     dict(diverged=bool(r.diverged[j]), batch=len(idx), batch_wall_s=wall))
    return results


def _block_vortex_metrics(N, k0, u, v, device):
    """Vortex metrics of a block of cavities [b, n, n] on the device: one handle for the block, one call."""
    import torch

    from .batched import BatchedSpectralLDC
    from .postprocess import vortex_metrics_device

    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    with BatchedSpectralLDC(N, N, len(u), k0["Lx"], k0["Ly"], str(k0["basis_type"]).lower(), device) as s:
        s.set_cases(100.0)          # (the streamfunction does not depend on the physics; the ABI wants the cases set)
        s.set_state(np.nan_to_num(u), np.nan_to_num(v), None)
        return vortex_metrics_device(s)


def _my_block(bounds, n):
    """[lo, hi) of the cavities this rank post-processes: its shard of the solve, or everything without a process group."""
    from .sharded import _dist

    dist = _dist()
    if dist is None or dist.get_world_size() == 1 or len(bounds) != dist.get_world_size():
        return 0, n
    return bounds[dist.get_rank()]


def _gather_records(verr_loc, vm_loc, bounds, n):
    """Per-trial dicts of every rank's block, concatenated in trial order on every rank."""
    from .sharded import _dist

    dist = _dist()
    if dist is None or dist.get_world_size() == 1 or len(bounds) != dist.get_world_size():
        return verr_loc, vm_loc
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, (verr_loc, vm_loc))
    verrs = [d for part in parts for d in part[0]]
    vms = [d for part in parts for d in part[1]]
    assert len(verrs) == n and len(vms) == n
    return verrs, vms


def _group_validation_errors(N, k0, per, u, v, p, root, device):
    """Batched L2 errors of one block of cavities: one `ldc_validation_errors` call per distinct integer Re and stored
    solution."""
    from .batched import BatchedSpectralLDC
    from .validation import REF_DIRS, StoredSolution, batched_validation_errors

    out = [dict() for _ in range(per["Re"].size)]
    todo = []
    for Re in sorted(set(int(x) for x in per["Re"])):
        for base, suffix in REF_DIRS:
            path = os.path.join(root, base, f"Re{Re}", "solution.vts")
            if os.path.exists(path):
                todo.append((Re, suffix, path))
    if not todo:
        return out
    import torch

    dev = device if device is not None else (torch.cuda.current_device() if torch.cuda.is_available() else 0)
    with BatchedSpectralLDC(N, N, per["Re"].size, k0["Lx"], k0["Ly"], str(k0["basis_type"]).lower(), dev) as b:
        b.set_cases(per["Re"])
        b.set_state(np.ascontiguousarray(u), np.ascontiguousarray(v), np.ascontiguousarray(p))
        for Re, suffix, path in todo:
            e = batched_validation_errors(b, StoredSolution.from_vts(path))
            for j in np.nonzero(per["Re"].astype(int) == Re)[0]:
                out[j][f"u_L2_error{suffix}"] = float(e[j, 0])
                out[j][f"v_L2_error{suffix}"] = float(e[j, 1])
    return out


class BatchObjective:
    """Ask-and-tell bridge for an Optuna study: `values = BatchObjective(base)(list_of_param_dicts)` evaluates the
    whole batch of suggested parameters in one batched solve.

        obj = BatchObjective({"solver": "spectral/fsg", "N": 30, "Re": 1000}, objective="botella_vortex")
        trials = [study.ask() for _ in range(64)]
        values = obj([{"solver.corner_smoothing": t.suggest_float("solver.corner_smoothing", 0.02, 0.35)} for t in trials])
        for t, v in zip(trials, values): study.tell(t, v)
    """

    def __init__(self, base: dict, objective: str = "fv_l2_error", root: str = ".", device=None):
        self.base, self.objective, self.root, self.device = dict(base), objective, root, device
        self.history = []

    def __call__(self, suggestions):
        trials = [{**self.base, **s} for s in suggestions]
        res = run_trials(trials, self.objective, self.root, self.device)
        self.history.extend(res)
        return [r.objective for r in res]


def log_to_mlflow(results, experiment: str = "LDC-Sweep", parent_name: str = "batched-sweep"):
    """Fan the records out as one MLflow child run per trial under a parent run (what MLflowSweepCallback,
    src/utilities/mlflow/callback.py:190-318, arranges for a Hydra sweep).  Needs MLflow; raises ImportError otherwise."""
    import mlflow

    mlflow.set_experiment(experiment)
    with mlflow.start_run(run_name=parent_name) as parent:
        for r in results:
            with mlflow.start_run(run_name=f"trial-{r.index}", nested=True):
                mlflow.log_params(r.params)
                mlflow.log_metrics({k: v for k, v in r.metrics.items() if isinstance(v, (int, float))})
                if r.validation_errors:
                    mlflow.log_metrics(r.validation_errors)
                if math.isfinite(r.objective):
                    mlflow.log_metric("objective", r.objective)
        return parent.info.run_id


def main(argv=None):
    import argparse

    ap = argparse.ArgumentParser(description="Batched Hydra-style sweep of spectral LDC solves (one batched solve per group)")
    ap.add_argument("overrides", nargs="*", help="Hydra overrides: solver=spectral/fsg N=32 Re=100,400 solver.corner_smoothing=...")
    ap.add_argument("--n-trials", type=int, default=0, help="samples per interval(...) override")
    ap.add_argument("--sampler", default="grid", choices=("grid", "random"))
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--root", default=".", help="directory holding data/validation/")
    ap.add_argument("--out", default=None, help="write the per-trial records as JSON")
    ap.add_argument("--mlflow", action="store_true", help="also log one MLflow child run per trial")
    args = ap.parse_args(argv)
    import torch
    import torch.distributed as dist

    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if "RANK" in os.environ and not dist.is_initialized():
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
    rank = dist.get_rank() if dist.is_initialized() else 0
    t0 = time.time()
    if dist.is_initialized():
        dist.barrier()   # brings the communicator up (seconds for NCCL on 8 GPUs) outside the sweep's own clock
    comm_setup = time.time() - t0
    objective = "fv_l2_error"
    overrides = []
    for ov in args.overrides:
        if ov.startswith("optuna.objective="):
            objective = ov.split("=", 1)[1]
        else:
            overrides.append(ov)
    trials = expand_overrides(overrides, args.n_trials, args.sampler, args.seed)
    t0 = time.time()
    res = run_trials(trials, objective, args.root)
    wall = time.time() - t0
    if rank == 0:
        best = min(res, key=lambda r: r.objective) if res else None
        summary = dict(trials=len(res), groups=len({r.group for r in res}), wall_s=wall, comm_setup_s=comm_setup, objective=objective,
                       converged=sum(r.metrics["converged"] for r in res),
                       best=dict(index=best.index, overrides=best.overrides, objective=best.objective) if best else None)
        print(json.dumps(summary))
        if args.out:
            with open(args.out, "w") as f:
                json.dump(dict(summary=summary, trials=[asdict(r) for r in res]), f, indent=1, default=float)
        if args.mlflow:
            log_to_mlflow(res)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return res


if __name__ == "__main__":
    main()
