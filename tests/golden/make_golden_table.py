"""Mint the validation-table fixture from the UNMODIFIED reference (base.py:890-964).

    python tests/golden/make_golden_table.py     # authoring container only (needs /root/reference)

The stub MLflow module gets an active run and a `log_table` that keeps the DataFrame; the metrics are set by hand (the
method only formats them), the CSV files are the reference's own (copied into the fixture as text so the test can rebuild
data/validation/botella in a scratch directory).  `table.json`: per Re, the CSV text, the metrics and the logged rows.
"""
from __future__ import annotations

import json
import os

from make_golden import HERE, REF, sg_kwargs, import_reference

METRICS = dict(psi_min=-0.1187654, omega_center=-2.0654321, psi_min_x=0.4688, psi_min_y=0.5661, psi_BL=2.29e-4, omega_BL=1.0987,
               psi_BL_x=0.1357, psi_BL_y=0.1121, psi_BR=1.7254e-3, omega_BR=1.1102, psi_BR_x=0.8645, psi_BR_y=0.1115)


def main():
    SGSolver, *_ = import_reference()
    from solvers import base

    kept = {}
    base.mlflow.active_run = lambda: True
    base.mlflow.log_table = lambda df, artifact_file: kept.update(rows=df.to_dict(orient="records"), file=artifact_file)
    os.chdir(REF)
    out = {}
    for Re in (100, 1000, 250):
        s = SGSolver(**sg_kwargs(12, float(Re)))
        for k, v in METRICS.items():
            setattr(s.metrics, k, v)
        kept.clear()
        s.mlflow_log_validation_table()
        path = f"data/validation/botella/botella_Re{Re}_vortex.csv"
        out[str(Re)] = dict(csv=open(path).read() if os.path.exists(path) else None, rows=kept.get("rows"), file=kept.get("file"))
        print(Re, kept.get("rows", None) and kept["rows"][:2])
    with open(os.path.join(HERE, "table.json"), "w") as f:
        json.dump(dict(metrics=METRICS, cases=out), f, indent=1, ensure_ascii=False)


if __name__ == "__main__":
    main()
