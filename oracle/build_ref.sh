#!/bin/bash
# oracle/_ref: the UNMODIFIED reference sources of the spectral path, staged beside the oracle so that they travel to the
# GPU box (oracle/_ref is git-ignored, not gpurun-ignored).  Test / measurement infrastructure only: tests/, bench.py's
# CPU legs and smoke() may import it (through oracle/ref_loader.py); the product package never does.
#
#   bash oracle/build_ref.sh [/root/reference]
#
# Copies <reference>/src/solvers (pure Python: base.py, datastructures.py, spectral/**, fv/**) byte for byte.  The three
# third-party modules the reference imports at module top and that are not installed here (pyvista, mlflow, pyamg:
# base.py:10-11, fv/linear_solvers/scipy_solver.py:4) are stubbed at import time by oracle/ref_loader.py, not written
# into the copy.  No file of the reference enters the repository's history.
set -euo pipefail
REF="${1:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
if [ ! -d "$REF/src/solvers" ]; then
    echo "build_ref: $REF/src/solvers not found (nothing staged)" >&2
    exit 0
fi
rm -rf "$HERE/_ref"
mkdir -p "$HERE/_ref/src"
cp -r "$REF/src/solvers" "$HERE/_ref/src/solvers"
find "$HERE/_ref" -name "__pycache__" -type d -prune -exec rm -rf {} +
( cd "$REF/src" && find solvers -name "*.py" -print0 | sort -z | xargs -0 sha256sum ) > "$HERE/_ref/SHA256SUMS"
echo "build_ref: staged $(find "$HERE/_ref/src" -name '*.py' | wc -l) files from $REF/src/solvers"
