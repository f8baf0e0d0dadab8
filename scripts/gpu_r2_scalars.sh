#!/bin/bash
# per-step scalars spread over lanes (step_scalars): timings of the resident kernels, then the GPU suite
mkdir -p gpurun_out
for nb in "64 256 200 resident_eo" "64 148 200 resident_eo" "30 1 4000 auto" "32 128 2000 auto" "15 1 4000 auto" "48 256 300 auto" "64 256 100 resident"; do
  timeout -s KILL 200 python scripts/probe_one.py $nb 2>&1 | tail -1 | cut -c1-170
done
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/sc_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/sc_pytest_gpu.log
