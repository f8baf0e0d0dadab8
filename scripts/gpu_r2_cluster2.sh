#!/bin/bash
# whole GPU suite with the cluster kernel in AUTO, then one rank's share of the config-4 sweep (tail hand-over)
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/cl_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/cl_pytest_gpu.log
timeout -s KILL 300 python scripts/r2_tail_switch.py 128 2>&1 | tail -6
