#!/bin/bash
# Round 2, GPU visit A: the persistent TMA-fed streamed kernel -- parity first, then timings, then the whole GPU suite.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/a_smi.txt
timeout -s KILL 600 python scripts/r2_stream_check.py parity > gpurun_out/a_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/a_parity.log
tail -25 gpurun_out/a_parity.log
timeout -s KILL 600 python scripts/r2_stream_check.py time > gpurun_out/a_time.log 2>&1; echo "time rc=$?" >> gpurun_out/a_time.log
tail -20 gpurun_out/a_time.log
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest_gpu.log
tail -15 gpurun_out/a_pytest_gpu.log
