#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest_gpu.log
tail -15 gpurun_out/e_pytest_gpu.log
