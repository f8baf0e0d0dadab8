#!/bin/bash
# whole GPU suite, then ncu --set full of one k_cluster_eo launch (plain run of the same command first)
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/cl_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/cl_pytest_gpu.log
timeout 300 python scripts/ncu_target.py 64 8 50 stream_tma > gpurun_out/plain_target_cluster.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_cluster_eo -s 1 -c 1 -f -o gpurun_out/prof_cluster_eo \
    python scripts/ncu_target.py 64 8 50 stream_tma > gpurun_out/ncu_target_cluster.log 2>&1
tail -3 gpurun_out/plain_target_cluster.log gpurun_out/ncu_target_cluster.log
ls -la gpurun_out/*.ncu-rep
