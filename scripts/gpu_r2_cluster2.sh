#!/bin/bash
# whole GPU suite with the cluster kernel in AUTO, then one rank's share of the config-4 sweep (tail hand-over) and config 5
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -m gpu -x -q > gpurun_out/cl_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/cl_pytest_gpu.log
timeout -s KILL 300 python scripts/r2_tail_switch.py 128 2>&1 | tail -4
timeout 600 python scripts/config5_demo.py --batch 8 --levels 2 --max-iter 500000 --out gpurun_out/config5_n1_b8_cluster.json > /dev/null 2>&1
python -c "
import json; d=json.load(open('gpurun_out/config5_n1_b8_cluster.json')); print({k: d[k] for k in ('wall_s','kernel_ms_max_over_ranks','iterations','converged')})"
