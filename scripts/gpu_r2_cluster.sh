#!/bin/bash
# k_cluster_eo: parity, timings, clock samples (each under its own timeout)
mkdir -p gpurun_out
timeout 400 python scripts/r2_cluster_check.py parity > gpurun_out/cluster_parity.log 2>&1; echo "parity rc=$?"
tail -4 gpurun_out/cluster_parity.log
timeout 300 python scripts/r2_cluster_check.py time > gpurun_out/cluster_time.log 2>&1; echo "time rc=$?"
grep "^3 " gpurun_out/cluster_time.log | cut -c1-150
timeout 200 python scripts/r2_ce_debug.py 64 8 200 gpurun_out/ce_clocks_N64_B8.txt 2>&1 | tail -11
