#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/ncu_target.py 64 256 100 > gpurun_out/plain_target_eo.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_resident_eo -s 1 -c 1 -f -o gpurun_out/prof_eo_r2 \
    python scripts/ncu_target.py 64 256 100 > gpurun_out/ncu_target_eo.log 2>&1
tail -3 gpurun_out/ncu_target_eo.log; ls -la gpurun_out/prof_eo_r2.ncu-rep
