#!/bin/bash
# ncu --set full of the two persistent streamed kernels (one capture each; plain run first)
mkdir -p gpurun_out
timeout 300 python scripts/ncu_target.py 128 8 20 stream_tma > gpurun_out/plain_target_stream.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stream_tma -s 1 -c 1 -f -o gpurun_out/prof_stream_tma \
    python scripts/ncu_target.py 128 8 20 stream_tma > gpurun_out/ncu_target_stream.log 2>&1
timeout 300 python scripts/ncu_target.py 128 64 10 stream_tma > gpurun_out/plain_target_wide.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stream_tma_wide -s 1 -c 1 -f -o gpurun_out/prof_stream_tma_wide \
    python scripts/ncu_target.py 128 64 10 stream_tma > gpurun_out/ncu_target_wide.log 2>&1
ls -la gpurun_out/*.ncu-rep
