#!/bin/bash
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_bw scripts/tma_bw.cu && timeout -s KILL 120 /tmp/tma_bw > gpurun_out/tma_bw.txt 2>&1
cat gpurun_out/tma_bw.txt
for cfg in "128 8" "128 64"; do
  set -- $cfg
  timeout -s KILL 200 python scripts/r2_s2_debug.py $1 $2 60 gpurun_out/s2_dbg_N$1_B$2.txt 2>&1 | tee gpurun_out/c_dbg_N$1_B$2.log
done
timeout -s KILL 300 python scripts/probe_one.py 128 8 200 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 128 64 100 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 128 8 200 stream_eo
timeout -s KILL 900 python -m pytest tests/test_gpu_round2.py -x -q 2>&1 | tail -15
