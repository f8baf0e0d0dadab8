#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python scripts/r2_stream_check.py parity > gpurun_out/d_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/d_parity.log
tail -12 gpurun_out/d_parity.log
for cfg in "128 8" "128 64"; do
  set -- $cfg
  timeout -s KILL 200 python scripts/r2_s2_debug.py $1 $2 60 gpurun_out/s2_dbg_N$1_B$2.txt 2>&1 | tee gpurun_out/d_dbg_N$1_B$2.log
done
timeout -s KILL 300 python scripts/probe_one.py 128 8 200 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 128 64 100 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 96 64 100 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 64 8 500 stream_tma
timeout -s KILL 300 python scripts/probe_one.py 64 37 300 stream_tma
timeout -s KILL 900 python -m pytest tests/test_gpu_round2.py -q 2>&1 | tail -15
