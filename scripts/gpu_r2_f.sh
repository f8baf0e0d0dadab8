#!/bin/bash
mkdir -p gpurun_out
LDC_STREAM_CW=2 timeout -s KILL 600 python scripts/r2_stream_check.py parity > gpurun_out/f_parity_cw2.log 2>&1; echo "parity cw2 rc=$?" >> gpurun_out/f_parity_cw2.log
tail -12 gpurun_out/f_parity_cw2.log
for cw in 1 2; do
  LDC_STREAM_CW=$cw timeout -s KILL 300 python scripts/probe_one.py 128 64 100 stream_tma
  LDC_STREAM_CW=$cw timeout -s KILL 300 python scripts/probe_one.py 128 18 100 stream_tma
  LDC_STREAM_CW=$cw timeout -s KILL 300 python scripts/probe_one.py 96 64 100 stream_tma
  LDC_STREAM_CW=$cw timeout -s KILL 300 python scripts/probe_one.py 64 64 300 stream_tma
done
timeout -s KILL 300 python scripts/probe_one.py 128 64 100 stream_tma
