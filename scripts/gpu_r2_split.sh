#!/bin/bash
# split K loop of k_stream_tma (4 x 4 tiles): parity, then timings with and without it
mkdir -p gpurun_out
LDC_STREAM_CW=1 timeout -s KILL 400 python scripts/r2_stream_check.py parity > gpurun_out/split_parity.log 2>&1; echo "parity rc=$?"
tail -9 gpurun_out/split_parity.log
for sp in 0 1; do
  for nb in "128 8 200" "128 1 200" "128 16 200" "100 8 200"; do
    LDC_STREAM_CW=1 LDC_STREAM_SPLIT=$sp timeout -s KILL 200 python scripts/probe_one.py $nb stream_tma 2>&1 | tail -1 | sed "s/^/split=$sp /"
  done
done
