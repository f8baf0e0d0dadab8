#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): bash scripts/gpu_r2_san.sh racecheck|memcheck
tool=$1
mkdir -p gpurun_out
timeout -s KILL 300 python scripts/sanitize_target.py > gpurun_out/san_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/san_plain.log; exit 1; }
tail -3 gpurun_out/san_plain.log
timeout -s KILL 1500 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_target.py > gpurun_out/san_$tool.log 2>&1
echo "rc=$?" >> gpurun_out/san_$tool.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Hazard|rc=|sanitize target OK|closed|not available" gpurun_out/san_$tool.log | head -40
