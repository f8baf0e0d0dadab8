#!/bin/bash
# The allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (which collects it last).
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/noalloc_alone.log
tail -5 gpurun_out/noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_suite.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/noalloc_suite.log
tail -8 gpurun_out/noalloc_suite.log
