#!/bin/bash
# GPU visit: parity tests, small-grid probe, bench, launch list of the bench command and one full ncu capture of the
# stepping kernel at the bench launch shape (256 cavities x 200 pseudo-steps) for roofline.traffic.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python scripts/gpu_probe.py small > gpurun_out/probe_small.log 2>&1; cp gpurun_out/probe.json gpurun_out/probe_small.json
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-time-to-tol > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-time-to-tol > gpurun_out/ncu_bench.log 2>&1
timeout 300 python scripts/ncu_target.py 64 256 200 > gpurun_out/plain_target.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_resident -s 1 -c 1 -f -o gpurun_out/prof_eo_bench \
    python scripts/ncu_target.py 64 256 200 > gpurun_out/ncu_target.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -22 gpurun_out/probe_small.log; cat gpurun_out/bench.json | cut -c1-600
