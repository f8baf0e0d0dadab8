#!/bin/bash
# One GPU-box visit: parity tests, FP64 probe, bench, ncu launch list + full capture of the resident kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python scripts/gpu_probe.py > gpurun_out/probe.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
if [ "$1" != "nocu" ]; then
timeout 300 python bench.py --steps 2 --warmup 3 --pseudo 20 --no-cpu-baseline --no-time-to-tol > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --pseudo 20 --no-cpu-baseline --no-time-to-tol > gpurun_out/ncu_bench.log 2>&1
timeout 300 python scripts/ncu_target.py 64 256 8 > gpurun_out/plain_target.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_resident -s 1 -c 1 -f -o gpurun_out/prof_resident \
    python scripts/ncu_target.py 64 256 8 > gpurun_out/ncu_target.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/probe.log | tail -12; cat gpurun_out/bench.json; tail -2 gpurun_out/bench.err; cat gpurun_out/bench_ref.json
