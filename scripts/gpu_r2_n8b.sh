#!/bin/bash
N=${1:-8}
bash scripts/gpu_r2_bench2.sh $N
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 \
    scripts/sweep_cli_demo.py --N 64 --grid 32 2>&1 | tail -4 | cut -c1-700
cp gpurun_out/sweep_cli_demo.json gpurun_out/sweep_cli_config4_n$N.json 2>/dev/null
