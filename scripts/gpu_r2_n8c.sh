#!/bin/bash
# 8 GPUs: bench (with sweep_to_tol), BASELINE config 5 and config 4 through the sweep front-end
N=${1:-8}
bash scripts/gpu_r2_bench2.sh $N
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    scripts/config5_demo.py --batch $((8 * N)) --levels 2 --max-iter 500000 --out gpurun_out/config5_n$N.json 2>&1 | tail -1 | cut -c1-420
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 \
    scripts/sweep_cli_demo.py --N 64 --grid 32 2>&1 | tail -3 | cut -c1-500
cp gpurun_out/sweep_cli_demo.json gpurun_out/sweep_cli_config4_n$N.json 2>/dev/null
