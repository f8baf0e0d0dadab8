#!/bin/bash
N=${1:-8}
bash scripts/gpu_r2_bench2.sh $N
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
    scripts/config5_demo.py --batch $((8 * N)) --levels 2 --max-iter 500000 --out gpurun_out/config5_n$N.json 2>&1 | tail -3 | cut -c1-600
