#!/bin/bash
# bench under torchrun on N GPUs (usage: bash scripts/gpu_r2_bench2.sh N), NCCL_DEBUG=INFO as the driver runs it
N=$1
mkdir -p gpurun_out
NCCL_DEBUG=INFO timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
grep -c "NCCL INFO" gpurun_out/bench_n$N.err; grep -m2 "nranks\|NVLS" gpurun_out/bench_n$N.err | cut -c1-200
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_n$N.json").read().strip().splitlines() if l.startswith("{")][-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "n_gpus", d["n_gpus"]); print("sweep", d.get("sweep_to_tol"))
PY
