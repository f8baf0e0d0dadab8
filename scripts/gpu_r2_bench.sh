#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
for k in ("value","ms_per_step","gpu_launches"): print(k, d[k])
print("e2e", d["e2e"]); print("roofline", {k:d["roofline"][k] for k in ("achieved","peak","frac","frac_algorithmic","dgemm_tflops","dfma_peak")})
print("cpu", d.get("cpu_baseline")); print("tt", d.get("time_to_tol")); print("sweep", d.get("sweep_to_tol")); print("clocks", d["clocks"])
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-400 gpurun_out/bench_ref.json
