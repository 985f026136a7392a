#!/bin/bash
# 8-GPU pass: the driver's launch line for N = 8 (weak line + strong sub-record)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 5 > gpurun_out/k_bench_8gpu.json 2> gpurun_out/k_bench_8gpu.err; echo "bench 8gpu rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/k_bench_8gpu.json"))
    print("8gpu weak: ms %.1f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), "strong:", {k: d["strong"][k] for k in ("ms_per_step", "value", "num_envs_per_gpu")}, d["strong"].get("e2e"), d["clocks"])
except Exception as e:
    print("8gpu bench parse failed", e)
PY
tail -c 600 gpurun_out/k_bench_8gpu.err
