#!/bin/bash
# usage: tools/r02_run_n_gpus.sh N  -- the driver's launch line for N GPUs (weak line + strong sub-record) into gpurun_out/z_bench_Ngpu.json
N=$1
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/z_bench_${N}gpu.json 2> gpurun_out/z_bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/z_bench_${N}gpu.json"))
print("${N}gpu weak: ms %.1f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), "strong:", {k: d["strong"][k] for k in ("ms_per_step", "value", "num_envs_per_gpu")}, d["strong"].get("e2e"), d["clocks"])
PY
