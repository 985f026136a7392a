#!/bin/bash
# 2-GPU pass: device selection, real NCCL parity, DP under graphs, weak + strong bench line
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/e_smi.txt
( time timeout 900 python -m pytest tests/test_gpu_graphs.py -m gpu -q ) > gpurun_out/e_pytest_graphs.txt 2>&1; echo "graphs tests rc=$?"; tail -4 gpurun_out/e_pytest_graphs.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_graph_check.py > gpurun_out/e_dp_graph_check.txt 2>&1; echo "dp_graph_check rc=$?"; grep dp_graph_check gpurun_out/e_dp_graph_check.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/e_bench_2gpu.json 2> gpurun_out/e_bench_2gpu.err; echo "bench 2gpu rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/e_bench_2gpu.json"))
    print("2gpu weak: ms %.1f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), "strong:", {k: d["strong"][k] for k in ("ms_per_step", "value", "num_envs_per_gpu")}, d["strong"].get("e2e"))
except Exception as e:
    print("2gpu bench parse failed", e)
PY
