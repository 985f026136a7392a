#!/bin/bash
# after enlarging the weight-preparation grids: GPU suite, 512-env (strong-scaling shard) bench, default bench
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/q_pytest_all.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/q_pytest_all.txt
timeout 300 python bench.py --num-envs 512 --steps 10 --warmup 5 --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/q_bench_512.json 2> gpurun_out/q_bench_512.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/q_bench_512.json"))
print("512 envs: ms_per_step %.2f e2e %.2f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), "profile pass %.2f sum kernels %.2f" % (d["profile_pass"]["ms_per_step"], d["profile_pass"]["sum_kernel_ms_per_step"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/q_bench.json"))
print("default: ms_per_step %.2f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["roofline"], d["clocks"])
PY
