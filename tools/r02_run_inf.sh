#!/bin/bash
# rollout inference after the small-launch tile rule: GPU suite, then the bench's inference records (4096 / 16384 / 65536 envs) and config 0
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/inf_pytest_all.txt 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/inf_pytest_all.txt | tail -2
timeout 600 python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/inf_bench.json 2> gpurun_out/inf_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/inf_bench.json"))
print("ms_per_step %.2f" % d["ms_per_step"], d["clocks"]["sm_mhz"])
for r in d["inference"]: print(r["num_envs"], "ms_per_call %.4f api %.4f tensor_pipe_frac %.3f" % (r["ms_per_call"], r["ms_per_call_api"], r["tensor_pipe_frac"]))
for c in d["configs"]: print(c["name"], "%.3f ms" % c["ms_per_step"], c["policy_step_ms"])
PY
