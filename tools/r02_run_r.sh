#!/bin/bash
# quad loss head: GPU suite, default bench, head ncu
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r_pytest_all.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r_pytest_all.txt
timeout 600 python bench.py --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/r_bench.json"))
print("default: ms_per_step %.2f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
timeout 300 python bench.py --num-envs 512 --steps 10 --warmup 5 --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/r_bench_512.json 2> gpurun_out/r_bench_512.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/r_bench_512.json"))
print("512 envs: ms_per_step %.2f e2e %.2f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
