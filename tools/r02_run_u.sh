#!/bin/bash
# four-column skinny wgrad + lean thin forward: GPU suite, default bench, launch list
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/u_pytest_all.txt 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/u_pytest_all.txt | tail -2
timeout 600 python bench.py --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/u_bench.json 2> gpurun_out/u_bench.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/u_bench.json"))
print("default: mu_per_step %.2f value %.4g e2e %.4g" % (d["mu_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 3) for k, v in d["kernel_mu_per_step"].items()}, d["clocks"])
PY
python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/u_plain.json 2> gpurun_out/u_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/u_launches.csv python bench.py --steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep "" > gpurun_out/u_ncu.log 2>&1; echo "ncu rc=$?"
