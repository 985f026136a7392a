#!/bin/bash
# usage: tools/r02_gpu_pass.sh <prefix> [tests] [launches]  -- default bench (no CPU baseline) into gpurun_out/<prefix>_bench.json;
# "tests": the GPU suite first; "launches": an ncu launch list of one plain-launch step afterwards
P=$1; shift
mkdir -p gpurun_out
if [[ " $* " == *" tests "* ]]; then
  ( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/${P}_pytest_all.txt 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/${P}_pytest_all.txt | tail -2
fi
timeout 600 python bench.py --no-cpu-baseline --no-configs --gae-sweep "" > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/${P}_bench.json"))
print("default: ms_per_step %.2f value %.4g e2e %.4g" % (d["ms_per_step"], d["value"], d["e2e"]["value"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d["clocks"])
PY
if [[ " $* " == *" launches "* ]]; then
  ARGS='--steps 1 --warmup 1 --prof-steps 1 --no-cpu-baseline --no-e2e --no-configs --no-graphs --gae-sweep ""'
  eval python bench.py $ARGS > gpurun_out/${P}_plain.json 2> gpurun_out/${P}_plain.err && eval ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${P}_launches.csv python bench.py $ARGS > gpurun_out/${P}_ncu.log 2>&1; echo "ncu rc=$?"
fi
